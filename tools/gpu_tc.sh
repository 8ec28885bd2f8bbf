mkdir -p gpurun_out
timeout 120 python tools/kernel_bench.py --only fuse_fwd,fuse_bwd --iters 5 > gpurun_out/tc_bench.log 2>&1; echo "tc_bench rc=$?"; tail -3 gpurun_out/tc_bench.log
timeout 600 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q > gpurun_out/pytest_fusion.log 2>&1; echo "pytest_fusion rc=$?"; tail -12 gpurun_out/pytest_fusion.log
bash tools/gpu_warp_quick.sh
