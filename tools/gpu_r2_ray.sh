mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q > gpurun_out/pytest_fusion.log 2>&1; echo "pytest_fusion rc=$?" > gpurun_out/rc.log
tail -4 gpurun_out/pytest_fusion.log
timeout 300 python tools/kernel_bench.py --only ray_fwd,ray_bwd,fuse_fwd,fuse_bwd > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?" >> gpurun_out/rc.log
cat gpurun_out/kernel_bench.log; cat gpurun_out/rc.log
