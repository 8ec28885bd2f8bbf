# quick GPU check: GPU tests (fusion first) + bench without the CPU arm
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q > gpurun_out/pytest_fusion.log 2>&1; echo "pytest_fusion rc=$?" > gpurun_out/rc.log
timeout 600 python -m pytest tests/test_warp_gpu.py -m gpu -x -q > gpurun_out/pytest_warp.log 2>&1; echo "pytest_warp rc=$?" >> gpurun_out/rc.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
cat gpurun_out/rc.log; tail -25 gpurun_out/pytest_fusion.log; tail -c 600 gpurun_out/bench.log
timeout 300 python tools/kernel_bench.py --only fuse_fwd,fuse_bwd,compact_fwd,compact_bwd > gpurun_out/kernel_bench.log 2>&1; cat gpurun_out/kernel_bench.log
