mkdir -p gpurun_out
timeout 300 python tools/kernel_bench.py --only ray_fwd,ray_bwd --iters 10 > gpurun_out/ray_bench.log 2>&1; tail -2 gpurun_out/ray_bench.log
timeout 900 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q 2>&1 | tail -2
