python -m pytest tests/test_warp_gpu.py -x -q > gpurun_out/pytest_warp.log 2>&1; echo "pytest_warp rc=$?" > gpurun_out/rc.log
python bench.py --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
