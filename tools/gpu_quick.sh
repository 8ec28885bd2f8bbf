# quick GPU check: fusion tests + bench without the CPU arms
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" > gpurun_out/rc.log
python bench.py --no-cpu-baseline > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
cat gpurun_out/rc.log; tail -15 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/bench.log
