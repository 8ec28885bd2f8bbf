# full GPU round: tests, smoke, bench (both arms), kernel bench, launch list, one ncu --set full over one step's kernels
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?" > gpurun_out/rc.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.log
python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "bench_ref rc=$?" >> gpurun_out/rc.log
python tools/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?" >> gpurun_out/rc.log
python tools/kernel_bench_aux.py > gpurun_out/kernel_bench_aux.log 2>&1; echo "kbench_aux rc=$?" >> gpurun_out/rc.log
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo "ncu_launch rc=$?" >> gpurun_out/rc.log
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:voxel_|ray_sample|warp_|smooth_|to_depth' -s 57 -c 19 -o gpurun_out/prof_step python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1; echo "ncu_full rc=$?" >> gpurun_out/rc.log
cat gpurun_out/rc.log; tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/kernel_bench.log
