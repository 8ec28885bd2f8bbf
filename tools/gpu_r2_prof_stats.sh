mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"warp_stats_kernel" -c 2 -o gpurun_out/prof_stats python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/ncu_stats.log 2>&1
echo "ncu rc=$?" > gpurun_out/rc.log; tail -3 gpurun_out/ncu_stats.log
cat gpurun_out/rc.log
