mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:warp_loss_ -s 6 -c 2 -o gpurun_out/prof_warp python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/ncu_warp.log 2>&1
echo rc=$?; tail -3 gpurun_out/ncu_warp.log
