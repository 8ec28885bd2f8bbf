# round-2 evidence: all GPU tests, smoke, kernel bench, full bench + reference arm, then (only after those exited 0) the ncu
# launch list of `bench.py --steps 2 --warmup 1` and one `--set full` capture of every kernel of one step
bash tools/gpu_r2_all.sh
if grep -q "rc=[1-9]" gpurun_out/rc.log; then echo "a stage failed: no ncu"; exit 1; fi
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extras > gpurun_out/ncu_launch.log 2>&1
echo "launches rc=$?" >> gpurun_out/rc.log
timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"warp_|ray_sample|voxel_fuse|dense_|voxel_count|smooth_|to_depth|wcat" -s 44 -c 22 -o gpurun_out/prof_step python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/ncu_step.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc.log; tail -3 gpurun_out/ncu_step.log
cat gpurun_out/rc.log
