mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_warp_gpu.py -m gpu -x -q -k "batch8" > gpurun_out/pytest_warp8.log 2>&1; echo "pytest_warp8 rc=$?" > gpurun_out/rc.log
tail -5 gpurun_out/pytest_warp8.log
timeout 900 python -m pytest tests/test_train_step_gpu.py -m gpu -x -q > gpurun_out/pytest_train.log 2>&1; echo "pytest_train rc=$?" >> gpurun_out/rc.log
tail -30 gpurun_out/pytest_train.log
timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph > gpurun_out/plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"warp_|ray_sample|voxel_fuse|dense_|voxel_count|smooth_|to_depth" -s 38 -c 19 -o gpurun_out/prof_step python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph > gpurun_out/ncu_step.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc.log; tail -3 gpurun_out/ncu_step.log
cat gpurun_out/rc.log
