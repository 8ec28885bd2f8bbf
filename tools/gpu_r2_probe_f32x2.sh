# packed-fp32 issue-rate probe + one ncu --set full capture (with source) of the loss and statistics kernels
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/f32x2_probe tools/f32x2_probe.cu > gpurun_out/f32x2_probe.log 2>&1
timeout 120 /tmp/f32x2_probe >> gpurun_out/f32x2_probe.log 2>&1; echo "probe rc=$?" > gpurun_out/rc.log
cat gpurun_out/f32x2_probe.log
timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"warp_loss_bwd_kernel|warp_stats_kernel" -c 2 -o gpurun_out/prof_loss python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/ncu_loss.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc.log; tail -3 gpurun_out/ncu_loss.log
cat gpurun_out/rc.log
