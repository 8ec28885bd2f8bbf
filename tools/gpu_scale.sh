# weak-scaling bench on one 8-GPU box: N = 8, 4, 2 (N = 1 is tools/gpu_final.sh); run with gpurun --gpus 8
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/scale_gpus.log 2>&1; nvidia-smi topo -m >> gpurun_out/scale_gpus.log 2>&1; nproc >> gpurun_out/scale_gpus.log; rm -f gpurun_out/scale_rc.log
for N in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "bench_n$N rc=$?" >> gpurun_out/scale_rc.log
  tail -1 gpurun_out/bench_n$N.log | cut -c1-220
done
cat gpurun_out/scale_rc.log
