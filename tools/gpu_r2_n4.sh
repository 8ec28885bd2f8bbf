# weak-scaling bench on one 4-GPU box: N = 4, 2 (run with gpurun --gpus 4)
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n4.txt 2>&1; rm -f gpurun_out/scale_rc.log
for N in 4 2; do
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "bench_n$N rc=$?" >> gpurun_out/scale_rc.log
  tail -1 gpurun_out/bench_n$N.log | cut -c1-200
done
cat gpurun_out/scale_rc.log
