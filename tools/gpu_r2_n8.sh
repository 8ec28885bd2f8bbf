# weak-scaling bench line at N GPUs of one box (run with gpurun --gpus N): N=${1:-8}
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "bench_n$N rc=$?"
tail -1 gpurun_out/bench_n$N.log | cut -c1-300
