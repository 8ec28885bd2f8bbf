mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/umma_mn_probe tools/umma_mn_probe.cu > gpurun_out/mn_probe.log 2>&1
timeout 60 /tmp/umma_mn_probe >> gpurun_out/mn_probe.log 2>&1; echo "mn_probe rc=$?" > gpurun_out/rc.log
cat gpurun_out/mn_probe.log
timeout 600 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q -k "fused_mlp or baseline_size or batch8" > gpurun_out/pytest_fuse.log 2>&1; echo "pytest_fuse rc=$?" >> gpurun_out/rc.log
tail -5 gpurun_out/pytest_fuse.log
timeout 300 python tools/kernel_bench.py --only fuse_fwd,fuse_bwd > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?" >> gpurun_out/rc.log
cat gpurun_out/kernel_bench.log
timeout 300 python tools/kernel_bench.py --only fuse_fwd,fuse_bwd --iters 2 --warmup 1 > gpurun_out/plain_fuse.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,lts__t_bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"voxel_fuse|dense_|wcat" -c 40 --csv --log-file gpurun_out/launches_fuse.csv python tools/kernel_bench.py --only fuse_fwd,fuse_bwd --iters 2 --warmup 1 > gpurun_out/ncu_fuse.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.log
cat gpurun_out/rc.log
