mkdir -p gpurun_out
python tools/kernel_bench.py --only fuse_fwd,fuse_bwd --iters 2 --warmup 1 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:voxel_fuse_ -s 3 -c 2 -o gpurun_out/prof_fuse python tools/kernel_bench.py --only fuse_fwd,fuse_bwd --iters 2 --warmup 1 > gpurun_out/ncu_fuse.log 2>&1
echo rc=$?; cat gpurun_out/plain.log; tail -5 gpurun_out/ncu_fuse.log
