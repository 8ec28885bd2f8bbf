# round 2: first GPU contact of the project-then-gather fused path (every step under its own timeout: a hung kernel
# must not take the box with it)
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
timeout 300 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q -k "fused_mlp_matches or degenerate" > gpurun_out/pytest_fuse_first.log 2>&1; echo "fuse_first rc=$?" > gpurun_out/rc.log
tail -30 gpurun_out/pytest_fuse_first.log
if grep -q "fuse_first rc=0" gpurun_out/rc.log; then
  timeout 1200 python -m pytest tests/test_fusion_gpu.py -m gpu -q > gpurun_out/pytest_fusion.log 2>&1; echo "pytest_fusion rc=$?" >> gpurun_out/rc.log
  tail -30 gpurun_out/pytest_fusion.log
  timeout 300 python tools/kernel_bench.py --only classify,fuse_fwd,fuse_bwd,ray_fwd,ray_bwd,pose_fwd,pose_bwd > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?" >> gpurun_out/rc.log
  cat gpurun_out/kernel_bench.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.log
  tail -8 gpurun_out/smoke.log
  timeout 300 python tools/kernel_bench.py --only fuse_fwd,fuse_bwd --iters 2 --warmup 1 > gpurun_out/plain_fuse.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_fuse.csv python tools/kernel_bench.py --only fuse_fwd,fuse_bwd --iters 2 --warmup 1 > gpurun_out/ncu_fuse.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.log
fi
cat gpurun_out/rc.log
