mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_warp_gpu.py -m gpu -x -q -k "smooth or dropin or to_depth or golden" > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?" > gpurun_out/rc.log
tail -3 gpurun_out/pytest_quick.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-e2e > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step']); print({k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"warp_loss_bwd_kernel|warp_stats_kernel|ray_sample_bwd|to_depth" -c 10 -o gpurun_out/prof_loss python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/ncu_loss.log 2>&1
echo "ncu rc=$?" >> gpurun_out/rc.log; tail -3 gpurun_out/ncu_loss.log
cat gpurun_out/rc.log
