mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_warp_gpu.py -m gpu -x -q -k "golden" > gpurun_out/pytest_canary.log 2>&1; rc=$?; echo "canary rc=$rc" > gpurun_out/rc.log
tail -3 gpurun_out/pytest_canary.log
if [ $rc -ne 0 ]; then cat gpurun_out/rc.log; exit 1; fi
timeout 600 python -m pytest tests/test_warp_gpu.py tests/test_projection_gpu.py -m gpu -q > gpurun_out/pytest_warp.log 2>&1; echo "pytest_warp rc=$?" >> gpurun_out/rc.log
tail -5 gpurun_out/pytest_warp.log
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['ms_per_step'])); print({k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
tail -3 gpurun_out/bench_quick.log | cut -c1-300
cat gpurun_out/rc.log
