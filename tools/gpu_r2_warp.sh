mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_warp_gpu.py -m gpu -x -q > gpurun_out/pytest_warp.log 2>&1; echo "pytest_warp rc=$?" > gpurun_out/rc.log
tail -6 gpurun_out/pytest_warp.log
timeout 600 python bench.py --no-cpu-baseline --no-extras --no-e2e > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step']); print({k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
cat gpurun_out/rc.log
