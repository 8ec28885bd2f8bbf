# warp/loss GPU tests + bench (no CPU arm); prints the per-ABI-call times
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_warp_gpu.py -m gpu -x -q > gpurun_out/pytest_warp.log 2>&1; echo "pytest_warp rc=$?"; tail -5 gpurun_out/pytest_warp.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'], 'glue', d['config'].get('torch_glue_ms_per_step'))
for k,v in d['kernels'].items(): print('%-24s %.3f ms'%(k,v['ms_per_step']))
PY
