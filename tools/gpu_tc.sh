mkdir -p gpurun_out
timeout 300 python tools/kernel_bench.py --only ray_fwd,ray_bwd --iters 10 > gpurun_out/ray_bench.log 2>&1; tail -2 gpurun_out/ray_bench.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_gpu rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'], 'glue', d['config'].get('torch_glue_ms_per_step'), 'launches', d['gpu_launches'])
for k,v in d['kernels'].items(): print('%-26s %.3f ms'%(k,v['ms_per_step']))
PY
