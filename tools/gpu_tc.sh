mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_warp_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'], d['e2e']['ms_per_step'], 'glue', d['config'].get('torch_glue_ms_per_step'), 'graph', d['config'].get('cuda_graph'))
for k,v in d['kernels'].items(): print('%-28s %.3f ms'%(k,v['ms_per_step']))
PY
