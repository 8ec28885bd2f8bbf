mkdir -p gpurun_out
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'], 'glue', d['config'].get('torch_glue_ms_per_step'), 'launches', d['gpu_launches'], 'graph', d['config'].get('cuda_graph'), 'traffic', d['roofline']['traffic'])
for k,v in d['kernels'].items(): print('%-26s %.3f ms'%(k,v['ms_per_step']))
PY
timeout 600 python bench.py --no-cpu-baseline --no-graph --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('eager', d['value'], d['ms_per_step'])"
timeout 900 python -m pytest tests/test_warp_gpu.py -m gpu -x -q 2>&1 | tail -2
