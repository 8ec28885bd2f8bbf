mkdir -p gpurun_out
for i in 1 2 3; do
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'], d['e2e']['ms_per_step'], 'glue', d['config'].get('torch_glue_ms_per_step'), 'graph', d['config'].get('cuda_graph'))
PY
done
