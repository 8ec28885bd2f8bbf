mkdir -p gpurun_out
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?" > gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step'], 'e2e', d['e2e'] and (d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['h2d_bytes_per_step'], d['e2e']['d2h_bytes_per_step'])); print({k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras --e2e-grads > gpurun_out/bench_quick2.log 2>&1; echo "bench2 rc=$?" >> gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick2.log'):
    if l.startswith('{'):
        d=json.loads(l); print('e2e-grads', d['e2e'] and (d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['h2d_bytes_per_step'], d['e2e']['d2h_bytes_per_step']))
PY
python tools/pcie_probe.py > gpurun_out/pcie_probe.log 2>&1; tail -8 gpurun_out/pcie_probe.log
cat gpurun_out/rc.log
