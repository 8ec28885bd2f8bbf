mkdir -p gpurun_out
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?" > gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step']); print({k:v for k,v in d.get('dropin_modules',{}).items() if k!='what'}); print({k:v for k,v in d.get('train_step',{}).items() if k in ('value','ms_per_step')})
PY
grep -i "failed\|error" gpurun_out/bench_quick.log | head -5
cat gpurun_out/rc.log
