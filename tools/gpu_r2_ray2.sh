mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q -k "golden" > gpurun_out/pytest_canary.log 2>&1; rc=$?; echo "canary rc=$rc" > gpurun_out/rc.log
tail -3 gpurun_out/pytest_canary.log
if [ $rc -ne 0 ]; then cat gpurun_out/rc.log; exit 1; fi
timeout 600 python -m pytest tests/test_fusion_gpu.py -m gpu -q > gpurun_out/pytest_fusion.log 2>&1; echo "pytest_fusion rc=$?" >> gpurun_out/rc.log
tail -5 gpurun_out/pytest_fusion.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-e2e --no-warp > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step']); print({k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
cat gpurun_out/rc.log
