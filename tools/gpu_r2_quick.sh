mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_warp_gpu.py tests/test_fusion_gpu.py tests/test_train_step_gpu.py -m gpu -x -q -k "golden or midsize or fused_mlp or full_size_standard or smooth or dropin or chained" > gpurun_out/pytest_quick.log 2>&1; echo "pytest_quick rc=$?" > gpurun_out/rc.log
tail -3 gpurun_out/pytest_quick.log
timeout 600 python bench.py --no-cpu-baseline --no-extras --no-e2e > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
python - <<'PY'
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step']); print({k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
cat gpurun_out/rc.log
