mkdir -p gpurun_out
: > gpurun_out/rayvar.log
run() {
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --no-e2e --no-warp > gpurun_out/bench_var.log 2>&1
  python - "$1" <<'PY' >> gpurun_out/rayvar.log
import json,sys
for l in open('gpurun_out/bench_var.log'):
    if l.startswith('{'):
        d=json.loads(l); print(sys.argv[1], 'ms',round(d['ms_per_step'],3), {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
}
timeout 240 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q -k "golden or baseline_size" > gpurun_out/pytest_canary.log 2>&1; rc=$?; echo "canary rc=$rc"
if [ $rc -ne 0 ]; then tail -20 gpurun_out/pytest_canary.log; exit 1; fi
run "bulk ring16 mb2 evict_first"
sed -i -e "s/st_stream_f4(o + v \* ostride, r);/st_stream_f4_hint(o + v * ostride, r, policy);/" -e "s/            int cur_base = INT_MIN;\n            float4 vv\[8\];/XX/" vfdepth_b200/csrc/ray_sample.cu
python - <<'PY'
p='vfdepth_b200/csrc/ray_sample.cu'
s=open(p).read()
s=s.replace("            const size_t ostride = (size_t)W * D * CV;\n","            const size_t ostride = (size_t)W * D * CV;\n            const uint64_t policy = l2_evict_first_policy();\n",1)
open(p,'w').write(s)
PY
run "fwd stores evict_first too"
cat gpurun_out/rayvar.log
tail -5 gpurun_out/bench_var.log | cut -c1-300
