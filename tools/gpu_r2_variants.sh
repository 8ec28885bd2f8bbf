# A/B of kernel-tuning variants in ONE box session: for every define set in VARIANTS (separated by ';') rebuild the library,
# run the warp parity tests named by TESTS (-k expression) and a short resident bench; print the per-call times.
#   VARIANTS="-DVFD_LOSS_THREADS=256;-DVFD_LOSS_THREADS=320" TESTS="golden or full_size" bash tools/gpu_r2_variants.sh
mkdir -p gpurun_out
: > gpurun_out/rc.log
IFS=';' read -ra VS <<< "${VARIANTS:-;}"
i=0
for v in "${VS[@]}"; do
  i=$((i+1))
  export VFD_NVCC_DEFINES="$v"
  echo "=== variant $i: '$v'" | tee -a gpurun_out/rc.log
  python -m vfdepth_b200.build > gpurun_out/build_$i.log 2>&1 || { echo "build $i failed" | tee -a gpurun_out/rc.log; tail -5 gpurun_out/build_$i.log; continue; }
  grep -A2 "warp_loss_bwd_kernelILb1" gpurun_out/build_$i.log | grep -E "spill|Used" 
  if [ -n "$TESTS" ]; then
    timeout 900 python -m pytest ${TESTFILES:-tests/test_warp_gpu.py} -m gpu -x -q -k "$TESTS" > gpurun_out/pytest_$i.log 2>&1; echo "pytest $i rc=$?" | tee -a gpurun_out/rc.log
    tail -3 gpurun_out/pytest_$i.log
  fi
  timeout 400 python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline --no-extras --no-e2e > gpurun_out/bench_$i.log 2>&1; echo "bench $i rc=$?" | tee -a gpurun_out/rc.log
  python - <<PY
import json
for l in open('gpurun_out/bench_$i.log'):
    if l.startswith('{'):
        d=json.loads(l); print('value',round(d['value'],1),'ms',round(d['ms_per_step'],3)); print({k.replace('vfd_',''):round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
done
cat gpurun_out/rc.log
