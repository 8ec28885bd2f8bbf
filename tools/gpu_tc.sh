mkdir -p gpurun_out
timeout 120 python tools/kernel_bench.py --only fuse_bwd --iters 5 > gpurun_out/tc_bench.log 2>&1; echo "tc_bench rc=$?"; tail -2 gpurun_out/tc_bench.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:voxel_fuse --csv --log-file gpurun_out/tc_launches.csv python tools/kernel_bench.py --only fuse_bwd --iters 3 > gpurun_out/tc_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/tc_launches.csv')) if len(r)>5]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value'); ui=h.index('Metric Unit')
for r in rows[1:][-6:]: print(r[ki][:60], r[vi], r[ui])
PY
