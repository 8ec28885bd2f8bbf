mkdir -p gpurun_out
for e in 0 1 2; do
VFD_EXP=$e timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:voxel_fuse_bwd --csv --log-file gpurun_out/tc_launches_$e.csv python tools/kernel_bench.py --only fuse_bwd --iters 2 > gpurun_out/tc_ncu.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/tc_launches_$e.csv')) if len(r)>5]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value')
for r in rows[1:][-2:]: print('exp $e', r[ki][:60], r[vi])
PY
done
