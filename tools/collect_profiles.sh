# copy the results of tools/gpu_final.sh (gpurun_out/) into profiles/ under this round's names
R=${1:-r01}
tail -1 gpurun_out/bench.log > profiles/${R}_bench_line.json
tail -1 gpurun_out/bench_ref.log > profiles/${R}_bench_reference_line.json
cp gpurun_out/kernel_bench.log profiles/${R}_kernel_bench.txt
[ -f gpurun_out/kernel_bench_aux.log ] && cp gpurun_out/kernel_bench_aux.log profiles/${R}_kernel_bench_aux.txt
cp gpurun_out/launches.csv profiles/${R}_launches_bench_steps2.csv
python tools/summarize_ncu.py gpurun_out/prof_step.ncu-rep profiles/${R}_ncu_full_summary.csv profiles/traffic.json 1 > /dev/null
[ -f gpurun_out/bench_n2.log ] && tail -1 gpurun_out/bench_n2.log > profiles/${R}_bench_line_n2.json
ls -la profiles/
