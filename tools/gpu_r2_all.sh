# round 2: GPU tests (fusion first, then warp, then the rest), smoke, kernel bench, bench -- each under its own timeout
mkdir -p gpurun_out
rm -f gpurun_out/parity_report.jsonl
timeout 900 python -m pytest tests/test_fusion_gpu.py -m gpu -x -q > gpurun_out/pytest_fusion.log 2>&1; echo "pytest_fusion rc=$?" > gpurun_out/rc.log
tail -3 gpurun_out/pytest_fusion.log
timeout 1500 python -m pytest tests/test_warp_gpu.py -m gpu -x -q > gpurun_out/pytest_warp.log 2>&1; echo "pytest_warp rc=$?" >> gpurun_out/rc.log
tail -3 gpurun_out/pytest_warp.log
timeout 900 python -m pytest tests/test_projection_gpu.py tests/test_aggregate_gpu.py tests/test_train_step_gpu.py -m gpu -x -q > gpurun_out/pytest_rest.log 2>&1; echo "pytest_rest rc=$?" >> gpurun_out/rc.log
tail -3 gpurun_out/pytest_rest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.log
tail -4 gpurun_out/smoke.log
timeout 300 python tools/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?" >> gpurun_out/rc.log
cat gpurun_out/kernel_bench.log
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.log
tail -c 1200 gpurun_out/bench.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "bench_ref rc=$?" >> gpurun_out/rc.log
tail -c 600 gpurun_out/bench_ref.log
cat gpurun_out/rc.log
