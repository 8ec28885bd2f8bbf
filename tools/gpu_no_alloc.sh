# the no-allocation entry point's test: alone in a fresh process, then after the whole GPU suite (pytest -x, as the round's end runs it)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?" > gpurun_out/rc_no_alloc.log
tail -4 gpurun_out/pytest_no_alloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_all.log 2>&1; echo "all rc=$?" >> gpurun_out/rc_no_alloc.log
tail -4 gpurun_out/pytest_all.log
cat gpurun_out/rc_no_alloc.log
