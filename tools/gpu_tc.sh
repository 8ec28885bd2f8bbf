mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_warp_gpu.py -m gpu -x -q -k "depth_syn or to_depth or smooth or dropin" > gpurun_out/pytest_new.log 2>&1; echo "pytest_new rc=$?"; tail -40 gpurun_out/pytest_new.log
