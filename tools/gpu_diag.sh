mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_train_step_gpu.py -m gpu -x -q > gpurun_out/pytest_train.log 2>&1; echo "pytest_train rc=$?"
grep -o "worst_param_grad_err[^}]*" gpurun_out/pytest_train.log | head -2; tail -5 gpurun_out/pytest_train.log
