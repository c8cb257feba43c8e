mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_optimizer_gpu.py -m gpu -x -q > gpurun_out/v2_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/v2_pytest.log
