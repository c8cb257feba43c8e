mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_loss.py tests/test_optimizer_gpu.py -m gpu -x -q > gpurun_out/v3_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/v3_pytest.log | cut -c1-220
timeout 300 python tools/step_parts.py 2>&1 | tee gpurun_out/v3_step_parts.json | tail -3
