mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/v_pytest.log
timeout 300 python tools/stage_times.py --cfg C3 --steps 20 2>&1 | tee gpurun_out/v_stages.log
timeout 300 python tools/stage_times.py --cfg C2 --steps 20 2>&1 | tee -a gpurun_out/v_stages.log
timeout 300 python tools/stage_times.py --cfg C3 --steps 20 --det 2>&1 | tee -a gpurun_out/v_stages.log
timeout 300 python tools/parity_report.py > gpurun_out/v_parity.log 2>&1; tail -15 gpurun_out/v_parity.log
