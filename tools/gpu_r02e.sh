mkdir -p gpurun_out
T=${TAG:-r02e}
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${T}_pytest.log
timeout 300 python tools/fast_mode_dev.py T1 C1 C2 C3 2>&1 | tee gpurun_out/${T}_fast_mode_dev.jsonl
for m in 2 3 4; do echo "PB_MINB=$m"; B200S_PB_MINB=$m timeout 300 python tools/stage_times.py --cfg C3 --steps 20 | cut -c1-330; done 2>&1 | tee gpurun_out/${T}_pb_minb.log
timeout 600 python tools/idxmap_times.py 2>&1 | tee gpurun_out/${T}_idxmap_times.jsonl
TAG=$T bash tools/sanitize.sh
