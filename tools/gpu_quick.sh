# quick 1-GPU iteration: tests (-x), stage times, optional extras via $EXTRA
mkdir -p gpurun_out
T=${TAG:-q}
timeout 1200 python -m pytest tests -m gpu -x -q ${PYTEST_ARGS:-} > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/${T}_pytest.log
for c in ${CFGS:-C3}; do timeout 300 python tools/stage_times.py --cfg $c --steps 20; done 2>&1 | tee gpurun_out/${T}_stages.log
timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only 2>&1 | tee -a gpurun_out/${T}_stages.log
if [ -n "$EXTRA" ]; then bash -c "$EXTRA"; fi
