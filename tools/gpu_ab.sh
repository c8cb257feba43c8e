# A/B of one env toggle on the stage timer:  VAR=B200S_TILE_ORDER VALS="0 1" CFGS="C3 C2" bash tools/gpu_ab.sh
mkdir -p gpurun_out
T=${TAG:-ab}
for v in ${VALS:-0 1}; do
  echo "${VAR}=$v"
  for c in ${CFGS:-C3}; do env ${VAR}=$v timeout 300 python tools/stage_times.py --cfg $c --steps 20 | cut -c1-330; done
done 2>&1 | tee gpurun_out/${T}_ab.log
if [ -n "$PYTEST" ]; then timeout 1500 python -m pytest tests -m gpu -q -x ${PYTEST_ARGS:-} > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${T}_pytest.log; fi
if [ -n "$EXTRA" ]; then bash -c "$EXTRA"; fi
