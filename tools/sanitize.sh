# compute-sanitizer passes over the tools/sanitize_target.py workload (T0/T1: every kernel family once).
# Outputs -> gpurun_out/${TAG}_sanitizer_{memcheck,racecheck,synccheck}.log ; summaries are copied to profiles/.
mkdir -p gpurun_out
T=${TAG:-r02}
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 \
    python tools/sanitize_target.py > gpurun_out/${T}_sanitizer_${tool}.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize target ok" gpurun_out/${T}_sanitizer_${tool}.log | tail -3
done
