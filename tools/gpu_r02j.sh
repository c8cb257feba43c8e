# Round-2 call j: grouped look-back / counting rank / (x,y) keys / fused expansion: parity + A/B + ncu of the sort kernels.
mkdir -p gpurun_out
T=${TAG:-r02j}
par() {  # name, env...
  name=$1; shift
  env "$@" timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_$name.log 2>&1; echo "$name pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_$name.log
}
echo "== parity"
par default B200S_DUMMY=0
par allnew B200S_RANK_COUNT=0xff B200S_XY_KEYS=1 B200S_FUSE_EXPAND=1
par count B200S_RANK_COUNT=0xff
par xy B200S_XY_KEYS=1
par xycount B200S_RANK_COUNT=0xff B200S_XY_KEYS=1
par fuse10 B200S_RANK_COUNT=0x10 B200S_XY_KEYS=1 B200S_FUSE_EXPAND=1
par flatlb B200S_LB_GROUPED=0
echo "== stage times per binning mode (LB_GROUPED RANK_COUNT XY_KEYS FUSE_EXPAND)"
for m in "0 0 0 0" "1 0 0 0" "1 0xff 0 0" "1 0x30 0 0" "1 0x0f 0 0" "1 0 1 0" "1 0xff 1 0" "1 0xff 1 1" "1 0x10 1 1" "1 0x1f 1 1"; do
  set -- $m
  echo "LB_GROUPED=$1 RANK_COUNT=$2 XY_KEYS=$3 FUSE_EXPAND=$4"
  B200S_LB_GROUPED=$1 B200S_RANK_COUNT=$2 B200S_XY_KEYS=$3 B200S_FUSE_EXPAND=$4 timeout 300 python tools/stage_times.py --cfg C3 --steps 20 | cut -c1-330
  B200S_LB_GROUPED=$1 B200S_RANK_COUNT=$2 B200S_XY_KEYS=$3 B200S_FUSE_EXPAND=$4 timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only | cut -c1-330
done 2>&1 | tee gpurun_out/${T}_binning_modes.log
echo "== gradient run-to-run noise"
timeout 600 python tools/grad_noise.py 2>&1 | tee gpurun_out/${T}_grad_noise.jsonl
echo "== ncu of the binning kernels, two modes"
for m in "match 0 0 0" "count 0xff 1 1"; do
  set -- $m
  B200S_RANK_COUNT=$2 B200S_XY_KEYS=$3 B200S_FUSE_EXPAND=$4 timeout 600 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_bin_$1 \
    -k regex:"sort_|sorted_scan|scan_compact|expand_|tile_ranges" --launch-skip 36 -c 12 python tools/stage_times.py --cfg C3 --steps 1 --fwd-only > gpurun_out/${T}_ncu_bin_$1.log 2>&1; echo "ncu $1 rc=$?"
  ncu -i gpurun_out/${T}_bin_$1.ncu-rep --page raw --csv > gpurun_out/${T}_raw.csv 2>/dev/null && python tools/ncu_summary.py gpurun_out/${T}_raw.csv > gpurun_out/${T}_ncu_bin_$1_summary.csv
  ncu -i gpurun_out/${T}_bin_$1.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${T}_src.csv 2>/dev/null && python tools/ncu_stalls.py gpurun_out/${T}_src.csv sort_ 24 > gpurun_out/${T}_ncu_bin_$1_stalls.txt
  rm -f gpurun_out/${T}_raw.csv gpurun_out/${T}_src.csv
done
ls -la gpurun_out | head -40
