# Round-2 call k: chain-free scans + flat look-back A/B of (x,y) keys / fused expansion; parity of every new path.
mkdir -p gpurun_out
T=${TAG:-r02k}
par() {  # name, env...
  name=$1; shift
  env "$@" timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_$name.log 2>&1; echo "$name pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_$name.log
}
echo "== parity"
par default B200S_DUMMY=0
par chainfree B200S_CHAINFREE=1
par chainfree_xy B200S_CHAINFREE=1 B200S_XY_KEYS=1
par chainfree_allnew B200S_CHAINFREE=1 B200S_RANK_COUNT=0x10 B200S_XY_KEYS=1 B200S_FUSE_EXPAND=1
echo "== stage times per binning mode (CHAINFREE RANK_COUNT XY_KEYS FUSE_EXPAND)"
for m in "0 0 0 0" "0 0 1 0" "0 0x10 1 1" "1 0 0 0" "1 0 1 0" "1 0x10 1 1" "1 0x20 1 0" "1 0x0f 1 0"; do
  set -- $m
  echo "CHAINFREE=$1 RANK_COUNT=$2 XY_KEYS=$3 FUSE_EXPAND=$4"
  for c in "C3" "C4 --fwd-only" "C2"; do
    B200S_CHAINFREE=$1 B200S_RANK_COUNT=$2 B200S_XY_KEYS=$3 B200S_FUSE_EXPAND=$4 timeout 300 python tools/stage_times.py --cfg $c --steps 20 | cut -c1-330
  done
done 2>&1 | tee gpurun_out/${T}_binning_modes.log
echo "== ncu launch times of the binning kernels, chain-free + (x,y) keys"
B200S_CHAINFREE=1 B200S_XY_KEYS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_ncu_launches_chainfree.csv \
  -k regex:"sort_|sorted_scan|scan_|run_|expand_|tile_ranges|preprocess_fwd" --launch-skip 40 -c 20 python tools/stage_times.py --cfg C3 --steps 1 --fwd-only > /dev/null 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02k_ncu_launches_chainfree.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]; h=rows[hi]
for r in rows[hi+1:]: print(r[h.index('Kernel Name')][:48].ljust(48), r[h.index('Grid Size')].rjust(14), r[h.index('Metric Value')])
PY
ls -la gpurun_out | head -30
