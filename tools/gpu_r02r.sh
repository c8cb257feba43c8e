# Round-2 call r: deeper look-back windows for the depth-sort passes only (all of their tiles are co-resident).
mkdir -p gpurun_out
T=${TAG:-r02r}
for v in default dd16 dd32; do
  echo "variant=$v"
  if [ "$v" = default ]; then unset B200S_LIB; else export B200S_LIB=$PWD/neuralgaussiansplatting_b200/_variants/libb200splat_$v.so; fi
  for c in "C3" "C4 --fwd-only" "C2"; do timeout 300 python tools/stage_times.py --cfg $c --steps 20 | cut -c1-330; done
  if [ "$v" != default ]; then timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py -m gpu -x -q 2>&1 | tail -1; fi
done 2>&1 | tee gpurun_out/${T}_lb_depth_ds.log
