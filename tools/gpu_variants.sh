# A/B of build variants on the stage timer: VARIANTS="default scan512 ..." CFGS="C3" bash tools/gpu_variants.sh
mkdir -p gpurun_out
T=${TAG:-var}
for v in ${VARIANTS:-default}; do
  echo "variant=$v"
  if [ "$v" = default ]; then unset B200S_LIB; else export B200S_LIB=$PWD/neuralgaussiansplatting_b200/_variants/libb200splat_$v.so; fi
  for c in ${CFGS:-C3}; do timeout 300 python tools/stage_times.py --cfg $c --steps 20 ${STAGE_ARGS:-} | cut -c1-330; done
  if [ -n "$FWD_CFGS" ]; then for c in $FWD_CFGS; do timeout 300 python tools/stage_times.py --cfg $c --steps 20 --fwd-only | cut -c1-330; done; fi
done 2>&1 | tee gpurun_out/${T}_variants.log
unset B200S_LIB
if [ -n "$EXTRA" ]; then bash -c "$EXTRA"; fi
