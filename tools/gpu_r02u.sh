# Round-2 call u: onesweep occupancy with the (x, y) keys: 2 / 3 (default) / 4 CTAs per SM.
mkdir -p gpurun_out
for v in default smb2 smb4; do
  echo "variant=$v"
  if [ "$v" = default ]; then unset B200S_LIB; else export B200S_LIB=$PWD/neuralgaussiansplatting_b200/_variants/libb200splat_$v.so; fi
  for c in "C3" "C4 --fwd-only"; do timeout 300 python tools/stage_times.py --cfg $c --steps 20 | cut -c1-330; done
done 2>&1 | tee gpurun_out/r02u_sort_minb.log
