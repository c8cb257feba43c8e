# Round-2 call v: blend kernels at other register / occupancy points (backward 10 / 12 (default) / 14 CTAs per SM, forward 12 / 16).
mkdir -p gpurun_out
for v in default bwd10 bwd14 fwd12; do
  echo "variant=$v"
  if [ "$v" = default ]; then unset B200S_LIB; else export B200S_LIB=$PWD/neuralgaussiansplatting_b200/_variants/libb200splat_$v.so; fi
  timeout 300 python tools/stage_times.py --cfg C3 --steps 20 | cut -c1-330
  if [ "$v" = fwd12 ]; then timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only | cut -c1-330; fi
done 2>&1 | tee gpurun_out/r02v_blend_minb.log
