mkdir -p gpurun_out
for v in "" lb16 lb32; do
  if [ -n "$v" ]; then export B200S_LIB=neuralgaussiansplatting_b200/_variants/libb200splat_$v.so; fi
  echo "variant=${v:-default(lb8)}"
  timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -1
  timeout 300 python tools/stage_times.py --cfg C3 --steps 20
  timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only
  timeout 300 python tools/stage_times.py --cfg C2 --steps 20
done 2>&1 | tee gpurun_out/v5_stages.log
