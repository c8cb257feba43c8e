python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for nw in 1 2 8; do
  echo "pytest NW=$nw"; B200S_BLEND_FWD_WARPS=$nw B200S_BLEND_BWD_WARPS=$nw python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -3
done
for nw in 1 2 4 8; do
  echo "NW=$nw"; B200S_BLEND_FWD_WARPS=$nw B200S_BLEND_BWD_WARPS=$nw python tools/stage_times.py --cfg C3 --steps 10
done
