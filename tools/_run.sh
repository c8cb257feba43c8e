for o in 1 0; do echo "ORDER=$o"; 
B200S_TILE_ORDER=$o python tools/stage_times.py --cfg C3 --steps 20
B200S_TILE_ORDER=$o python tools/stage_times.py --cfg C2 --steps 20
B200S_TILE_ORDER=$o python tools/stage_times.py --cfg C4 --steps 20 --fwd-only
done
