for m in 2 3 4; do echo "PB_MINB=$m"; B200S_PB_MINB=$m python tools/stage_times.py --cfg C3 --steps 10; done
