python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/stage_times.py --cfg C3 --steps 10
python tools/stage_times.py --cfg C2 --steps 10
python tools/stage_times.py --cfg C4 --steps 10 --fwd-only
