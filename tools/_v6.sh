mkdir -p gpurun_out
timeout 300 python tools/train_synthetic.py --cfg T1 --iters 60 2>&1 | tail -3
timeout 300 python tools/train_synthetic.py --cfg C1 --iters 100 2>&1 | tail -3
