# round 2, call A (1 GPU): tests, SM peaks, both bench arms, per-view times
mkdir -p gpurun_out
T=${TAG:-r02a}
tools/_bin/sm_peaks > gpurun_out/${T}_sm_peaks.json 2>&1; cat gpurun_out/${T}_sm_peaks.json
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_ours.json 2> gpurun_out/${T}_bench_ours.err; echo "ours rc=$?"; tail -5 gpurun_out/${T}_bench_ours.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"; tail -5 gpurun_out/${T}_bench_ref.err
cut -c1-600 gpurun_out/${T}_bench_ref.json; cut -c1-3000 gpurun_out/${T}_bench_ours.json
for v in 0 1 2 3 4 5 6 7; do timeout 120 python tools/stage_times.py --cfg C3 --steps 10 --view $v; done 2>&1 | tee gpurun_out/${T}_views.log | cut -c1-120
