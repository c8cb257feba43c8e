# One fat GPU call: tests, both bench arms, stage times, ncu launch list + full capture. Outputs -> gpurun_out/
mkdir -p gpurun_out
T=${TAG:-r01b}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_ours.json 2> gpurun_out/${T}_bench_ours.err; echo "ours rc=$?"
cat gpurun_out/${T}_bench_ref.json | cut -c1-400; cat gpurun_out/${T}_bench_ours.json | cut -c1-1500
for c in C2 C3; do timeout 300 python tools/stage_times.py --cfg $c --steps 20; done 2>&1 | tee gpurun_out/${T}_stages.log
timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only 2>&1 | tee -a gpurun_out/${T}_stages.log
timeout 300 python tools/stage_times.py --cfg C3 --steps 20 --det 2>&1 | tee -a gpurun_out/${T}_stages.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_ncu_launches.csv \
  python bench.py --steps 2 --warmup 1 > gpurun_out/${T}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_full \
  -k regex:"blend_|preprocess_|sort_|sorted_scan|scan_compact|expand_kernel|tile_ranges" --launch-skip 40 -c 40 python tools/stage_times.py --cfg C3 --steps 1 > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
