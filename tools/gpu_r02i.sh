# Round-2 call i: counting-rank sort A/B + parity, then the round's validation (tests, both bench arms, ncu).
mkdir -p gpurun_out
T=${TAG:-r02i}
echo "== parity with the counting-rank sort on every pass"
B200S_RANK_COUNT=0xff timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_api_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_rankcount.log 2>&1; echo "rank-count pytest rc=$?"; tail -3 gpurun_out/${T}_pytest_rankcount.log
echo "== stage times per ranking mode"
for m in 0 0xff 0x30 0x0f; do
  echo "RANK_COUNT=$m"
  B200S_RANK_COUNT=$m timeout 300 python tools/stage_times.py --cfg C3 --steps 20 | cut -c1-330
  B200S_RANK_COUNT=$m timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only | cut -c1-330
done 2>&1 | tee gpurun_out/${T}_rank_modes.log
for w in 4; do
  echo "BLEND_BWD_WARPS=$w"; B200S_BLEND_BWD_WARPS=$w timeout 300 python tools/stage_times.py --cfg C3 --steps 20 | cut -c1-330
  echo "BLEND_FWD_WARPS=$w"; B200S_BLEND_FWD_WARPS=$w timeout 300 python tools/stage_times.py --cfg C3 --steps 20 | cut -c1-330
done 2>&1 | tee gpurun_out/${T}_blend_warps.log
echo "== full GPU suite (default build)"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_ours.json 2> gpurun_out/${T}_bench_ours.err; echo "ours rc=$?"
cut -c1-300 gpurun_out/${T}_bench_ref.json; cut -c1-600 gpurun_out/${T}_bench_ours.json
for c in C2 C3; do timeout 300 python tools/stage_times.py --cfg $c --steps 20; done 2>&1 | tee gpurun_out/${T}_stage_times.jsonl
timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only 2>&1 | tee -a gpurun_out/${T}_stage_times.jsonl
timeout 300 python tools/stage_times.py --cfg C3 --steps 20 --det 2>&1 | tee -a gpurun_out/${T}_stage_times.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_ncu_launches_C3_fwdbwd.csv \
  python bench.py --steps 2 --warmup 1 > gpurun_out/${T}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_full \
  -k regex:"blend_|preprocess_|sort_|sorted_scan|scan_compact|expand_kernel|tile_ranges|tile_order|bwd_prologue" --launch-skip 40 -c 24 python tools/stage_times.py --cfg C3 --steps 1 > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_raw.csv 2>/dev/null && python tools/ncu_summary.py gpurun_out/${T}_ncu_raw.csv > gpurun_out/${T}_ncu_full_summary_C3.csv
ncu -i gpurun_out/${T}_full.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${T}_ncu_source.csv 2>/dev/null && python tools/ncu_lines.py gpurun_out/${T}_ncu_source.csv 25 > gpurun_out/${T}_ncu_source_hotlines.txt
rm -f gpurun_out/${T}_ncu_source.csv
ls -la gpurun_out | head -40
