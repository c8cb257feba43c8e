# Round-2 call l: end-of-round validation of the default build ((x, y) tile keys): full GPU suite (+ the checked build),
# both bench arms, stage times, ncu launch list + full capture with summaries.
mkdir -p gpurun_out
T=${TAG:-r02l}
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
B200S_LIB=$PWD/neuralgaussiansplatting_b200/_variants/libb200splat_checked.so timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_binning_modes_gpu.py > gpurun_out/${T}_checked_build_pytest.log 2>&1; echo "checked pytest rc=$?"; tail -3 gpurun_out/${T}_checked_build_pytest.log
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_ours.json 2> gpurun_out/${T}_bench_ours.err; echo "ours rc=$?"
cut -c1-300 gpurun_out/${T}_bench_ref.json; cut -c1-600 gpurun_out/${T}_bench_ours.json
for c in C2 C3; do timeout 300 python tools/stage_times.py --cfg $c --steps 20; done 2>&1 | tee gpurun_out/${T}_stage_times.jsonl
timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only 2>&1 | tee -a gpurun_out/${T}_stage_times.jsonl
timeout 300 python tools/stage_times.py --cfg C3 --steps 20 --det 2>&1 | tee -a gpurun_out/${T}_stage_times.jsonl
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_ncu_launches_C3_fwdbwd.csv \
  python bench.py --steps 2 --warmup 1 > gpurun_out/${T}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_full \
  -k regex:"blend_|preprocess_|sort_|sorted_scan|scan_compact|expand_|tile_ranges|tile_order|bwd_prologue" --launch-skip 40 -c 24 python tools/stage_times.py --cfg C3 --steps 1 > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_raw.csv 2>/dev/null && python tools/ncu_summary.py gpurun_out/${T}_ncu_raw.csv > gpurun_out/${T}_ncu_full_summary_C3.csv
ncu -i gpurun_out/${T}_full.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${T}_ncu_source.csv 2>/dev/null
for k in sort_onesweep sorted_scan expand_kernel scan_compact blend_fwd blend_bwd preprocess_bwd preprocess_fwd; do python tools/ncu_stalls.py gpurun_out/${T}_ncu_source.csv $k 14 | cut -c1-180; echo; done > gpurun_out/${T}_ncu_stalls.txt
rm -f gpurun_out/${T}_ncu_source.csv gpurun_out/${T}_ncu_raw.csv gpurun_out/${T}_full.ncu-rep
ls -la gpurun_out | head -40
