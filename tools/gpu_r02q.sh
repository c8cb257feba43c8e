# Round-2 call q: last check of the final tree: full GPU suite + one short bench line.
mkdir -p gpurun_out
T=${TAG:-r02q}
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_ours.json 2> gpurun_out/${T}_bench_ours.err; echo "ours rc=$?"; cut -c1-260 gpurun_out/${T}_bench_ours.json
