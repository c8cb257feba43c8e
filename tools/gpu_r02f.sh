mkdir -p gpurun_out
T=${TAG:-r02f}
for o in 0 1; do echo "TILE_ORDER=$o"; B200S_TILE_ORDER=$o timeout 300 python tools/stage_times.py --cfg C3 --steps 20 | cut -c1-330; B200S_TILE_ORDER=$o timeout 300 python tools/stage_times.py --cfg C4 --steps 20 --fwd-only | cut -c1-330; B200S_TILE_ORDER=$o timeout 300 python tools/stage_times.py --cfg C2 --steps 20 | cut -c1-330; done 2>&1 | tee gpurun_out/${T}_tile_order.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${T}_pytest.log
B200S_LIB=$PWD/neuralgaussiansplatting_b200/_variants/libb200splat_checked.so timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_checked_build_pytest.log 2>&1; echo "checked pytest rc=$?"; tail -4 gpurun_out/${T}_checked_build_pytest.log
timeout 300 python tools/fast_mode_dev.py T1 C1 C2 C3 2>&1 | tee gpurun_out/${T}_fast_mode_dev.jsonl
