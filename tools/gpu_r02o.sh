# Round-2 call o: fast dL/dsh reconstruction: parity of the factored path + kernel time alone (exact / fast, 4-6 CTAs per SM).
mkdir -p gpurun_out
T=${TAG:-r02o}
timeout 600 python -m pytest tests/test_factored_sh_gpu.py tests/test_abi.py -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
for v in default sr5 sr6; do
  echo "variant=$v"
  if [ "$v" = default ]; then unset B200S_LIB; else export B200S_LIB=$PWD/neuralgaussiansplatting_b200/_variants/libb200splat_$v.so; fi
  timeout 300 python tools/time_reconstruct.py
done 2>&1 | tee gpurun_out/${T}_time_reconstruct.log
