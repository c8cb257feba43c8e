mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v4_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/v4_pytest.log | cut -c1-200
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/v4_bench.json 2> gpurun_out/v4_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/v4_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'full', d['full_iteration']['ms_per_step'], d['roofline']['traffic'])" || tail -5 gpurun_out/v4_bench.err
