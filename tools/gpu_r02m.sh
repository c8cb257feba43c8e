# Round-2 call m (2 GPUs): the two-rank GPU tests that skip on a one-GPU lease + the driver-shaped bench at N = 2, both arms.
mkdir -p gpurun_out
T=${TAG:-r02m}
timeout 600 python -m pytest tests/test_api_gpu.py tests/test_trainer_gpu.py tests/test_factored_sh_gpu.py -m gpu -q -k "two_rank or two_ranks or peer or factored" > gpurun_out/${T}_pytest_2gpu.log 2>&1; echo "2-GPU pytest rc=$?"; tail -3 gpurun_out/${T}_pytest_2gpu.log
for impl in ours reference; do
  extra=""; [ $impl = reference ] && extra="--impl reference"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 20 --warmup 5 $extra > gpurun_out/${T}_bench_n2_$impl.json 2> gpurun_out/${T}_bench_n2_$impl.err
  echo "$impl rc=$?"; tail -2 gpurun_out/${T}_bench_n2_$impl.err | cut -c1-300
done
python - <<'PY'
import json
for impl in ("ours", "reference"):
    try:
        d = json.loads(open(f"gpurun_out/r02m_bench_n2_{impl}.json").read().strip().splitlines()[-1])
        print(impl, d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "render", d.get("render_only", {}).get("value"),
              {k: v for k, v in d.get("collective_check", {}).items() if k != "what"})
    except Exception as e:
        print(impl, "parse failed", e)
PY
ls -la gpurun_out | head
