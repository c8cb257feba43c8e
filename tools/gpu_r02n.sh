# Round-2 call n (8 GPUs): the driver-shaped bench at N = 8 on the end-of-round build.
mkdir -p gpurun_out
T=${TAG:-r02n}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/${T}_bench_n8_ours.json 2> gpurun_out/${T}_bench_n8_ours.err
echo "ours rc=$?"; tail -2 gpurun_out/${T}_bench_n8_ours.err | cut -c1-300
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02n_bench_n8_ours.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "render", d.get("render_only", {}).get("value"), d["config"].get("collective"),
      {k: v for k, v in d.get("collective_check", {}).items() if k != "what"}, d["clocks"])
PY
