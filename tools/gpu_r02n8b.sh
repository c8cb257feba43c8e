mkdir -p gpurun_out
T=${TAG:-r02n8b}
run() { name=$1; np=$2; port=$3; shift 3
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $port bench.py --gpus $np --steps 20 --warmup 5 --no-cpu-baseline "$@" > gpurun_out/${T}_$name.json 2> gpurun_out/${T}_$name.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${T}_$name.json").read().strip().splitlines()[-1])
    print("$name", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d.get("collective_check",{}).get("all_ranks_ok"), (d.get("render_only") or {}).get("value"), ((d.get("render_only") or {}).get("e2e") or {}).get("value"))
except Exception as e: print("$name parse failed", e)
PY
}
run n8 8 29601
run n8_ctas8 8 29602 --no-render-only --peer-ctas 8
run n8_ctas32 8 29603 --no-render-only --peer-ctas 32
run n8_pull 8 29604 --no-render-only --gather-pull
run n4 4 29605 --no-render-only
run n2 2 29606 --no-render-only
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29607 tools/peer_check.py --iters 10 --only-factored > gpurun_out/${T}_peer_check.json 2> gpurun_out/${T}_peer_check.err; echo "peer_check rc=$?"; tail -1 gpurun_out/${T}_peer_check.json | cut -c1-1500
