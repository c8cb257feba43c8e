mkdir -p gpurun_out
T=${TAG:-r02n4}
NP=${NP:-4}
run() { name=$1; port=$2; shift 2
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port $port bench.py --gpus $NP --steps 40 --warmup 5 --no-cpu-baseline --no-render-only "$@" > gpurun_out/${T}_$name.json 2> gpurun_out/${T}_$name.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${T}_$name.json").read().strip().splitlines()[-1])
    print("$name", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d.get("collective_check",{}).get("all_ranks_ok"), d.get("collective_check",{}).get("max_rel_diff_vs_nccl"))
except Exception as e: print("$name parse failed", e)
PY
}
run default 29701
run mm2 29702 --peer-ctas 2
run exact6 29703 --exact-collective --peer-ctas 6
