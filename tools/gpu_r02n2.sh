mkdir -p gpurun_out
T=${TAG:-r02n2}
run() { name=$1; port=$2; shift 2
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $port bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline --no-render-only "$@" > gpurun_out/${T}_$name.json 2> gpurun_out/${T}_$name.err
  echo "$name rc=$?"; tail -2 gpurun_out/${T}_$name.err | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${T}_$name.json").read().strip().splitlines()[-1])
    print("$name", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], {k:v for k,v in d.get("collective_check",{}).items() if k!="what"})
except Exception as e: print("$name parse failed", e)
PY
}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29700 tools/peer_check.py --quick > gpurun_out/${T}_peer_quick.json 2> gpurun_out/${T}_peer_quick.err; echo "peer quick rc=$?"; tail -1 gpurun_out/${T}_peer_quick.json | cut -c1-600; tail -3 gpurun_out/${T}_peer_quick.err | cut -c1-300
run push16 29701
run pull16 29702 --gather-pull
run push16_cprio 29703 --compute-priority
run push8 29704 --peer-ctas 8
run push32 29705 --peer-ctas 32
run push8_cprio 29706 --peer-ctas 8 --compute-priority
