# round 2, call B (2 GPUs): 2-GPU tests, peer_check timings, bench N=2 both arms + variants
mkdir -p gpurun_out
T=${TAG:-r02b}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_api_gpu.py::test_peer_allreduce_two_ranks_exact tests/test_trainer_gpu.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${T}_pytest.log
timeout 600 $TR --master-port 29541 tools/peer_check.py --iters 10 > gpurun_out/${T}_peer_check.json 2> gpurun_out/${T}_peer_check.err; echo "peer_check rc=$?"; tail -3 gpurun_out/${T}_peer_check.err; cat gpurun_out/${T}_peer_check.json
timeout 600 $TR --master-port 29542 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err; echo "n2 rc=$?"; tail -3 gpurun_out/${T}_bench_n2.err
timeout 600 $TR --master-port 29543 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline --no-render-only --dense-collective > gpurun_out/${T}_bench_n2_dense.json 2> gpurun_out/${T}_bench_n2_dense.err; echo "n2 dense rc=$?"
timeout 600 $TR --master-port 29544 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline --no-render-only --collective nccl > gpurun_out/${T}_bench_n2_nccl.json 2> gpurun_out/${T}_bench_n2_nccl.err; echo "n2 nccl rc=$?"
timeout 600 $TR --master-port 29545 bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref_n2.json 2> gpurun_out/${T}_bench_ref_n2.err; echo "ref n2 rc=$?"
for f in n2 n2_dense n2_nccl ref_n2; do python - <<PY
import json
d=json.load(open("gpurun_out/${T}_bench_$f.json"))
print("$f", d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"].get("collective","")[:60], d.get("collective_check"), (d.get("render_only") or {}).get("value"), ((d.get("render_only") or {}).get("e2e") or {}).get("value"))
PY
done
