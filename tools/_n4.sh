mkdir -p gpurun_out
N=${N:-4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 200 $TR tools/peer_check.py --quick > gpurun_out/peer_check_quick_n$N.json 2> gpurun_out/peer_check_quick_n$N.err; echo "peer_check rc=$?"; cut -c1-400 gpurun_out/peer_check_quick_n$N.json
timeout 200 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/n${N}_default.json 2> gpurun_out/n${N}_default.err; echo "bench rc=$?"
python -c "
import json,sys
d=json.loads(open('gpurun_out/n${N}_default.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('collective'))" || grep -v "^\*\|OMP" gpurun_out/n${N}_default.err | tail -8
