mkdir -p gpurun_out
N=${N:-8}
for mode in overlap blocking; do
  NCCL_DEBUG=${NCCL_DEBUG:-WARN} timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --allreduce $mode > gpurun_out/n${N}_$mode.json 2> gpurun_out/n${N}_$mode.err; echo "$mode rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/n${N}_$mode.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d.get('clocks'))"
done
grep -i "nvls\|algo" gpurun_out/n${N}_overlap.err | head -5
