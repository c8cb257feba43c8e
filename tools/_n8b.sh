mkdir -p gpurun_out
N=${N:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for mode in "--peer-ctas 8 " "--peer-ctas 12 " "--peer-ctas 8 --multimem"; do
  tag=$(echo $mode | tr -d ' -')
  timeout 200 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline $mode > gpurun_out/n${N}_$tag.json 2> gpurun_out/n${N}_$tag.err; echo "$mode rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/n${N}_$tag.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('collective'))" || tail -5 gpurun_out/n${N}_$tag.err
done
