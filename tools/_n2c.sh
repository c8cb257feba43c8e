mkdir -p gpurun_out
N=${N:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR tools/peer_check.py > gpurun_out/peer_check_n$N.json 2> gpurun_out/peer_check_n$N.err; echo "peer_check rc=$?"
cat gpurun_out/peer_check_n$N.json; tail -5 gpurun_out/peer_check_n$N.err
for mode in "--peer-ctas 16" "--peer-ctas 16 --sparse-collective"; do
  tag=$(echo $mode | tr -d ' -')
  timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline $mode > gpurun_out/n${N}_$tag.json 2> gpurun_out/n${N}_$tag.err; echo "$mode rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/n${N}_$tag.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('collective'))" || tail -5 gpurun_out/n${N}_$tag.err
done
