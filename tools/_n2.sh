mkdir -p gpurun_out
for impl in reference ours; do
  extra=""; [ $impl = reference ] && extra="--impl reference"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 $extra > gpurun_out/n2_$impl.json 2> gpurun_out/n2_$impl.err; echo "$impl rc=$?"
  cut -c1-900 gpurun_out/n2_$impl.json; tail -3 gpurun_out/n2_$impl.err
done
