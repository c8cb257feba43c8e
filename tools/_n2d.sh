mkdir -p gpurun_out
N=${N:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR tools/peer_check.py > gpurun_out/peer_check_n$N.json 2> gpurun_out/peer_check_n$N.err; echo "peer_check rc=$?"
cat gpurun_out/peer_check_n$N.json; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/peer_check_n$N.err | tail -5
