#!/bin/bash
# One driver for every GPU lease recipe (replaces the one-shot tools/gpu_r02*.sh / _n*.sh scripts of rounds 1-2; their
# command lines live on in `git log -- tools/`). Run on the GPU box through gpurun, from the repo root:
#
#     gpurun --timeout 1500 -- 'TAG=r03d bash tools/lease.sh suite bench ncu_aux'
#
# Every recipe writes gpurun_out/${TAG}_*; copy what should be judged into profiles/. Recipes:
#   suite      full `pytest -m gpu` run                         bench      both bench arms, N = 1
#   noalloc    tests/test_zz_no_alloc.py alone, fresh process   stages     tools/stage_times.py at C2 / C3 / C4 / C3 det
#   launches   ncu launch list of `bench.py --steps 2`          ncu_hot    ncu --set full of the K1..K9 kernels at C3
#   ncu_aux    ncu --set full of the kernels either side: idxmap (K11..K16), fused Adam, fused loss, dL/dsh rebuild
#   nrank      N-rank bench, both arms (NGPU=2|4|8; needs gpurun --gpus N)
#   twogpu     the two-rank pytest cases (needs gpurun --gpus 2)
set -u
mkdir -p gpurun_out
T=${TAG:-lease}
STEPS=${STEPS:-20}
NGPU=${NGPU:-2}

summarise() {   # $1 = report stem (gpurun_out/${T}_xxx), $2 = summary suffix; rest = kernel-name fragments for the stall listing
    local stem=$1 suffix=$2; shift 2
    ncu -i ${stem}.ncu-rep --page raw --csv > ${stem}_raw.csv 2>/dev/null && python tools/ncu_summary.py ${stem}_raw.csv > gpurun_out/${T}_ncu_full_summary_${suffix}.csv
    ncu -i ${stem}.ncu-rep --page source --csv --print-source cuda,sass > ${stem}_source.csv 2>/dev/null
    for k in "$@"; do python tools/ncu_stalls.py ${stem}_source.csv $k 14 | cut -c1-180; echo; done > gpurun_out/${T}_ncu_stalls_${suffix}.txt
    rm -f ${stem}_source.csv ${stem}_raw.csv ${stem}.ncu-rep
}

for recipe in "$@"; do
case $recipe in
suite)
    timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/${T}_pytest.log ;;
noalloc)
    timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${T}_noalloc_alone.log 2>&1; echo "noalloc rc=$?" | tee -a gpurun_out/${T}_noalloc_alone.log ;;
bench)
    timeout 600 python bench.py --impl reference --steps $STEPS --warmup 5 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
    timeout 600 python bench.py --steps $STEPS --warmup 5 > gpurun_out/${T}_bench_ours.json 2> gpurun_out/${T}_bench_ours.err; echo "ours rc=$?" ;;
stages)
    { for c in C2 C3; do timeout 300 python tools/stage_times.py --cfg $c --steps $STEPS; done
      timeout 300 python tools/stage_times.py --cfg C4 --steps $STEPS --fwd-only
      timeout 300 python tools/stage_times.py --cfg C3 --steps $STEPS --det; } > gpurun_out/${T}_stage_times.jsonl 2>&1 ;;
launches)
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_ncu_launches_C3_fwdbwd.csv \
        python bench.py --steps 2 --warmup 1 --no-render-only --no-cpu-baseline > gpurun_out/${T}_ncu_launches.log 2>&1; echo "launches rc=$?" ;;
ncu_hot)
    timeout 300 python tools/stage_times.py --cfg C3 --steps 1 > /dev/null 2>&1 || { echo "target failed without ncu"; continue; }
    timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_full \
        -k regex:"blend_|preprocess_|sort_|sorted_scan|scan_compact|expand_|tile_ranges|tile_order|bwd_prologue" --launch-skip 40 -c 24 \
        python tools/stage_times.py --cfg C3 --steps 1 > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu_hot rc=$?"
    summarise gpurun_out/${T}_full C3 sort_onesweep sorted_scan expand_kernel scan_compact blend_fwd blend_bwd preprocess_bwd preprocess_fwd ;;
ncu_aux)
    timeout 300 python tools/idxmap_times.py --cfg C3 --steps 2 > gpurun_out/${T}_idxmap_times.jsonl 2>&1 || { echo "idxmap target failed without ncu"; continue; }
    timeout 300 python tools/step_parts.py > gpurun_out/${T}_step_parts.json 2>&1 || { echo "step_parts target failed without ncu"; continue; }
    timeout 300 python tools/time_reconstruct.py > gpurun_out/${T}_reconstruct.jsonl 2>&1 || echo "reconstruct target failed without ncu"
    timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_idx -k regex:"idx_" --launch-skip 10 -c 3 \
        python tools/idxmap_times.py --cfg C3 --steps 2 > gpurun_out/${T}_ncu_idx.log 2>&1; echo "ncu idx rc=$?"
    summarise gpurun_out/${T}_idx idxmap_C3 idx_splat idx_resolve idx_feature_backward
    timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_aux -k regex:"adam_step|loss_fwd|loss_bwd" --launch-skip 48 -c 6 \
        python tools/step_parts.py > gpurun_out/${T}_ncu_aux.log 2>&1; echo "ncu adam/loss rc=$?"
    summarise gpurun_out/${T}_aux adam_loss_C3 adam_step loss_fwd loss_bwd
    timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${T}_shr -k regex:"sh_grad_reconstruct" --launch-skip 3 -c 2 \
        python tools/time_reconstruct.py > gpurun_out/${T}_ncu_shr.log 2>&1; echo "ncu reconstruct rc=$?"
    summarise gpurun_out/${T}_shr sh_reconstruct_C3 sh_grad_reconstruct ;;
nrank)
    for impl in reference ours; do
        timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29541 \
            bench.py --impl $impl --gpus $NGPU --steps $STEPS --warmup 5 > gpurun_out/${T}_bench_n${NGPU}_${impl}.json 2> gpurun_out/${T}_bench_n${NGPU}_${impl}.err
        echo "n${NGPU} $impl rc=$?"
    done ;;
twogpu)
    timeout 900 python -m pytest tests -m gpu -q -k "two_rank or two_gpu or peer" > gpurun_out/${T}_pytest_2gpu.log 2>&1; echo "twogpu rc=$?" | tee -a gpurun_out/${T}_pytest_2gpu.log ;;
*)
    echo "unknown recipe: $recipe" ;;
esac
done
ls -la gpurun_out | tail -30
