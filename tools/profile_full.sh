#!/bin/bash
# gpurun call 2 of 2: plain bench, then ncu --set full of one whole site update (the hot launches of one step).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3"
$CMD > gpurun_out/bench_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_(kl_grad|kl_value|marg_tiled|mean_tiled|nat2ssm_tiled|chol_tiled|invband_tiled|ssm_to_naturals|patch_naturals)' -s 20 -c 10 \
    -f -o gpurun_out/prof_update $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
