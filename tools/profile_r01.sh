#!/bin/bash
# One gpurun call: plain bench -> ncu launch list -> ncu --set full of one whole site update (12 hot launches).
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3"
$CMD > gpurun_out/bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
$CMD > gpurun_out/bench_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_(kl_grad|kl_value|marg_tiled|nat2ssm_tiled|chol_tiled|invband_tiled|ssm_to_naturals)' -s 24 -c 12 \
    -f -o gpurun_out/prof_update $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
