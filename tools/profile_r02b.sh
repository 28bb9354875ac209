#!/bin/bash
# gpurun calls for the round-2 (second half) profiles.  A (default): plain quick bench, ncu launch list of the same
# command, ncu --set full of the fused forward kernel.  B ("sweeps"): ncu --set full of the three sweep kernels.
# Results land in gpurun_out/ (kept under 64 MiB per call).
mkdir -p gpurun_out
CMD="python bench.py --quick --no-check --steps 2 --warmup 3"
if [ "$1" = "sweeps" ]; then
    $CMD > gpurun_out/r02b_plain3.log 2>&1 &&
    ncu --set full --clock-control none -k 'regex:k_(mean_tiled|nat2ssm_tiled|chol_tiled)' -s 12 -c 4 \
        -f -o gpurun_out/r02b_sweeps $CMD > gpurun_out/r02b_ncu_sweeps.log 2>&1
    # (the report of four kernels is > 64 MiB: keep its raw page only)
    ncu -i gpurun_out/r02b_sweeps.ncu-rep --page raw --csv > gpurun_out/r02b_sweeps_raw.csv 2>/dev/null
    rm -f gpurun_out/r02b_sweeps.ncu-rep
    tail -2 gpurun_out/r02b_ncu_sweeps.log; ls -la gpurun_out | tail -4
    exit 0
fi
$CMD > gpurun_out/r02b_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches.csv $CMD > gpurun_out/r02b_ncu_launch.log 2>&1
$CMD > gpurun_out/r02b_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:k_fwd_fused' -s 3 -c 1 \
    -f -o gpurun_out/r02b_fused $CMD > gpurun_out/r02b_ncu_fused.log 2>&1
tail -2 gpurun_out/r02b_ncu_fused.log; ls -la gpurun_out | tail -6
