#!/bin/bash
# gpurun call A (default) / B (argument "quad").  A: plain quick bench, ncu launch list of the same command, ncu --set
# full of the hot kernels of one update.  B: the D >= 2 quadrature evidence (plain, then ncu --set full of k_kl_grad
# for D = 2 and D = 3).  Results land in gpurun_out/ (kept under 64 MiB per call).
mkdir -p gpurun_out
if [ "$1" = "quad" ]; then
    python tools/quad_profile.py > gpurun_out/r02_quad.json 2> gpurun_out/r02_quad.err &&
    ncu --set full --clock-control none --import-source on -k 'regex:k_kl_grad' -s 2 -c 2 \
        -f -o gpurun_out/r02_prof_quad python tools/quad_profile.py > gpurun_out/r02_ncu_quad.log 2>&1
    tail -2 gpurun_out/r02_ncu_quad.log; ls -la gpurun_out
    exit 0
fi
CMD="python bench.py --quick --no-check --steps 2 --warmup 3"
$CMD > gpurun_out/r02_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu_launch.log 2>&1
$CMD > gpurun_out/r02_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_(kl_grad|mean_tiled|nat2ssm_tiled|chol_tiled)' -s 20 -c 7 \
    -f -o gpurun_out/r02_prof_update $CMD > gpurun_out/r02_ncu_full.log 2>&1
tail -2 gpurun_out/r02_ncu_full.log; ls -la gpurun_out
