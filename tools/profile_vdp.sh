#!/bin/bash
# plain VDP bench, then its ncu launch list
mkdir -p gpurun_out
CMD="python bench.py --workload vdp --steps 2 --warmup 3"
$CMD > gpurun_out/bench_vdp.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_vdp.csv $CMD > gpurun_out/ncu_vdp.log 2>&1
tail -1 gpurun_out/bench_vdp.log | cut -c1-300
