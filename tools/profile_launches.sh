#!/bin/bash
# gpurun call 1 of 2: plain bench, then the ncu launch list of the same command.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3"
$CMD > gpurun_out/bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
tail -2 gpurun_out/bench_plain.log | cut -c1-400
