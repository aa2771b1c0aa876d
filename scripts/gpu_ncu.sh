#!/bin/bash
# ncu evidence for bench.py (1 GPU): plain run first, then the launch list, then one full capture of the top kernel.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch_run.log 2>&1
$CMD > gpurun_out/ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:posterior_tc -s 1 -c 1 -o gpurun_out/prof_tc $CMD > gpurun_out/ncu_full_run.log 2>&1
tail -2 gpurun_out/ncu_plain.log; tail -3 gpurun_out/ncu_full_run.log; ls -la gpurun_out | tail -8
