#!/bin/bash
mkdir -p gpurun_out
CMD="python scripts/fit_only.py 1024"
$CMD > gpurun_out/fit_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_fit.csv $CMD > gpurun_out/fit_ncu.log 2>&1
cat gpurun_out/fit_plain.log
