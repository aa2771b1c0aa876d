#!/bin/bash
# Tensor-pipe activity of the posterior kernel at C4's observation count (16 384), block-sparse and dense walk, on a
# slice of C4's grid (8192 x 272 cells); plain run first, then one ncu pass with a handful of metrics.
mkdir -p gpurun_out
CMD="python scripts/scale_probe.py 16384"
timeout 300 $CMD > gpurun_out/c4_plain.log 2>&1 || { tail -5 gpurun_out/c4_plain.log; exit 1; }
M=gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.avg.per_second,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed.sum
timeout 900 ncu --metrics $M --clock-control none -k regex:posterior_tc -c 2 --csv --log-file gpurun_out/ncu_c4_skip.csv $CMD > gpurun_out/ncu_c4_skip.log 2>&1
SBO_TC_SKIP=0 timeout 900 ncu --metrics $M --clock-control none -k regex:posterior_tc -c 2 --csv --log-file gpurun_out/ncu_c4_dense.csv $CMD > gpurun_out/ncu_c4_dense.log 2>&1
cat gpurun_out/c4_plain.log; cut -d, -f5,13- gpurun_out/ncu_c4_skip.csv | tail -16; cut -d, -f5,13- gpurun_out/ncu_c4_dense.csv | tail -16
