#!/bin/bash
# Rebuild posterior_tc.cu with each macro set on the GPU box and probe its accuracy.
mkdir -p gpurun_out; : > gpurun_out/variants.log
cd safe_bayesian_optimization_b200/csrc
for v in "" "-DSBO_TC_EXP=1" "-DSBO_TC_EXP=2" "-DSBO_TC_EXP=2 -DSBO_TC_MU64" "-DSBO_TC_EXP=2 -DSBO_TC_MU64 -DSBO_TC_TERMS=4" "-DSBO_TC_EXP=2 -DSBO_TC_TERMS=1"; do
  rm -f posterior_tc.o; make -s EXTRA="$v" 2>&1 | grep -E "error" ; 
  (cd ../.. && timeout 300 python scripts/acc_probe.py "[$v]" 256,1024,2048,4096 2>&1 | tail -8 | tee -a gpurun_out/variants.log)
done
rm -f posterior_tc.o; make -s
cd ../.. && timeout 300 python scripts/acc_probe.py "[simt]" 256,1024,2048,4096 simt 2>&1 | tail -5 | tee -a gpurun_out/variants.log
