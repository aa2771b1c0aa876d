#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/vprobe.log
cd safe_bayesian_optimization_b200/csrc
for v in "" "-DSBO_TC_TERMS=1" "-DSBO_TC_TERMS=4"; do
  rm -f posterior_tc.o; make -s EXTRA="$v" 2>&1 | grep -E " error" ;
  (cd ../.. && timeout 300 python scripts/v_probe.py "[$v]" 2>&1 | tail -4 | tee -a gpurun_out/vprobe.log)
done
rm -f posterior_tc.o; make -s
