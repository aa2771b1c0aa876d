#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/sortprobe.log
for s in none morton rows; do SORT=$s timeout 300 python scripts/acc_probe.py "[sort=$s]" 256,1024,2048,4096 2>&1 | tail -4 | tee -a gpurun_out/sortprobe.log; done
