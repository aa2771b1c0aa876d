#!/bin/bash
# One gpurun call: accuracy probe of the fp16x3 tcgen05 kernel as a single CTA and as a CTA pair,
# then the GPU suite and a short bench.  Every step under its own timeout.
mkdir -p gpurun_out
for cg in 1 2; do
  echo "== accuracy CG=$cg"
  SBO_TC_CTA_GROUP=$cg timeout 240 python scripts/acc_probe.py "[cg$cg]" 256,1024,2048,4096 tcgen05 2>&1 | tail -6 | tee gpurun_out/accuracy_cg$cg.log
done
echo "== accuracy simt"; timeout 240 python scripts/acc_probe.py "[simt]" 256,1024 simt 2>&1 | tail -3 | tee gpurun_out/accuracy_simt.log
for cg in 1 2; do
  echo "== bench CG=$cg"; SBO_TC_CTA_GROUP=$cg timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -2 | tee gpurun_out/bench_cg$cg.log
done
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -25 | tee gpurun_out/pytest_gpu.log
