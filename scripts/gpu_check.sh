#!/bin/bash
# One gpurun call: the whole GPU suite, accuracy probe, smoke, short bench; logs under gpurun_out/.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -25 | tee gpurun_out/pytest_gpu.log
echo "== accuracy"; timeout 300 python scripts/acc_probe.py "[lib]" 256,1024,2048,4096 tcgen05,simt 2>&1 | tail -8 | tee gpurun_out/accuracy.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -6 | tee gpurun_out/smoke.log
echo "== bench"; timeout 600 python bench.py --steps 10 --warmup 3 2>&1 | tail -3 | tee gpurun_out/bench.log
