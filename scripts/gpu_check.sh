#!/bin/bash
# One gpurun call: staged GPU checks with per-stage timeouts; logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== sets/select/factor"; timeout 600 python -m pytest tests/test_gpu_sets.py tests/test_gpu_factor.py -q -m gpu -x 2>&1 | tail -25 | tee gpurun_out/t1.log
echo "== posterior simt"; timeout 900 python -m pytest tests/test_gpu_posterior.py -q -m gpu -k "simt and not tcgen05 and test_step_matches" -s 2>&1 | grep -E "^\[|passed|failed|Error|error|assert" | tail -40 | tee gpurun_out/t2.log
echo "== posterior tcgen05"; timeout 600 python -m pytest tests/test_gpu_posterior.py -q -m gpu -k "tcgen05 and test_step_matches" -s 2>&1 | grep -E "^\[|    mask|passed|failed|Error|error|assert|Timeout" | tail -60 | tee gpurun_out/t3.log
echo "== rest"; timeout 900 python -m pytest tests/test_gpu_posterior.py -q -m gpu -k "not test_step_matches" 2>&1 | tail -30 | tee gpurun_out/t4.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -12 | tee gpurun_out/smoke.log
echo "== bench"; timeout 600 python bench.py --steps 5 --warmup 3 2>&1 | tail -5 | tee gpurun_out/bench.log
