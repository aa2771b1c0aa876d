#!/bin/bash
# factorisation checks: GPU factor tests, fit timings, launch list of one fit at n = 1024
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_factor.py -q -m gpu -x 2>&1 | tail -5 | tee gpurun_out/pytest_factor.log
timeout 600 python scripts/scale_probe.py 1024,4096,16384 2>&1 | tail -4 | tee gpurun_out/scale_probe.log
