#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_posterior.py -q -m gpu -k "test_step_matches" -s 2>&1 | grep -E "fp32\]|fp64\]|mask mism|passed|failed" | tee gpurun_out/probe.log
