#!/bin/bash
# Round 2, one GPU, last pass: P planes by the row recurrence (two exp per observation and 16 global rows) -- the whole suite
# as the driver runs it, then the steady-state steps at C4 / C3.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/pytest_gpu_r02w.log
timeout 200 python scripts/inc_probe.py C4 C3 2>&1 | tee gpurun_out/inc_probe_r02w.log | grep -E "full step|rep 3"
