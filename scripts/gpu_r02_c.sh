#!/bin/bash
# Round 2: the allocation-free entry point alone in a fresh process, the group tests, then the whole GPU suite.
mkdir -p gpurun_out
echo "== zz alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -15 | tee gpurun_out/pytest_zz_alone_r02c.log
echo "== group"; timeout 900 python -m pytest tests/test_gpu_group.py -q -m gpu -x -s 2>&1 | tail -30 | tee gpurun_out/pytest_group_r02c.log
echo "== full suite"; timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -30 | tee gpurun_out/pytest_gpu_r02c.log
