#!/bin/bash
# Round 2, one GPU: block-sparse walk of the fp64 steady-state (thin 2-D) kernel -- its tests, C5's 50-step loop.
mkdir -p gpurun_out
echo "== thin tests"; timeout 900 python -m pytest tests/test_gpu_posterior.py tests/test_gpu_large.py -q -m gpu -x -s -k "thin or hundreds or c5 or policy" 2>&1 | grep -E "incremental posterior ms|passed|failed|Error|assert|chain" | tail -12 | tee gpurun_out/pytest_thin_r02l.log
echo "== C5"; timeout 1200 python bench.py --workload C5 --extras '' --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c5_r02l.json 2> gpurun_out/bench_c5_r02l.err
echo "rc=$?"; tail -3 gpurun_out/bench_c5_r02l.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_c5_r02l.json").read().strip().splitlines()[-1])
print("C5 ms/step %.1f"%d["ms_per_step"]); print("inc", {k:v for k,v in d["e2e_incremental"].items() if k!="includes"})
PY
echo "== C5 loop, every chunk (tc_skip=0)"; timeout 1200 python bench.py --workload C5 --extras '' --steps 1 --warmup 0 --no-cpu-baseline --no-parity --options tc_skip=0 > gpurun_out/bench_c5_dense_r02l.json 2>> gpurun_out/bench_c5_r02l.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_c5_dense_r02l.json").read().strip().splitlines()[-1])
print("C5 dense ms/step %.1f"%d["ms_per_step"]); print("inc", {k:v for k,v in d["e2e_incremental"].items() if k in ("ms_per_step","p50_ms","p99_ms","end_parity")})
PY
