#!/bin/bash
# Round 2, one GPU: block-sparse walk of the fp64 kernel -- its test, the whole GPU suite, C5 as the bench workload.
mkdir -p gpurun_out
echo "== fp64 skip test"; timeout 900 python -m pytest tests/test_gpu_posterior.py -q -m gpu -x -s -k "f64_chunk_skipping" 2>&1 | grep -E "executed flops|passed|failed|Error|assert" | tail -8 | tee gpurun_out/pytest_f64skip_r02k.log
echo "== full suite"; timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_r02k.log
echo "== C5"; timeout 1200 python bench.py --workload C5 --extras '' --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c5_r02k.json 2> gpurun_out/bench_c5_r02k.err
echo "rc=$?"; tail -3 gpurun_out/bench_c5_r02k.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_c5_r02k.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("C5 ms/step %.1f value %.4g e2e %s | exec TF/s %.2f frac %.3f alg x%.2f | clocks %s"%(d["ms_per_step"], d["value"], d["e2e"] and round(d["e2e"]["ms_per_step"],1), r["achieved"], r["frac"], r.get("algorithmic_speedup",0), d["clocks"]["sm_mhz"]))
print("parity", d["parity"]); print("inc", {k:v for k,v in d["e2e_incremental"].items() if k!="includes"})
PY
