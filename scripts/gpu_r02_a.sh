#!/bin/bash
# Round 2, first GPU pass: suite, smoke, default bench (C4 strong-scaling line + extras), A/B of the new walk options.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -40 | tee gpurun_out/pytest_gpu_r02a.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -6 | tee gpurun_out/smoke_r02a.log
echo "== bench C4 default"; timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c4_r02a.json 2> gpurun_out/bench_c4_r02a.err; tail -c 3000 gpurun_out/bench_c4_r02a.json; tail -5 gpurun_out/bench_c4_r02a.err
for opt in "tc_graded=0,tc_strip=0" "tc_graded=1,tc_strip=0" "tc_graded=0,tc_strip=4" "tc_graded=1,tc_strip=2" "tc_graded=1,tc_strip=8"; do
  echo "== A/B $opt"
  timeout 600 python bench.py --workload C4 --extras '' --no-parity --no-cpu-baseline --steps 2 --warmup 1 --options "$opt" > gpurun_out/ab_${opt//[=,]/_}_r02a.json 2>> gpurun_out/ab_r02a.err
  python - "$opt" <<'PY'
import json,sys
opt=sys.argv[1]
try:
    d=json.loads(open("gpurun_out/ab_"+opt.replace("=","_").replace(",","_")+"_r02a.json").read().strip().splitlines()[-1])
    r=d["roofline"]
    print(opt, "ms_post %.1f"%d["phases_ms"]["posterior"], "exec TF/s %.0f frac %.3f alg x%.2f"%(r["achieved"], r["frac"], r["algorithmic_speedup"]), "e2e ms %.1f"%d["e2e"]["ms_per_step"], "clk", d["clocks"]["sm_mhz"], d["clocks"]["reasons"], "plan", r["plan"])
except Exception as e:
    print(opt, "failed", e)
PY
done
echo "== fit probe"; timeout 300 python scripts/scale_probe.py 4096,16384 2>&1 | tail -3 | tee gpurun_out/scale_probe_r02a.log
