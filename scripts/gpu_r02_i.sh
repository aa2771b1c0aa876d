#!/bin/bash
# Round 2, one GPU: large fits under the look-ahead schedule on one GPU (W = 1), the pass-based skip decision at C2, the
# whole GPU suite, and the default bench line.
mkdir -p gpurun_out
echo "== factor + group tests"; timeout 900 python -m pytest tests/test_gpu_factor.py tests/test_gpu_group.py -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/pytest_factor_group_r02i.log
echo "== fit timings"; timeout 600 python scripts/scale_probe.py 4096,8192,16384 2>&1 | tail -3 | tee gpurun_out/scale_probe_r02i.log
echo "== C2, skip decision by passes (default) vs dense"
for opt in "" "tc_skip=0"; do
  timeout 600 python bench.py --workload C2 --extras '' --no-cpu-baseline --steps 20 --warmup 5 --options "$opt" > gpurun_out/bench_c2_${opt//[=,]/_}_r02i.json 2>> gpurun_out/bench_c2_r02i.err
  python - "$opt" <<'PY'
import json,sys
opt=sys.argv[1]
d=json.loads(open("gpurun_out/bench_c2_"+opt.replace("=","_").replace(",","_")+"_r02i.json").read().strip().splitlines()[-1])
r=d["roofline"]; p=d["parity"]
print("C2 [%s] ms/step %.4f e2e %.4f exec TF/s %.0f frac %.3f alg x%.3f plan %s | parity mu %.2e s2 %.2e mask out %d"%(opt, d["ms_per_step"], d["e2e"]["ms_per_step"], r["achieved"], r["frac"], r["algorithmic_speedup"], r["plan"], p["mu_rel_max"], p["s2_abs_max_over_sf2"], p["mask_mismatch_outside_band"]))
PY
done
echo "== full suite"; timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_r02i.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -4 | tee gpurun_out/smoke_r02i.log
echo "== default bench"; T0=$(date +%s); timeout 1700 python bench.py > gpurun_out/bench_default_r02i.json 2> gpurun_out/bench_default_r02i.err
echo "rc=$? wall=$(( $(date +%s) - T0 )) s"; tail -2 gpurun_out/bench_default_r02i.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default_r02i.json").read().strip().splitlines()[-1])
print("C4 ms/step %.1f e2e %.1f fit %.1f clk %s launches %s traffic %s frac %.3f"%(d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["fit_ms"], d["clocks"]["sm_mhz"], d["gpu_launches"], d["roofline"]["traffic"], d["roofline"]["frac"]))
print("cpu", d["cpu_baseline"]["value"], "parity", d["parity"]["s2_abs_max_over_sf2"], d["parity"]["mu_rel_max"], "inc", d["e2e_incremental"]["ms_per_step"])
for k,v in d["extras"].items(): print(k, "ms %.3f e2e %s inc %.3f"%(v["ms_per_step"], v["e2e"] and round(v["e2e"]["ms_per_step"],3), v["e2e_incremental"]["ms_per_step"]), v["e2e_incremental"].get("drift_checks"), v["e2e_incremental"].get("refreshed_steps"))
PY
