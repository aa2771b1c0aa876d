#!/bin/bash
# Round 2, one GPU, final state: what the driver runs at round end -- pytest -m gpu -x, smoke(), the default bench, the reference arm.
mkdir -p gpurun_out
echo "== pytest -m gpu -x"; timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_r02r.log
echo "== smoke"; timeout 600 python __graft_entry__.py smoke 2>&1 | tail -5 | tee gpurun_out/smoke_r02r.log
echo "== default bench"; SECONDS=0; timeout 1500 python bench.py > gpurun_out/bench_default_r02r.json 2> gpurun_out/bench_default_r02r.err; echo "rc=$?"; echo "wall ${SECONDS} s"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_default_r02r.json").read().strip().splitlines()[-1])
print("C4 ms/step %.1f e2e %.1f frac %.3f clocks %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["clocks"]))
print("parity", {k: d["parity"][k] for k in ("mu_rel_max", "s2_abs_max_over_sf2", "s2_h1_max", "mask_mismatch_outside_band")})
for name, v in [("C4", d)] + list(d.get("extras", {}).items()):
    i = v["e2e_incremental"]; c = i.get("changed_segments_only", {})
    print(name, "step %.2f ms | inc %.3f ms (p99 %.3f) end_parity %s | delta %.3f ms, %.0f of %d mask bytes, equal %s" % (
        v["ms_per_step"], i["ms_per_step"], i["p99_ms"], i.get("end_parity"), c.get("ms_per_step", -1),
        c.get("mask_d2h_bytes_per_step", -1), c.get("mask_bytes_full", -1), c.get("mask_equals_device_mask")))
PY
echo "== reference arm"; timeout 900 python bench.py --impl reference > gpurun_out/bench_reference_r02r.json 2> gpurun_out/bench_reference_r02r.err; echo "rc=$?"; cut -c1-400 gpurun_out/bench_reference_r02r.json
