#!/bin/bash
# Round 2, one GPU, final state: the default bench again (pass r called /usr/bin/time, which the box does not have).
mkdir -p gpurun_out
echo "== default bench"; SECONDS=0; timeout 1500 python bench.py > gpurun_out/bench_default_r02r.json 2> gpurun_out/bench_default_r02r.err; echo "rc=$? wall ${SECONDS} s"; tail -2 gpurun_out/bench_default_r02r.err
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
