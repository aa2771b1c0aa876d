#!/bin/bash
# Round 2, one GPU: after batching the incremental GEMM epilogue loads, the block-sparse P planes and the one-CTA append reductions --
# the suite, the steady-state step taken apart at C4 / C3 / C5 (plain, then the ncu launch list of the C4 steps), the default bench.
mkdir -p gpurun_out
echo "== full suite first"; timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -6 | tee gpurun_out/pytest_gpu_r02o.log
echo "== steady-state step taken apart"; timeout 900 python scripts/inc_probe.py C4 C3 C5 2>&1 | tee gpurun_out/inc_probe_r02o.log | tail -30
echo "== launch list of the C4 steady-state steps"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"append|tc_inc|pplanes|sets_kernel|planes_kernel|row_l1|pack_obs|bbox|thin" -c 600 --csv --log-file gpurun_out/launches_inc_c4_r02o.csv python scripts/inc_probe.py C4 > gpurun_out/ncu_inc.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/launches_inc_c4_r02o.csv", errors="ignore")) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = r[4].split("(")[0][-60:]; v = float(r[-1].replace(",", "")); u = r[-2]
    v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v   # -> us
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]: print(f"{t:12.1f} us {c:5d} x {k}")
PY
echo "== default bench"; timeout 1500 python bench.py > gpurun_out/bench_default_r02o.json 2> gpurun_out/bench_default_r02o.err; echo "rc=$?"; tail -2 gpurun_out/bench_default_r02o.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_default_r02o.json").read().strip().splitlines()[-1])
print("C4 ms/step %.1f e2e %.1f frac %.3f wall %.0f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d.get("wall_s", 0)))
for name, v in [("C4", d)] + list(d.get("extras", {}).items()):
    i = v["e2e_incremental"]; c = i.get("changed_segments_only", {})
    print(name, "step %.2f ms | inc %.3f ms (p99 %.3f) | delta %.3f ms, %.0f of %d mask bytes in %.1f copies, equal %s" % (
        v["ms_per_step"], i["ms_per_step"], i["p99_ms"], c.get("ms_per_step", -1), c.get("mask_d2h_bytes_per_step", -1), c.get("mask_bytes_full", -1),
        c.get("copies_per_step", -1), c.get("mask_equals_device_mask")))
PY

