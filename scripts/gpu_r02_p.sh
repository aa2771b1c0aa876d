#!/bin/bash
# Round 2, one GPU: fp64 one-new-row step on the fp64 tensor path (posterior_thin2d_dmma_kernel) and the P planes with one
# thread per observation -- the suite, the steady-state steps at C4 / C5 (C5 also with the DFMA tile, thin_2d=2), launch lists,
# one ncu --set full capture each of the incremental GEMM kernel (C4) and the DMMA tile (C5), the default bench.
mkdir -p gpurun_out
echo "== thin tests first"; timeout 900 python -m pytest tests/test_gpu_posterior.py -q -m gpu -x -s -k "thin or incremental or gemm" 2>&1 | grep -E "incremental posterior ms|passed|failed|Error|assert" | tail -8
echo "== full suite"; timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -5 | tee gpurun_out/pytest_gpu_r02p.log
echo "== steady-state steps"; timeout 900 python scripts/inc_probe.py C4 C5 2>&1 | tee gpurun_out/inc_probe_r02p.log | grep -E "full step|rep 3"
timeout 600 python scripts/inc_probe.py C5 thin_2d=2 2>&1 | tee -a gpurun_out/inc_probe_r02p.log | grep -E "full step|rep 3"
echo "== launch lists"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"append|tc_inc|pplanes|sets_kernel|row_l1|thin" -c 600 --csv --log-file gpurun_out/launches_inc_c4_r02p.csv python scripts/inc_probe.py C4 > gpurun_out/ncu_inc.log 2>&1; echo "ncu rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"append|tc_inc|pplanes|sets_kernel|row_l1|thin" -c 600 --csv --log-file gpurun_out/launches_inc_c5_r02p.csv python scripts/inc_probe.py C5 > gpurun_out/ncu_inc5.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
for f in ("gpurun_out/launches_inc_c4_r02p.csv", "gpurun_out/launches_inc_c5_r02p.csv"):
    rows = [r for r in csv.reader(open(f, errors="ignore")) if len(r) > 10 and r[0].isdigit()]
    agg = collections.OrderedDict()
    for r in rows:
        name = r[4].split("(")[0][-44:]; v = float(r[-1].replace(",", "")) / 1e3
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
    print(f)
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:10]: print(f"{t / c:10.1f} us x {c:3d}  {k}")
PY
echo "== ncu --set full: incremental GEMM (C4), DMMA tile (C5)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_inc -s 2 -c 1 -f -o gpurun_out/ncu_inc_c4_r02p python scripts/inc_probe.py C4 > gpurun_out/ncu_full4.log 2>&1; echo "ncu rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:thin2d_dmma -s 2 -c 1 -f -o gpurun_out/ncu_thin2d_c5_r02p python scripts/inc_probe.py C5 > gpurun_out/ncu_full5.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
echo "== default bench"; timeout 1500 python bench.py > gpurun_out/bench_default_r02p.json 2> gpurun_out/bench_default_r02p.err; echo "rc=$?"; tail -2 gpurun_out/bench_default_r02p.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_default_r02p.json").read().strip().splitlines()[-1])
print("C4 ms/step %.1f e2e %.1f frac %.3f" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"]))
for name, v in [("C4", d)] + list(d.get("extras", {}).items()):
    i = v["e2e_incremental"]; c = i.get("changed_segments_only", {})
    print(name, "step %.2f ms | inc %.3f ms (p99 %.3f) end_parity %s | delta %.3f ms, %.0f of %d mask bytes, equal %s" % (
        v["ms_per_step"], i["ms_per_step"], i["p99_ms"], i.get("end_parity", {}).get("s2_abs_max_over_sf2"), c.get("ms_per_step", -1),
        c.get("mask_d2h_bytes_per_step", -1), c.get("mask_bytes_full", -1), c.get("mask_equals_device_mask")))
PY
