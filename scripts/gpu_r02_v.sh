#!/bin/bash
# Round 2, one GPU, final state: launch list of the C4 steady-state steps (every kernel of append + step), and one
# ncu --set full capture of the incremental GEMM with sixteen new rows (C3).
mkdir -p gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"append|tc_inc|pplanes|sets_kernel|planes_kernel|row_l1|pack_obs|bbox|diag_chol" -c 400 --csv --log-file gpurun_out/launches_inc_c4_r02v.csv python scripts/inc_probe.py C4 > gpurun_out/ncu_inc.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/launches_inc_c4_r02v.csv", errors="ignore")) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows[-200:]:      # the steady-state steps (the fit's diagonal blocks come first)
    name = r[4].split("(")[0][-44:]; v = float(r[-1].replace(",", "")) / 1e3
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]): print(f"{t / c:10.1f} us x {c:3d}  {k}")
PY
timeout 300 ncu --set full --clock-control none --import-source on -k regex:tc_inc -s 4 -c 1 -f -o gpurun_out/ncu_inc_c3_r02v python scripts/inc_probe.py C3 > gpurun_out/ncu_full3.log 2>&1; echo "ncu rc=$?"
