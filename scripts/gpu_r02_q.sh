#!/bin/bash
# Round 2, one GPU: L2 prefetch in the epilogue of the incremental GEMM (it was bound by the DRAM latency of its
# read-modify-write) -- its tests, the steady-state steps at C4 / C3, the launch list of the C4 steps.
mkdir -p gpurun_out
echo "== incremental tests"; timeout 900 python -m pytest tests/test_gpu_posterior.py tests/test_gpu_large.py tests/test_gpu_group.py tests/test_gpu_mask_delta.py -q -m gpu -x -k "incremental or gemm or policy or hundreds or ingest or group or delta" 2>&1 | tail -3 | tee gpurun_out/pytest_inc_r02q.log
echo "== steady-state steps"; timeout 900 python scripts/inc_probe.py C4 C3 2>&1 | tee gpurun_out/inc_probe_r02q.log | grep -E "full step|rep 3"
echo "== launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tc_inc|pplanes" -c 100 --csv --log-file gpurun_out/launches_inc_c4_r02q.csv python scripts/inc_probe.py C4 C3 > gpurun_out/ncu_inc.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/launches_inc_c4_r02q.csv", errors="ignore")) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = r[4].split("(")[0][-44:] + r[8]; v = float(r[-1].replace(",", "")) / 1e3
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
for k, (c, t) in agg.items(): print(f"{t / c:10.1f} us x {c:3d}  {k}")
PY
echo "== ncu --set full: incremental GEMM (C4)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_inc -s 2 -c 1 -f -o gpurun_out/ncu_inc_c4_r02q python scripts/inc_probe.py C4 > gpurun_out/ncu_full4.log 2>&1; echo "ncu rc=$?"
