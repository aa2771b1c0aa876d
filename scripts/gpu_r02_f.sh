#!/bin/bash
# Round 2, one GPU: the phased full refresh (SBO_OPT_TC_PHASES) -- bit-identity test, timing sweep on the C4 slice,
# DRAM / L2 metrics of the default (auto) phase count, then the C4 headline with it.
mkdir -p gpurun_out
echo "== phased refresh test"; timeout 600 python -m pytest tests/test_gpu_posterior.py -q -m gpu -x -k "phased or eager or chunk_skipping" 2>&1 | tail -5 | tee gpurun_out/pytest_phased_r02f.log
for f in 1 2 4 8 16 32 0; do
  echo "== phases $f"; timeout 300 python scripts/scale_probe.py 16384 "tc_phases=$f" 2>&1 | tail -1 | tee -a gpurun_out/phases_sweep_r02f.log
done
M=gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_elapsed.avg.per_second,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum
echo "== ncu, auto phases, C4 slice"
timeout 900 ncu --metrics $M --clock-control none -k regex:posterior_tc_pair -c 40 --csv --log-file gpurun_out/ncu_c4_phased_r02f.csv python scripts/scale_probe.py 16384 > gpurun_out/ncu_c4_phased_r02f.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/ncu_c4_phased_r02f.csv")) if len(r)>10 and r[0].isdigit()]
agg=collections.defaultdict(float); ids=set()
for r in rows:
    ids.add(r[0]); agg[r[-3]]+=float(r[-1].replace(",",""))
print(len(ids),"launches captured (2 posteriors x phases)"); [print("  sum", k, "%.4g"%v) for k,v in agg.items()]
PY
echo "== C4 headline, default options"; timeout 900 python bench.py --steps 3 --warmup 3 --extras C3 --no-cpu-baseline > gpurun_out/bench_c4_r02f.json 2> gpurun_out/bench_c4_r02f.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_c4_r02f.json").read().strip().splitlines()[-1])
print("C4 ms/step %.1f e2e %.1f fit %.1f clk %s launches %s"%(d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["fit_ms"], d["clocks"]["sm_mhz"], d["gpu_launches"]))
print("parity", d["parity"]["s2_abs_max_over_sf2"], d["parity"]["mu_rel_max"], d["parity"]["mask_mismatch_outside_band"], "checksums", d["checksums"])
c3=d["extras"]["C3"]; print("C3 ms/step %.1f e2e %.1f"%(c3["ms_per_step"], c3["e2e"]["ms_per_step"]), c3["checksums"])
PY
tail -3 gpurun_out/bench_c4_r02f.err
