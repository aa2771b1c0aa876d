#!/bin/bash
# Round 2, one GPU: the DMMA tile routine of the fit (factor_dmma.cuh) -- factor / group tests, fit timings, then the whole
# GPU suite; DRAM traffic of the phased full-grid C4 posterior for profiles/ncu_traffic_r02.json.
mkdir -p gpurun_out
echo "== factor + group tests"; timeout 900 python -m pytest tests/test_gpu_factor.py tests/test_gpu_group.py -q -m gpu -x 2>&1 | tail -6 | tee gpurun_out/pytest_factor_group_r02h.log
echo "== fit timings"; timeout 600 python scripts/scale_probe.py 4096,8192,16384 2>&1 | tail -3 | tee gpurun_out/scale_probe_r02h.log
echo "== shared fit probe (one device, 2 members)"; timeout 600 python scripts/shared_fit_probe.py 4096,16384 2 2>&1 | tail -3 | tee gpurun_out/shared_fit_1dev_r02h.log
echo "== full suite"; timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -8 | tee gpurun_out/pytest_gpu_r02h.log
M=gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_elapsed.avg.per_second,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum
echo "== traffic of the phased full-grid C4 posterior"
timeout 900 ncu --metrics $M --clock-control none -k regex:posterior_tc_pair -c 5 --csv --log-file gpurun_out/ncu_c4_full_phased_r02h.csv python bench.py --steps 1 --warmup 0 --extras '' --no-parity --no-cpu-baseline > gpurun_out/ncu_c4_full_phased_r02h.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/ncu_c4_full_phased_r02h.csv")) if len(r)>10 and r[0].isdigit()]
agg=collections.defaultdict(float); ids=set()
for r in rows:
    ids.add(r[0]); agg[r[-3]]+=float(r[-1].replace(",",""))
print(len(ids),"launches (phases of one posterior)"); [print("  sum", k, "%.5g"%v) for k,v in agg.items()]
PY
