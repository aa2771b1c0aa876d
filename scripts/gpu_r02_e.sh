#!/bin/bash
# Round 2, one GPU: the shared fit through the group (members on one device: schedule + exchange code, bit-identity),
# the new group tests, the ncu launch list of a short default-workload bench, DRAM traffic of the full-grid C4 launch,
# and one --set full capture of the pair kernel at C4's observation count.
mkdir -p gpurun_out
echo "== group tests"; timeout 900 python -m pytest tests/test_gpu_group.py -q -m gpu -x 2>&1 | tail -15 | tee gpurun_out/pytest_group_r02e.log
echo "== shared fit probe (one device)"; timeout 600 python scripts/shared_fit_probe.py 1500,4096,16384 2 2>&1 | tail -6 | tee gpurun_out/shared_fit_1dev_r02e.log
timeout 300 python scripts/shared_fit_probe.py 8192 4 2>&1 | tail -3 | tee -a gpurun_out/shared_fit_1dev_r02e.log
CMD="python bench.py --steps 2 --warmup 1 --extras C2 --no-parity --no-cpu-baseline"
echo "== launch list"; timeout 900 $CMD > gpurun_out/ncu_plain_r02e.log 2>&1 &&
timeout 1700 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_bench_c4_r02e.csv $CMD > gpurun_out/ncu_launch_run_r02e.log 2>&1
echo "rc=$?"; grep -c . gpurun_out/launches_bench_c4_r02e.csv
M=gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.avg.per_second,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_bytes.sum
echo "== traffic of the full-grid C4 launch"
timeout 900 ncu --metrics $M --clock-control none -k regex:posterior_tc_pair -c 1 --csv --log-file gpurun_out/ncu_c4_full_r02e.csv python bench.py --steps 1 --warmup 0 --extras '' --no-parity --no-cpu-baseline > gpurun_out/ncu_c4_full_r02e.log 2>&1
cut -d, -f5,13- gpurun_out/ncu_c4_full_r02e.csv | tail -9
echo "== traffic of the C2 / C3 launches"
for w in C2 C3; do
timeout 600 ncu --metrics $M --clock-control none -k regex:posterior_tc_pair -c 1 --csv --log-file gpurun_out/ncu_${w}_full_r02e.csv python bench.py --workload $w --steps 1 --warmup 0 --extras '' --no-parity --no-cpu-baseline > gpurun_out/ncu_${w}_full_r02e.log 2>&1
cut -d, -f5,13- gpurun_out/ncu_${w}_full_r02e.csv | tail -9
done
echo "== set full, C4 slice"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:posterior_tc_pair -c 1 -o gpurun_out/prof_c4_pair_r02e python scripts/scale_probe.py 16384 > gpurun_out/ncu_setfull_r02e.log 2>&1
ls -la gpurun_out/*.ncu-rep
