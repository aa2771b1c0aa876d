#!/bin/bash
# Round 2, evidence pass on one GPU: the default bench line with its wall time, the ncu launch list of the same
# command (short form), and the C4-geometry ncu metrics (tensor pipe, DRAM bytes) of the pair kernel: block-sparse walk
# in strip-major order (default), raster order, and the dense walk.
mkdir -p gpurun_out
echo "== default bench"; T0=$(date +%s); timeout 1700 python bench.py > gpurun_out/bench_default_r02d.json 2> gpurun_out/bench_default_r02d.err
echo "rc=$? wall=$(( $(date +%s) - T0 )) s" | tee gpurun_out/bench_default_r02d.time; tail -3 gpurun_out/bench_default_r02d.err
echo "== reference arm"; T0=$(date +%s); timeout 900 python bench.py --impl reference > gpurun_out/bench_reference_r02d.json 2> gpurun_out/bench_reference_r02d.err
echo "rc=$? wall=$(( $(date +%s) - T0 )) s" | tee gpurun_out/bench_reference_r02d.time; cat gpurun_out/bench_reference_r02d.json | cut -c1-600
CMD="python bench.py --steps 2 --warmup 1 --extras C2 --no-parity --no-cpu-baseline"
echo "== launch list"; timeout 900 $CMD > gpurun_out/ncu_plain_r02d.log 2>&1 &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_c4_r02d.csv $CMD > gpurun_out/ncu_launch_run_r02d.log 2>&1
echo "rc=$?"; tail -n 3 gpurun_out/launches_bench_c4_r02d.csv | cut -c1-300
M=gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.avg.per_second,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__inst_executed.sum
for v in "strip:" "raster:tc_strip=0" "dense:tc_skip=0"; do
  tag=${v%%:*}; opt=${v#*:}
  echo "== ncu C4 slice $tag"
  timeout 300 python scripts/scale_probe.py 16384 "$opt" > gpurun_out/c4_plain_${tag}_r02d.log 2>&1 || { tail -5 gpurun_out/c4_plain_${tag}_r02d.log; continue; }
  cat gpurun_out/c4_plain_${tag}_r02d.log
  timeout 900 ncu --metrics $M --clock-control none -k regex:posterior_tc -c 2 --csv --log-file gpurun_out/ncu_c4_${tag}_r02d.csv python scripts/scale_probe.py 16384 "$opt" > gpurun_out/ncu_c4_${tag}_r02d.log 2>&1
  cut -d, -f5,13- gpurun_out/ncu_c4_${tag}_r02d.csv | tail -18
done
