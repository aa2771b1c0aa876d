#!/bin/bash
# Round 2, one GPU: block-sparse walk of the incremental GEMM's items -- its tests, incremental-step latency at C3 / C4
# (default vs every chunk), then the whole suite.
mkdir -p gpurun_out
echo "== incremental tests"; timeout 900 python -m pytest tests/test_gpu_posterior.py tests/test_gpu_large.py tests/test_gpu_subgoal.py -q -m gpu -x -k "incremental or gemm or policy or hundreds or ingest" 2>&1 | tail -4 | tee gpurun_out/pytest_inc_r02m.log
for opt in "" "tc_skip=0"; do
  echo "== C4 + C3 incremental loops [$opt]"
  timeout 900 python bench.py --workload C4 --extras C3 --steps 1 --warmup 1 --no-cpu-baseline --no-parity --options "$opt" > gpurun_out/bench_inc_${opt//[=,]/_}_r02m.json 2>> gpurun_out/bench_inc_r02m.err
  python - "$opt" <<'PY'
import json,sys
opt=sys.argv[1]
d=json.loads(open("gpurun_out/bench_inc_"+opt.replace("=","_")+"_r02m.json").read().strip().splitlines()[-1])
for name,v in (("C4",d),("C3",d["extras"]["C3"])):
    i=v["e2e_incremental"]; print(name, "[%s] step %.1f ms | inc ms %.3f p50 %.3f p99 %.3f kernel %s end_parity %s"%(opt, v["ms_per_step"], i["ms_per_step"], i["p50_ms"], i["p99_ms"], i["kernel_used"], i.get("end_parity")))
PY
done
echo "== full suite"; timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_r02m.log
