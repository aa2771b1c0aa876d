#!/bin/bash
# Round 2, one GPU: C3 through bench.py with the end-of-chain check now on for every workload (C3's sixteen-row steps generate and
# contract P in two bands: the block-sparse P planes of every band against a full refresh at full size).
mkdir -p gpurun_out
timeout 600 python bench.py --workload C3 --extras C2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_r02u.json 2> gpurun_out/bench_c3_r02u.err; echo "rc=$?"; tail -2 gpurun_out/bench_c3_r02u.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_c3_r02u.json").read().strip().splitlines()[-1])
for name, v in [("C3", d)] + list(d.get("extras", {}).items()):
    i = v["e2e_incremental"]
    print(name, "step %.2f ms | inc %.3f ms k=%d end_parity %s" % (v["ms_per_step"], i["ms_per_step"], i["k"], i.get("end_parity")))
PY
