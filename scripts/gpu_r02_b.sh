#!/bin/bash
# Round 2: the strong-scaled C4 line on N GPUs of one box (usage: gpu_r02_b.sh N [extras]).
N=${1:-2}; EX=${2:-C3}
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 3 --warmup 3 --extras "$EX" > gpurun_out/bench_c4_${N}gpu_r02b.json 2> gpurun_out/bench_c4_${N}gpu_r02b.err
echo "rc=$?"; tail -c 1500 gpurun_out/bench_c4_${N}gpu_r02b.err
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_c4_{N}gpu_r02b.json").read().strip().splitlines()[-1])
print("value %.4g cells/s  ms/step %.1f  e2e ms %.1f  spread %s" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["ms_posterior_spread"]))
for r in d["per_rank"]: print("  ", r)
print("parity", d["parity"]); print("checksums", d["checksums"]); print("limiter", d["limiter"])
for k,v in d["extras"].items(): print(k, "value %.4g ms %.2f e2e %.2f" % (v["value"], v["ms_per_step"], v["e2e"]["ms_per_step"] if v["e2e"] else -1), v["checksums"], [ (r["rows"], round(r["ms_posterior"],2)) for r in v["per_rank"]])
PY
