#!/bin/bash
# Round 2, N GPUs of one box (usage: gpu_r02_g.sh N): shared fit over NCCL (bit-identity + timing), group tests and the
# shared fit through peer copies on distinct devices, then the strong-scaled C4 bench line.
N=${1:-2}
mkdir -p gpurun_out
echo "== shared fit over NCCL, $N ranks"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
  scripts/dist_fit_check.py 300,1500,2500 8192,16384 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -8 | tee gpurun_out/dist_fit_nccl_${N}gpu_r02g.log
echo "== group tests on $N devices"; timeout 900 python -m pytest tests/test_gpu_group.py -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/pytest_group_${N}gpu_r02g.log
echo "== shared fit through the group (peer copies), $N devices"
timeout 600 python scripts/shared_fit_probe.py 8192,16384 $N 2>&1 | tail -3 | tee gpurun_out/shared_fit_group_${N}gpu_r02g.log
echo "== bench, $N GPUs"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 3 --warmup 3 --extras C3 > gpurun_out/bench_c4_${N}gpu_r02g.json 2> gpurun_out/bench_c4_${N}gpu_r02g.err
echo "rc=$?"; tail -c 600 gpurun_out/bench_c4_${N}gpu_r02g.err
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_c4_{N}gpu_r02g.json").read().strip().splitlines()[-1])
print("value %.4g cells/s  ms/step %.1f  e2e ms %.1f (fit %.1f ms shared by %s)  spread %s" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["fit_ms"], d["e2e"]["fit_shared_ranks"], d["ms_posterior_spread"]))
for r in d["per_rank"]: print("  ", r)
print("parity", d["parity"]); print("checksums", d["checksums"]); print("limiter", d["limiter"])
for k,v in d["extras"].items(): print(k, "value %.4g ms %.2f e2e %.2f" % (v["value"], v["ms_per_step"], v["e2e"]["ms_per_step"] if v["e2e"] else -1), v["checksums"], [ (r["rows"], round(r["ms_posterior"],2)) for r in v["per_rank"]])
PY
