#!/bin/bash
# Round 2: exactly what the driver runs for the scaling step at N GPUs (default steps / warm-up / extras), with wall time;
# (the reference arm was run at N = 2 in pass j; left out here to save GPU minutes).
N=${1:-2}
mkdir -p gpurun_out
T0=$(date +%s)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
  bench.py --gpus $N > gpurun_out/bench_default_${N}gpu_r02t.json 2> gpurun_out/bench_default_${N}gpu_r02t.err
echo "rc=$? wall=$(( $(date +%s) - T0 )) s"; tail -c 400 gpurun_out/bench_default_${N}gpu_r02t.err
python - $N <<'PY'
import json,sys
N=sys.argv[1]
d=json.loads(open(f"gpurun_out/bench_default_{N}gpu_r02t.json").read().strip().splitlines()[-1])
print("C4 value %.4g ms/step %.1f e2e %.1f fit %.1f shared %s spread %s band_plan %s"%(d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["e2e"]["fit_ms"], d["e2e"]["fit_shared_ranks"], d["ms_posterior_spread"], d.get("band_plan_ms")))
print("checksums", d["checksums"]); print("inc", d["e2e_incremental"]["ms_per_step"])
for k,v in d["extras"].items():
    print(k, "value %.4g ms %.3f e2e %s inc %.3f spread %s"%(v["value"], v["ms_per_step"], v["e2e"] and round(v["e2e"]["ms_per_step"],3), v["e2e_incremental"]["ms_per_step"], v["ms_posterior_spread"]), v["checksums"], v["parity"] and (v["parity"]["mu_rel_max"], v["parity"]["s2_abs_max_over_sf2"], v["parity"]["mask_mismatch_outside_band"]), v["e2e_incremental"].get("end_parity"))
PY
