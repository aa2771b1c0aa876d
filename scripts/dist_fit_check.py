"""Shared fit over NCCL (one process per GPU): run under torch.distributed.run with N >= 2 ranks.
Checks that every rank's L^-1 / alpha equal the replicated fit's bit for bit, then times both at the given sizes.
usage: python -m torch.distributed.run --nproc-per-node N scripts/dist_fit_check.py [n_check,...] [n_time,...]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n_check = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 and sys.argv[1] else [300, 1500, 2500])]
n_time = [int(x) for x in (sys.argv[2].split(",") if len(sys.argv) > 2 and sys.argv[2] else [])]
ok = True
for n in n_check + n_time:
    p = S.make_problem("C4" if n > 8192 else "C3", n=n, W=256, H=16)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision="fp32", device=local, max_obs=n, max_cells=p.m) as ctx:
        uid = [ctx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)
        out = {}
        for mode in (0, 1):
            ctx.set_option("dist_fit", mode)
            ctx.set_observations(p.xy, p.y)           # warm (allocations, NCCL channels)
            dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            ctx.set_observations(p.xy, p.y)
            wall = (time.perf_counter() - t0) * 1e3
            ms, shared = ctx.fit_info()
            assert shared == (world if mode else 0), (shared, mode)
            t = torch.tensor([ms, wall], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            out[mode] = (t.tolist(), ctx.linv() if n in n_check else None, ctx.alpha())
        same = True
        if n in n_check:
            same = bool(np.array_equal(out[0][1], out[1][1]) and np.array_equal(out[0][2], out[1][2]))
        else:
            same = bool(np.array_equal(out[0][2], out[1][2]))
        flag = torch.tensor([0 if same else 1], device="cuda")
        dist.all_reduce(flag)
        ok = ok and flag.item() == 0
        if rank == 0:
            print(f"n={n:6d} world={world} replicated fit {out[0][0][0]:8.2f} ms (host {out[0][0][1]:8.2f}) | shared fit {out[1][0][0]:8.2f} ms "
                  f"(host {out[1][0][1]:8.2f}) | bit-identical on every rank: {flag.item() == 0}", flush=True)
dist.destroy_process_group()
if rank == 0:
    print("dist_fit_check OK" if ok else "dist_fit_check FAILED", flush=True)
sys.exit(0 if ok else 1)
