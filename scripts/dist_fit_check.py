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
        # bit for bit where the replicated fit runs the same 128 x 128 DMMA tiles (n > 2944); to rounding against the 64 x 64
        # FMA tiles of a small replicated fit
        def eq(a, b):
            return bool(np.array_equal(a, b)) if n > 2944 else bool(np.abs(a - b).max() <= 1e-9 * np.abs(b).max())
        same = eq(out[1][2], out[0][2]) and (eq(out[1][1], out[0][1]) if n in n_check else True)
        flag = torch.tensor([0 if same else 1], device="cuda")
        dist.all_reduce(flag)
        ok = ok and flag.item() == 0
        if rank == 0:
            print(f"n={n:6d} world={world} replicated fit {out[0][0][0]:8.2f} ms (host {out[0][0][1]:8.2f}) | shared fit {out[1][0][0]:8.2f} ms "
                  f"(host {out[1][0][1]:8.2f}) | {'bit-identical' if n > 2944 else 'equal to rounding'} on every rank: {flag.item() == 0}", flush=True)
# a block that is not positive definite in the middle of a shared fit: every rank must report the same failure and go back
# to the observations it held before (the status travels in the header of the panel message)
p = S.make_problem("C3", n=600, W=256, H=12)
p.xy = p.xy * 12.0
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=1e-20, prior_mean=p.prior_mean,
                    precision="fp64", device=local, max_obs=p.n + 4, max_cells=p.m) as ctx:
    uid = [ctx.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(uid[0], rank, world)
    ctx.set_option("dist_fit", 1)
    ctx.set_observations(p.xy, p.y)
    linv_a = ctx.linv()
    status, msg = 0, ""
    try:
        ctx.set_observations(np.concatenate([p.xy, p.xy[300:302]]), np.concatenate([p.y, p.y[300:302]]))
    except pkg.SboError as e:
        status, msg = e.status, str(e)
    good = status == -4 and ctx.n_obs == p.n and np.array_equal(ctx.linv(), linv_a) and ctx.fit_info()[1] == world
    msgs = [None] * world
    dist.all_gather_object(msgs, (good, msg.split(" (the previous")[0]))
    same_msg = all(m[1] == msgs[0][1] for m in msgs)
    ok = ok and all(m[0] for m in msgs) and same_msg
    if rank == 0:
        print(f"not positive definite inside a shared fit: status {status} on every rank: {all(m[0] for m in msgs)}, same message: {same_msg} ({msgs[0][1]!r})", flush=True)
dist.destroy_process_group()
if rank == 0:
    print("dist_fit_check OK" if ok else "dist_fit_check FAILED", flush=True)
sys.exit(0 if ok else 1)
