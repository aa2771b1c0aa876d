"""Shared fit inside one process (sbo_group): members on distinct GPUs when the box has them, else all on device 0
(same schedule and exchange code, no speed-up to expect).  Checks L^-1 / alpha of every member against the replicated
fit bit for bit and prints the fit times.  usage: shared_fit_probe.py n[,n...] members"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S

ns = [int(x) for x in sys.argv[1].split(",")]
members = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ndev = torch.cuda.device_count()
devs = [i % ndev for i in range(members)]
ok = True
for n in ns:
    p = S.make_problem("C4" if n > 8192 else "C3", n=n, W=256, H=16)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision="fp32", max_obs=n, max_cells=p.m) as ctx:
        ctx.set_observations(p.xy, p.y)
        t0 = time.perf_counter(); ctx.set_observations(p.xy, p.y); w1 = (time.perf_counter() - t0) * 1e3
        ms1 = ctx.fit_info()[0]
        linv1, alpha1 = ctx.linv(), ctx.alpha()
    with pkg.SboGroup(devs, sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                      precision="fp32", max_obs=n, max_cells=p.m) as g:
        g.set_option("dist_fit", 1)
        g.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        g.set_observations(p.xy, p.y)
        t0 = time.perf_counter(); g.set_observations(p.xy, p.y); w2 = (time.perf_counter() - t0) * 1e3
        same, ms2 = True, 0.0
        for i in range(members):
            linv, alpha, (ms, shared) = g.member_factor(i)
            same = same and shared == members and np.array_equal(linv, linv1) and np.array_equal(alpha, alpha1)
            ms2 = max(ms2, ms)
            del linv
    ok = ok and same
    print(f"n={n:6d} members={members} on devices {devs}: replicated fit {ms1:8.2f} ms (host {w1:8.2f}) | shared fit {ms2:8.2f} ms "
          f"(host {w2:8.2f}) | every member bit-identical: {same}", flush=True)
print("shared_fit_probe OK" if ok else "shared_fit_probe FAILED")
sys.exit(0 if ok else 1)
