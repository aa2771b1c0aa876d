"""Incremental (rank-k append) step timings through the thin kernel: fp32 at C2/C4 sizes, fp64 at a C5 slice."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S

def run(name, n, W, H, precision, ks):
    p = S.make_problem(name, n=n, W=W, H=H)
    with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                        precision=precision, max_obs=n + 64, max_cells=p.m) as ctx:
        ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        ctx.set_observations(p.xy, p.y)
        ctx.step(p.beta, p.f_min, want_mask=False)
        t = 0
        for k in ks:
            for rep in range(2):
                xy2, y2 = S.extra_observations(p, k, t); t += 1
                ctx.append_observations(xy2, y2)
                inc, _ = ctx.step(p.beta, p.f_min, incremental=True, want_mask=False)
            print(f"{precision} n={n} grid {W}x{H} k={k}: incremental posterior {inc.ms_posterior:8.3f} ms "
                  f"({p.m * float(n) / inc.ms_posterior / 1e6:.1f} G cell-obs/s) kernel={inc.kernel_used}", flush=True)

run("C3", 4096, 4096, 4096, "fp32", [1, 16])   # incremental GEMM kernel (>= one item per CTA pair)
run("C2", 1024, 1024, 1024, "fp32", [1, 2, 4, 8])
run("C4", 16384, 8192, 512, "fp32", [1])
run("C4", 16384, 8192, 2048, "fp32", [1, 4, 16])
run("C5", 8192, 4096, 256, "fp64", [1, 2, 4, 8])
run("C5", 8192, 4096, 1024, "fp64", [1])   # >= one 128 x 128 tile per SM: the 2-D tile kernel
os.environ["SBO_THIN_2D"] = "0"
run("C5", 8192, 4096, 1024, "fp64", [1])
run("C4", 16384, 8192, 512, "fp32", [1])
run("C4", 16384, 8192, 2048, "fp32", [1, 4, 16])
