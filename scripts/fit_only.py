"""Three fits at n (default 1024) -- target of the ncu launch list for the factorisation kernels."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
p = S.make_problem("C2", n=n, W=256, H=16)
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp32", max_obs=n, max_cells=p.m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    for i in range(3):
        t0 = time.perf_counter(); ctx.set_observations(p.xy, p.y); print(f"fit n={n}: {(time.perf_counter()-t0)*1e3:.3f} ms")
