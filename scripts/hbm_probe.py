"""HBM-bound kernels at C4's grid size (8192 x 8192 = 64 Mi cells): sets (CI + mask + argmax), safety image gather,
safe / border compaction.  Times with CUDA events around the library calls' kernels are not exposed, so wall-clock of
the synchronous C-ABI calls (includes the D2H of the result) and, for the sets kernel, the library's own event timing."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
W = H = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
p = S.make_problem("C4", n=64, W=W, H=H)
m = W * H
with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                    precision="fp32", max_obs=64, max_cells=m) as ctx:
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, W, H)
    ctx.set_observations(p.xy, p.y)
    for _ in range(3):
        res = ctx.step_device(p.beta, p.f_min)
    print(f"sets kernel: {res.ms_sets:.3f} ms for {m} cells = {9.0*m/res.ms_sets/1e6:.0f} GB/s algorithmic (8 B read + 1 B written per cell); "
          f"posterior (n=64) {res.ms_posterior:.2f} ms; safe {res.safe_count}")
    for border in (False, True):
        for rep in range(3):
            t0 = time.perf_counter(); idx, cnt = ctx.compact_cells(border=border, cap=0); dt = time.perf_counter() - t0
        print(f"compact_cells(border={border}): count {cnt}, {dt*1e3:.3f} ms wall (count only, 2 passes over the mask = {2.0*m/dt/1e9:.0f} GB/s + index writes {4.0*cnt/1e6:.0f} MB)")
    for rep in range(2):
        t0 = time.perf_counter(); img, c, r = ctx.frontier_image(); dt = time.perf_counter() - t0
    print(f"frontier_image: {dt*1e3:.2f} ms wall incl. {m/1e6:.0f} MB D2H to pageable memory")
