"""Parity at the large configs (C3: 4096 observations, C4: 16 384) with the configs' real geometry: three bands of 8
grid rows of the full W x H grid (bottom edge, the clustered start region at H/4, mid-domain), GPU fp32 step (default
plan: block-sparse walk) against an fp64 scipy restatement evaluated on a random sample of the bands' cells.  Also a
rank-8 append + incremental step against a refit.  Prints M6-style lines."""
import sys, os, time
import numpy as np
from scipy.linalg import solve_triangular, cho_solve
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import safe_bayesian_optimization_b200 as pkg
from safe_bayesian_optimization_b200 import synthetic as S
from oracle import oracle as O

names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["C3", "C4"]
nsample = int(sys.argv[2]) if len(sys.argv) > 2 else 1536
rng = np.random.default_rng(7)
for name in names:
    p = S.make_problem(name)
    n = p.xy.shape[0]
    k = 8
    xy2, y2 = S.extra_observations(p, k, 0)
    xy_all, y_all = np.concatenate([p.xy, xy2]), np.concatenate([p.y, y2])
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    t0 = time.perf_counter()
    refs = []
    for xy, y in ((p.xy, p.y), (xy_all, y_all)):
        K = O.rbf_np(xy, xy, hp) + hp.noise_var * np.eye(xy.shape[0])
        Lm = np.linalg.cholesky(K); del K
        alpha = cho_solve((Lm, True), y - hp.prior_mean)
        refs.append((xy, Lm, alpha))
    print(f"{name}: n={n} grid {p.W}x{p.H}; fp64 host factorisations {time.perf_counter()-t0:.1f} s", flush=True)
    for band, row0 in (("edge", 0), ("start", p.H // 4 - 4), ("mid", p.H // 2)):
        cells_local = np.sort(rng.choice(8 * p.W, nsample, replace=False))
        cx, cy = g.cell_xy(row0 * p.W + cells_local)
        with pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                            precision="fp32", max_obs=n + k, max_cells=8 * p.W) as ctx:
            ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H, row0, row0 + 8)
            ctx.set_observations(p.xy, p.y)
            for stage, (xy, Lm, alpha) in zip(("full", "incremental(+8)"), refs):
                if stage == "full":
                    res, mask = ctx.step(p.beta, p.f_min)
                else:
                    ctx.append_observations(xy2, y2)
                    res, mask = ctx.step(p.beta, p.f_min, incremental=True)
                cg, skipping, kept = ctx.tc_plan()
                mu, s2 = ctx.mu()[cells_local], ctx.sigma2()[cells_local]
                Ks = O.rbf_np(np.stack([cx, cy], axis=1), xy, hp)
                mu_ref = hp.prior_mean + Ks @ alpha
                V = solve_triangular(Lm, Ks.T, lower=True)
                s2_ref = hp.sigma_f ** 2 - np.einsum("ij,ij->j", V, V)
                rel_mu = np.abs(mu - mu_ref) / np.maximum(np.abs(mu_ref), 1e-3 * np.abs(mu_ref).max())
                rel_s2 = np.abs(s2 - s2_ref) / np.maximum(np.abs(s2_ref), 1e-3 * np.abs(s2_ref).max())
                lcb = mu_ref - p.beta * np.sqrt(np.maximum(s2_ref, 0.0))
                mref = (lcb > p.f_min)
                mm = (mask[cells_local] != 0) != mref
                tol = 1e-5 * max(1.0, np.abs(mu_ref).max())
                inside = int(np.sum(mm & (np.abs(lcb - p.f_min) <= tol)))
                print(f"  {band:5s} rows {row0}-{row0+8} {stage:15s} kernel={res.kernel_used} skip={int(skipping)} kept={kept:.2f} "
                      f"mu rel max={rel_mu.max():.2e} mean={rel_mu.mean():.2e} | s2 rel max={rel_s2.max():.2e} abs max={np.abs(s2-s2_ref).max():.2e} "
                      f"| mask mismatches {int(mm.sum())} (inside band {inside}, outside {int(mm.sum())-inside}) of {nsample} | "
                      f"posterior {res.ms_posterior:.2f} ms", flush=True)
