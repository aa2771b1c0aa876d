"""Generates tests/golden/*.npz from the CPU oracle (C restatement, cross-checked against scipy and
scikit-learn in tests/test_oracle.py).  The reference ships no golden vectors and
cannot be built or imported here, so these fixtures are oracle outputs, not reference outputs.
Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from safe_bayesian_optimization_b200 import synthetic as S  # noqa: E402

CASES = {
    # name: (config, n, W, H)
    "c1_small": ("C1", 48, 40, 30),
    "c2_small": ("C2", 200, 96, 48),
    "c3_small": ("C3", 300, 128, 40),
}


def main():
    co = O.COracle()
    for name, (cfg, n, W, H) in CASES.items():
        p = S.make_problem(cfg, n=n, W=W, H=H)
        hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
        g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
        _, Linv, alpha = co.gp_fit(p.xy, p.y, hp)
        mu, s2 = co.posterior(p.xy, hp, Linv, alpha, g)
        std = np.sqrt(np.maximum(s2, 0.0))
        q0, q1 = co.confidence_intervals(mu, std, p.beta)
        mask, cnt = co.safe_set(q0, p.f_min)
        ucb_idx, ucb_val = co.argmax_safe(q1, mask)
        wid_idx, wid_val = co.argmax_safe(q1 - q0, mask)
        fr = O.find_safety_contour_indices(mask, g)
        sub = co.get_next_subgoal(fr, g, q0, q1, p.goal)
        np.savez_compressed(
            os.path.join(os.path.dirname(__file__), name + ".npz"),
            cfg=cfg, n=n, W=W, H=H, xy=p.xy, y=p.y, grid=np.array([p.x0, p.y0, p.rx, p.ry]),
            hyper=np.array([p.sigma_f, p.length, p.noise_var, p.prior_mean]), beta=p.beta, f_min=p.f_min,
            goal=np.array(p.goal), alpha=alpha, mu=mu, s2=s2, q0=q0, q1=q1, mask=mask, safe_count=cnt,
            ucb=np.array([ucb_idx, ucb_val]), width=np.array([wid_idx, wid_val]), frontier=fr, subgoal=sub)
        print(name, "n", n, "m", W * H, "safe", cnt, "frontier", fr.size, "subgoal", sub)


if __name__ == "__main__":
    main()
