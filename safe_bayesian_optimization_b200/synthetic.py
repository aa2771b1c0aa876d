"""Seeded synthetic terrains, observation sets and grids for the named configs.

Stands in for the reference's missing ``data/terrain.csv``
(/root/reference/.MISSING_LARGE_BLOBS:1; loaded at
src/turtlesim_spatial_publisher.py:47-75) and mimics its observation
generator: 30 initial points within 1 m of the start pose (ibid. :111-149),
then further points (ibid. :151-183), values normalised to roughly 0-10
(ibid. :94-100).  GP hyper-parameters are those of config/lpsc.yaml:35-37,
beta of config/safe_bayesian_optimization.yaml:4.  Recipe: SURVEY.md 8(d) M2.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

LENGTH_SCALE = 0.4   # config/lpsc.yaml:36
SIGMA_F = 1.0        # config/lpsc.yaml:37
NOISE_VAR = 0.1      # config/lpsc.yaml:35
BETA = 2.0           # config/safe_bayesian_optimization.yaml:4

# name -> (n observations, W, H, precision, config number)
CONFIGS = {
    "C1": dict(n=64, W=200, H=200, precision="fp64", num=1, res=0.1),
    "C2": dict(n=1024, W=1024, H=1024, precision="fp32", num=2),
    "C3": dict(n=4096, W=4096, H=4096, precision="fp32", num=3, append=16),
    "C4": dict(n=16384, W=8192, H=8192, precision="fp32", num=4),
    "C5": dict(n=8192, W=4096, H=4096, precision="fp64", num=5, append=1, steps=50),
}


@dataclass
class Terrain:
    Lx: float
    Ly: float
    modes: np.ndarray = field(repr=False)  # (8, 4): kx, ky, phase, amplitude

    def __call__(self, x, y):
        x = np.asarray(x, dtype=np.float64); y = np.asarray(y, dtype=np.float64)
        f = 5.0 + 2.5 * np.sin(2 * np.pi * x / self.Lx * 3) * np.cos(2 * np.pi * y / self.Ly * 2)
        f = f + 1.5 * np.exp(-((x - 0.3 * self.Lx) ** 2 + (y - 0.6 * self.Ly) ** 2) / (0.02 * self.Lx * self.Ly))
        for kx, ky, ph, amp in self.modes:
            f = f + amp * np.sin(2 * np.pi * (kx * x / self.Lx + ky * y / self.Ly) + ph)
        return f


@dataclass
class Problem:
    name: str
    xy: np.ndarray          # (n, 2) fp64 observation coordinates
    y: np.ndarray           # (n,)  fp64 noisy observations
    x0: float
    y0: float
    rx: float
    ry: float
    W: int
    H: int
    beta: float
    f_min: float
    goal: tuple
    prior_mean: float
    precision: str
    terrain: Terrain = field(repr=False, default=None)
    sigma_f: float = SIGMA_F
    length: float = LENGTH_SCALE
    noise_var: float = NOISE_VAR

    @property
    def n(self) -> int:
        return self.xy.shape[0]

    @property
    def m(self) -> int:
        return self.W * self.H


def make_problem(name: str = "C2", n: int | None = None, W: int | None = None, H: int | None = None,
                 seed: int | None = None, precision: str | None = None) -> Problem:
    """Build the seeded synthetic problem for a named config; n / W / H override
    the config's sizes (used by the parity tests to shrink a config while keeping
    its recipe: observation density of about 8 per length-scale^2)."""
    cfg = dict(CONFIGS[name])
    n = cfg["n"] if n is None else n
    W = cfg["W"] if W is None else W
    H = cfg["H"] if H is None else H
    rng = np.random.default_rng((1234 + cfg["num"]) if seed is None else seed)
    if "res" in cfg:                       # C1: node default resolution 0.1 (reference :50-51)
        rx = ry = cfg["res"]
        Lx, Ly = W * rx, H * ry
    else:                                  # ~8 observations per length-scale^2
        Lx = Ly = LENGTH_SCALE * math.sqrt(max(n, 8) / 8.0)
        rx, ry = Lx / W, Ly / H
    modes = np.stack([rng.integers(1, 6, 8).astype(np.float64), rng.integers(1, 6, 8).astype(np.float64),
                      rng.uniform(0, 2 * np.pi, 8), np.full(8, 0.5 / math.sqrt(8.0))], axis=1)
    terrain = Terrain(Lx, Ly, modes)
    start = np.array([0.25 * Lx, 0.25 * Ly])
    n0 = min(30, n)
    rad = min(1.0, 0.2 * min(Lx, Ly)) * np.sqrt(rng.uniform(0, 1, n0))
    ang = rng.uniform(0, 2 * np.pi, n0)
    first = start + np.stack([rad * np.cos(ang), rad * np.sin(ang)], axis=1)
    rest = np.stack([rng.uniform(0, Lx, n - n0), rng.uniform(0, Ly, n - n0)], axis=1)
    xy = np.concatenate([first, rest], axis=0)
    y = terrain(xy[:, 0], xy[:, 1]) + rng.normal(0.0, math.sqrt(NOISE_VAR), n)
    # f_min: 40th percentile of the terrain on a coarse lattice (so ~half the domain is safe)
    gx, gy = np.meshgrid(np.linspace(0, Lx, 96), np.linspace(0, Ly, 96))
    f_min = float(np.percentile(terrain(gx, gy), 40.0))
    return Problem(name=name, xy=xy, y=y, x0=0.0, y0=0.0, rx=rx, ry=ry, W=W, H=H, beta=BETA, f_min=f_min,
                   goal=(0.9 * Lx, 0.9 * Ly), prior_mean=float(np.mean(y)) if n > 0 else 5.0,
                   precision=cfg["precision"] if precision is None else precision, terrain=terrain)


def extra_observations(p: Problem, k: int, step: int):
    """k further observations (for the rank-k append path), seeded by step."""
    rng = np.random.default_rng(977 + 31 * step + p.n)
    Lx, Ly = p.terrain.Lx, p.terrain.Ly
    xy = np.stack([rng.uniform(0, Lx, k), rng.uniform(0, Ly, k)], axis=1)
    y = p.terrain(xy[:, 0], xy[:, 1]) + rng.normal(0.0, math.sqrt(p.noise_var), k)
    return xy, y
