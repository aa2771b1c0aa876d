"""CPU oracle for the safe-BO hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module, and only as the checker or the timed
CPU baseline.  The product (safe_bayesian_optimization_b200/) never does.

Three independent statements of the same arithmetic live here:

* ``COracle``  -- ctypes binding of libsbo_oracle.so (oracle/sbo_oracle.c,
  oracle/sbo_oracle_pinned.c), plain C fp64;
* ``gp_fit_scipy`` / ``gp_posterior_scipy`` -- scipy/OpenBLAS ``cho_factor`` +
  ``solve_triangular`` + GEMM (also the multi-threaded CPU baseline);
* ``gp_posterior_sklearn`` -- scikit-learn ``GaussianProcessRegressor``.

Parity status: confidence intervals / safe set / subgoal selection are pinned
by /root/reference/src/safe_bayesian_optimization_node.cpp:409-416, :499-550;
the GP posterior is PARITY UNPINNED (the reference obtains it from the
external ``mapping_package`` node, ibid. :568-644) -- see sbo_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libsbo_oracle.so")


def build(force: bool = False) -> str:
    """Compile libsbo_oracle.so with oracle/Makefile (gcc); returns its path."""
    srcs = [os.path.join(_HERE, f) for f in ("sbo_oracle.c", "sbo_oracle_pinned.c", "sbo_oracle.h", "Makefile")]
    stale = force or not os.path.exists(_LIB) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _LIB


class _Hyper(C.Structure):
    _fields_ = [("sigma_f", C.c_double), ("length", C.c_double),
                ("noise_var", C.c_double), ("prior_mean", C.c_double)]


class _Grid(C.Structure):
    _fields_ = [("x0", C.c_double), ("y0", C.c_double), ("rx", C.c_double), ("ry", C.c_double),
                ("W", C.c_int32), ("H", C.c_int32)]


@dataclass
class Hyper:
    sigma_f: float = 1.0      # config/lpsc.yaml:37
    length: float = 0.4       # config/lpsc.yaml:36
    noise_var: float = 0.1    # config/lpsc.yaml:35
    prior_mean: float = 0.0

    def c(self) -> _Hyper:
        return _Hyper(self.sigma_f, self.length, self.noise_var, self.prior_mean)


@dataclass
class Grid:
    x0: float
    y0: float
    rx: float
    ry: float
    W: int
    H: int

    @property
    def m(self) -> int:
        return self.W * self.H

    def c(self) -> _Grid:
        return _Grid(self.x0, self.y0, self.rx, self.ry, self.W, self.H)

    def cell_xy(self, cells=None):
        """fp64 coordinates, product and sum rounded separately (no FMA)."""
        if cells is None:
            cells = np.arange(self.m, dtype=np.int64)
        cells = np.asarray(cells, dtype=np.int64)
        iy = cells // self.W
        ix = cells - iy * self.W
        return self.x0 + ix.astype(np.float64) * self.rx, self.y0 + iy.astype(np.float64) * self.ry


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class COracle:
    """ctypes face of libsbo_oracle.so."""

    def __init__(self):
        self.lib = C.CDLL(build())
        L = self.lib
        d, i32, i64 = C.POINTER(C.c_double), C.c_int32, C.c_int64
        L.oracle_gp_fit.restype = i32
        L.oracle_gp_fit.argtypes = [d, d, i32, C.POINTER(_Hyper), d, d, d]
        L.oracle_posterior.restype = None
        L.oracle_posterior.argtypes = [d, i32, C.POINTER(_Hyper), d, d, C.POINTER(_Grid), i64, i64, d, d]
        L.oracle_confidence_intervals.restype = None
        L.oracle_confidence_intervals.argtypes = [d, d, C.c_double, i64, d, d]
        L.oracle_safe_set.restype = i64
        L.oracle_safe_set.argtypes = [d, C.c_double, i64, C.POINTER(C.c_uint8)]
        L.oracle_get_next_subgoal.restype = i32
        L.oracle_get_next_subgoal.argtypes = [C.POINTER(i32), i64, C.POINTER(_Grid), d, d, C.c_double, C.c_double]
        L.oracle_argmax_safe.restype = i64
        L.oracle_argmax_safe.argtypes = [d, C.POINTER(C.c_uint8), i64, d]
        L.oracle_num_threads.restype = i32

    def num_threads(self) -> int:
        return int(self.lib.oracle_num_threads())

    def gp_fit(self, xy, y, hp: Hyper):
        xy = np.ascontiguousarray(xy, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = xy.shape[0]
        Lm = np.empty((n, n)); Linv = np.empty((n, n)); alpha = np.empty(n)
        h = hp.c()
        st = self.lib.oracle_gp_fit(_dp(xy), _dp(y), n, C.byref(h), _dp(Lm), _dp(Linv), _dp(alpha))
        if st != 0:
            raise np.linalg.LinAlgError(f"oracle cholesky: non-positive pivot at row {st - 1}")
        return Lm, Linv, alpha

    def posterior(self, xy, hp: Hyper, Linv, alpha, grid: Grid, cell0=0, cell1=None):
        xy = np.ascontiguousarray(xy, dtype=np.float64)
        Linv = np.ascontiguousarray(Linv, dtype=np.float64)
        alpha = np.ascontiguousarray(alpha, dtype=np.float64)
        if cell1 is None:
            cell1 = grid.m
        mu = np.empty(cell1 - cell0); s2 = np.empty(cell1 - cell0)
        h, g = hp.c(), grid.c()
        self.lib.oracle_posterior(_dp(xy), xy.shape[0], C.byref(h), _dp(Linv), _dp(alpha), C.byref(g),
                                  cell0, cell1, _dp(mu), _dp(s2))
        return mu, s2

    def confidence_intervals(self, mu, std, beta):
        mu = np.ascontiguousarray(mu, dtype=np.float64); std = np.ascontiguousarray(std, dtype=np.float64)
        q0 = np.empty_like(mu); q1 = np.empty_like(mu)
        self.lib.oracle_confidence_intervals(_dp(mu), _dp(std), float(beta), mu.size, _dp(q0), _dp(q1))
        return q0, q1

    def safe_set(self, q0, f_min):
        q0 = np.ascontiguousarray(q0, dtype=np.float64)
        mask = np.empty(q0.size, dtype=np.uint8)
        cnt = self.lib.oracle_safe_set(_dp(q0), float(f_min), q0.size, mask.ctypes.data_as(C.POINTER(C.c_uint8)))
        return mask, int(cnt)

    def get_next_subgoal(self, frontier, grid: Grid, q0, q1, goal_xy):
        fr = np.ascontiguousarray(frontier, dtype=np.int32)
        q0 = np.ascontiguousarray(q0, dtype=np.float64); q1 = np.ascontiguousarray(q1, dtype=np.float64)
        g = grid.c()
        return int(self.lib.oracle_get_next_subgoal(fr.ctypes.data_as(C.POINTER(C.c_int32)), fr.size, C.byref(g),
                                                    _dp(q0), _dp(q1), float(goal_xy[0]), float(goal_xy[1])))

    def argmax_safe(self, val, mask):
        val = np.ascontiguousarray(val, dtype=np.float64); mask = np.ascontiguousarray(mask, dtype=np.uint8)
        best = C.c_double(0.0)
        i = self.lib.oracle_argmax_safe(_dp(val), mask.ctypes.data_as(C.POINTER(C.c_uint8)), val.size, C.byref(best))
        return int(i), float(best.value)


# ---- numpy restatements of the pinned arithmetic (second, independent statement) ----

def confidence_intervals_np(mu, std, beta):
    """reference :411-416"""
    conf = np.float64(beta) * np.asarray(std, dtype=np.float64)
    return mu - conf, mu + conf


def safe_set_np(q0, f_min):
    """reference :409 (strict >), bytes as at :454"""
    s = np.asarray(q0) > f_min
    return np.where(s, 255, 0).astype(np.uint8), int(s.sum())


def get_next_subgoal_np(frontier, grid: Grid, q0, q1, goal_xy):
    """reference :499-550, pure numpy (lexsort on (dist, position))."""
    fr = np.asarray(frontier, dtype=np.int64)
    if fr.size == 0:
        return -1
    x, y = grid.cell_xy(fr)
    width = q1[fr] - q0[fr]
    dx = x - goal_xy[0]; dy = y - goal_xy[1]
    dist = np.sqrt(dx * dx + dy * dy)
    order = np.lexsort((np.arange(fr.size), dist))
    top = max(1, fr.size // 4)
    best, best_w = -1, -1.0
    for p in order[:top]:
        if width[p] > best_w:
            best_w, best = width[p], p
    return int(fr[best]) if best >= 0 else -1


# ---- scipy / sklearn GP (independent cross-checks; also the BLAS CPU baseline) ----

def rbf_np(A, B, hp: Hyper):
    d2 = ((A[:, None, 0] - B[None, :, 0]) ** 2 + (A[:, None, 1] - B[None, :, 1]) ** 2)
    return hp.sigma_f ** 2 * np.exp(-d2 / (2.0 * hp.length ** 2))


def gp_fit_scipy(xy, y, hp: Hyper):
    from scipy.linalg import cho_factor, cho_solve, solve_triangular
    xy = np.asarray(xy, dtype=np.float64)
    n = xy.shape[0]
    K = rbf_np(xy, xy, hp) + hp.noise_var * np.eye(n)
    Lm = np.linalg.cholesky(K)
    Linv = solve_triangular(Lm, np.eye(n), lower=True)
    alpha = cho_solve((Lm, True), np.asarray(y, dtype=np.float64) - hp.prior_mean)
    return Lm, Linv, alpha


def gp_posterior_scipy(xy, hp: Hyper, Linv, alpha, grid: Grid, cell0=0, cell1=None, block=8192):
    """mu, sigma^2 by blocked GEMM (OpenBLAS, all threads)."""
    if cell1 is None:
        cell1 = grid.m
    xy = np.asarray(xy, dtype=np.float64)
    mu = np.empty(cell1 - cell0); s2 = np.empty(cell1 - cell0)
    LinvT = np.ascontiguousarray(Linv.T)
    for b0 in range(cell0, cell1, block):
        b1 = min(b0 + block, cell1)
        cx, cy = grid.cell_xy(np.arange(b0, b1))
        Ks = rbf_np(np.stack([cx, cy], axis=1), xy, hp)
        mu[b0 - cell0:b1 - cell0] = hp.prior_mean + Ks @ alpha
        V = Ks @ LinvT
        s2[b0 - cell0:b1 - cell0] = hp.sigma_f ** 2 - np.einsum("ij,ij->i", V, V)
    return mu, s2


def gp_posterior_sklearn(xy, y, hp: Hyper, grid: Grid, cells):
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel
    k = ConstantKernel(hp.sigma_f ** 2, "fixed") * RBF(hp.length, "fixed")
    gpr = GaussianProcessRegressor(kernel=k, alpha=hp.noise_var, optimizer=None, normalize_y=False)
    gpr.fit(np.asarray(xy, dtype=np.float64), np.asarray(y, dtype=np.float64) - hp.prior_mean)
    cx, cy = grid.cell_xy(cells)
    mu, std = gpr.predict(np.stack([cx, cy], axis=1), return_std=True)
    return mu + hp.prior_mean, std ** 2


# ---- frontier (reference :418-497), host stage that consumes the mask ("next" row N1) ----

def safety_image(mask, grid: Grid):
    """The image and the pixel -> cell table of FindSafetyContourIndices, cell by cell as the reference does it:
    int-truncated bounds (:431-434), pixel mapping (:450-451), drop rule (:453), last-writer-wins (:454, :468-475).
    Returns (img uint8 H x W, lut int64 W x H with -1 where no cell landed)."""
    m = grid.m
    x, y = grid.cell_xy()
    min_x, max_x = int(x.min()), int(x.max())   # static_cast<int>: truncation toward zero
    min_y, max_y = int(y.min()), int(y.max())
    if max_x == min_x or max_y == min_y:
        raise ZeroDivisionError("reference divides by (max-min)==0 here (:450-451)")
    with np.errstate(invalid="ignore"):
        px = ((x - min_x) / (max_x - min_x) * grid.W).astype(np.int64)   # truncation, values >= 0
        py = ((y - min_y) / (max_y - min_y) * grid.H).astype(np.int64)
    ok = (px >= 0) & (px < grid.W) & (py >= 0) & (py < grid.H)
    img = np.zeros((grid.H, grid.W), dtype=np.uint8)
    lut = -np.ones((grid.W, grid.H), dtype=np.int64)
    idx = np.arange(m)[ok]
    # sequential "last writer wins": later cells overwrite earlier ones
    img[py[ok], px[ok]] = np.where(np.asarray(mask)[ok] != 0, 255, 0)
    lut[px[ok], py[ok]] = idx
    return img, lut


def find_safety_contour_indices(mask, grid: Grid):
    """FindSafetyContourIndices restated with cv2.findContours (same Suzuki-Abe
    border following as the C++ call at :461-462) on the image of safety_image()."""
    import cv2
    img, lut = safety_image(mask, grid)
    contours, _ = cv2.findContours(img, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_NONE)
    out = []
    for c in contours:
        pts = c.reshape(-1, 2)
        li = lut[pts[:, 0], pts[:, 1]]
        out.extend(li[li >= 0].tolist())
    return np.asarray(out, dtype=np.int32)


def compact_cells(mask, grid: Grid, border=False):
    """Ascending indices of the safe cells (reference :944-949 gathers exactly these for the alpha shape); border=True
    keeps only safe cells with an unsafe cell or the grid frame among their 8 neighbours."""
    s = (np.asarray(mask).reshape(grid.H, grid.W) != 0)
    if border:
        p = np.zeros((grid.H + 2, grid.W + 2), dtype=bool)
        p[1:-1, 1:-1] = s
        full = np.ones_like(s)
        for dy in (0, 1, 2):
            for dx in (0, 1, 2):
                full &= p[dy:dy + grid.H, dx:dx + grid.W]
        s = s & ~full
    return np.flatnonzero(s.reshape(-1)).astype(np.int32)


# ---- N3: goal-in-polygon and polydist (host arithmetic of the reference, restated independently) ----

def point_within_polygon(rings, p):
    """bg::within(point, polygon) (reference src/safe_bayesian_optimization_node.cpp:657): strictly inside the outer
    ring and strictly outside every hole.  Even-odd ray casting plus an explicit on-boundary test -- a different
    algorithm from the library's winding count, on purpose."""
    def on_segment(a, b, q):
        cross = (b[0] - a[0]) * (q[1] - a[1]) - (b[1] - a[1]) * (q[0] - a[0])
        scale = max(1.0, abs(b[0] - a[0]), abs(b[1] - a[1]), abs(q[0] - a[0]), abs(q[1] - a[1]))
        if abs(cross) > np.finfo(np.float64).eps * scale:
            return False
        return min(a[0], b[0]) <= q[0] <= max(a[0], b[0]) and min(a[1], b[1]) <= q[1] <= max(a[1], b[1])

    def ring_state(r, q):          # -1 outside, 0 boundary, 1 inside
        r = [tuple(map(float, v)) for v in r]
        if len(r) < 3:
            return -1
        if r[0] != r[-1]:
            r = r + [r[0]]
        inside = False
        for a, b in zip(r[:-1], r[1:]):
            if on_segment(a, b, q):
                return 0
            if (a[1] > q[1]) != (b[1] > q[1]):
                xi = a[0] + (q[1] - a[1]) * (b[0] - a[0]) / (b[1] - a[1])
                if xi > q[0]:
                    inside = not inside
        return 1 if inside else -1

    q = (float(p[0]), float(p[1]))
    if len(rings) == 0 or ring_state(rings[0], q) != 1:
        return False
    return all(ring_state(h, q) == -1 for h in rings[1:])


def polydist(ring, p):
    """polydist (reference src/libraries/polygeom_lib.cpp:400-474) after the node's bg::correct on a closed,
    counter-clockwise polygon type (node :677-683, include/polygeom_lib.h:60).  Line-by-line restatement, including
    VertexListRolled[j] = VertexList[j] (:437), which makes the result the nearest vertex.  Returns ((x, y), dist)."""
    import math
    V = [[float(v[0]), float(v[1])] for v in ring]
    if len(V) == 0:
        return (0.0, 0.0), 100000000.0                                      # :419-423
    if len(V) > 2 and V[0] != V[-1]:                                        # bg::correct: close ...
        V.append(list(V[0]))
    area2 = sum(V[i][0] * V[i + 1][1] - V[i + 1][0] * V[i][1] for i in range(len(V) - 1))
    if area2 < 0.0:                                                         # ... and make counter-clockwise
        V.reverse()
    V.pop()                                                                 # :427
    n = len(V)
    if n == 0:
        return (0.0, 0.0), 100000000.0
    rolled = [[0.0, 0.0] for _ in range(n)]; dxy = [[0.0, 0.0] for _ in range(n)]; norm = [0.0] * n
    for i in range(n):                                                      # :433-443
        j = (i + 1) % n
        rolled[j] = [V[j][0], V[j][1]]
        dxy[j] = [V[j][0] - V[i][0], V[j][1] - V[i][1]]
        norm[j] = math.sqrt(dxy[j][0] * dxy[j][0] + dxy[j][1] * dxy[j][1])
        if norm[j] == 0.0:
            norm[j] = 1.0
    dtemp = []; ctemp = []
    for i in range(n):                                                      # :449-459
        n2 = norm[i] ** 2
        w_temp = (p[0] - V[i][0]) * (dxy[i][0] / n2) + (p[1] - V[i][1]) * (dxy[i][1] / n2)
        w = max(min(w_temp, 1.0), 0.0)
        c = [(1 - w) * V[i][0] + w * rolled[i][0], (1 - w) * V[i][1] + w * rolled[i][1]]
        ctemp.append(c)
        dtemp.append(math.sqrt((p[0] - c[0]) ** 2 + (p[1] - c[1]) ** 2))
    k = min(range(n), key=lambda i: (dtemp[i], i))                          # std::min_element: first minimum
    return (ctemp[k][0], ctemp[k][1]), dtemp[k]
