"""Host side of the safe-BO step above the C-ABI.

``SboContext`` is a thin RAII wrapper of ``sbo_ctx``.  ``OptimizerCore`` mirrors
the step interface of the reference's ``OptimizerNode``
(/root/reference/src/safe_bayesian_optimization_node.cpp): same member names
(``mu_``, ``std_``, ``S_``, ``Q_``, ``beta_``, ``f_min_``), same method names
(``process_terrain_map``, ``ComputeSets``, ``FindSafetyContourIndices``,
``GetNextSubgoal``) and the same "warn and return -1 / empty" error behaviour
(:419-422, :503-506), so the parity tests read like tests of the node.  All
arithmetic of the path runs in libsbo.so on the GPU; only the contour trace
(OpenCV, reference :461-462) stays on the host, as in the reference.
"""
from __future__ import annotations

import ctypes as C
import logging
from dataclasses import dataclass

import numpy as np

from . import _lib as L

log = logging.getLogger("safe_bayesian_optimization_b200")


@dataclass
class StepResult:
    n_cells: int
    safe_count: int
    argmax_ucb: tuple      # (cell, value); cell == -1 when the safe set is empty
    argmax_width: tuple
    clamped_sigma2: int
    n_obs: int
    kernel_used: int
    ms_posterior: float
    ms_sets: float
    ms_comm: float
    ms_total: float
    low_sigma2: int = 0      # cells with sigma^2 < 0.02 sigma_f^2
    refreshed: int = 0       # an incremental step that the refresh policy turned into a full posterior
    drift: float = -1.0      # result of this step's drift check (-1: none)

    @staticmethod
    def from_c(o: L.SboStepOut) -> "StepResult":
        return StepResult(o.n_cells, o.safe_count, (o.argmax_ucb.idx, o.argmax_ucb.value),
                          (o.argmax_width.idx, o.argmax_width.value), o.clamped_sigma2, o.n_obs, o.kernel_used,
                          o.ms_posterior, o.ms_sets, o.ms_comm, o.ms_total, o.low_sigma2, o.refreshed, o.drift)


def _dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _u8ptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _pack_rings(rings):
    rings = [np.ascontiguousarray(r, dtype=np.float64).reshape(-1, 2) for r in rings]
    sizes = np.array([r.shape[0] for r in rings], dtype=np.int32)
    xy = np.ascontiguousarray(np.concatenate(rings, axis=0)) if rings else np.zeros((0, 2))
    return xy, sizes


def point_within_polygon(rings, p) -> bool:
    """bg::within(point, polygon) of reference :657 (host arithmetic in libsbo.so, no GPU involved)."""
    lib = L.load()
    xy, sizes = _pack_rings(rings)
    w = C.c_int32(0)
    st = lib.sbo_point_within_polygon(_dptr(xy), sizes.ctypes.data_as(C.POINTER(C.c_int32)), sizes.size,
                                      float(p[0]), float(p[1]), C.byref(w))
    if st != L.SBO_OK:
        raise ValueError(f"sbo_point_within_polygon: status {st}")
    return bool(w.value)


def polydist(ring, p):
    """polydist of reference polygeom_lib.cpp:400-474 as the node calls it: ((x, y), dist)."""
    lib = L.load()
    ring = np.ascontiguousarray(ring, dtype=np.float64).reshape(-1, 2)
    out, d = np.zeros(2), C.c_double(0.0)
    st = lib.sbo_polydist(_dptr(ring), ring.shape[0], float(p[0]), float(p[1]), _dptr(out), C.byref(d))
    if st != L.SBO_OK:
        raise ValueError(f"sbo_polydist: status {st}")
    return (float(out[0]), float(out[1])), float(d.value)


class SboContext:
    """Owns one ``sbo_ctx`` (one GPU, one shard of grid rows)."""

    def __init__(self, *, sigma_f=1.0, length_scale=0.4, noise_var=0.1, prior_mean=0.0, precision="fp32",
                 kernel="auto", device=0, max_obs=1024, max_cells=1 << 20, include_noise_in_sigma=False):
        self.lib = L.load()
        cfg = L.SboConfig(sigma_f, length_scale, noise_var, prior_mean, int(include_noise_in_sigma),
                          {"fp32": L.SBO_FP32, "fp64": L.SBO_FP64}[precision],
                          {"auto": L.SBO_KERNEL_AUTO, "simt": L.SBO_KERNEL_SIMT, "tcgen05": L.SBO_KERNEL_TCGEN05}[kernel],
                          device, max_obs, 0, max_cells)
        self.precision = precision
        self._h = C.c_void_p()
        st = self.lib.sbo_create(C.byref(cfg), C.byref(self._h))
        if st != L.SBO_OK:
            msg = self.lib.sbo_last_error(None).decode()
            self._h = None
            raise L.SboError(st, msg)
        self.m_local = 0
        self.cell0 = 0

    # -- plumbing ---------------------------------------------------------
    def _check(self, st):
        if st != L.SBO_OK:
            raise L.SboError(st, self.lib.sbo_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            self.lib.sbo_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def handle(self):
        return self._h

    def set_stream(self, cuda_stream_ptr: int | None):
        self._check(self.lib.sbo_set_stream(self._h, C.c_void_p(cuda_stream_ptr or 0)))

    def synchronize(self):
        self._check(self.lib.sbo_synchronize(self._h))

    def set_option(self, name: str, value: int):
        """sbo_set_option by name (``_lib.OPTIONS``): the A/B and policy switches; nothing is read from the environment."""
        self._check(self.lib.sbo_set_option(self._h, L.OPTIONS[name], int(value)))

    def get_option(self, name: str) -> int:
        v = C.c_int32(0)
        self._check(self.lib.sbo_get_option(self._h, L.OPTIONS[name], C.byref(v)))
        return int(v.value)

    def device_info(self):
        sm, ok = C.c_int32(), C.c_int32()
        self._check(self.lib.sbo_device_info(self._h, C.byref(sm), C.byref(ok)))
        return sm.value, bool(ok.value)

    @property
    def launch_count(self) -> int:
        return int(self.lib.sbo_launch_count(self._h))

    # -- grid / observations ----------------------------------------------
    def set_grid(self, x0, y0, rx, ry, W, H, row0=0, row1=None):
        row1 = H if row1 is None else row1
        self._check(self.lib.sbo_set_grid(self._h, x0, y0, rx, ry, W, H, row0, row1))
        self.W, self.H, self.row0, self.row1 = W, H, row0, row1
        self.m_local = (row1 - row0) * W
        self.cell0 = row0 * W

    def set_observations(self, xy, y):
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1, 2)
        y = np.ascontiguousarray(y, dtype=np.float64).reshape(-1)
        if xy.shape[0] != y.shape[0]:
            raise ValueError("xy and y disagree on n")
        self._check(self.lib.sbo_set_observations(self._h, _dptr(xy), _dptr(y), xy.shape[0]))

    def append_observations(self, xy, y):
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1, 2)
        y = np.ascontiguousarray(y, dtype=np.float64).reshape(-1)
        self._check(self.lib.sbo_append_observations(self._h, _dptr(xy), _dptr(y), xy.shape[0]))

    @property
    def n_obs(self) -> int:
        return int(self.lib.sbo_num_observations(self._h))

    def fit_info(self):
        """(device ms of the last full fit, ranks that shared it; 0 = this context alone) -- sbo_fit_info."""
        ms, w = C.c_float(0.0), C.c_int32(0)
        self._check(self.lib.sbo_fit_info(self._h, C.byref(ms), C.byref(w)))
        return float(ms.value), int(w.value)

    def mask_copy_info(self):
        """(mask bytes the last step copied to the host, contiguous runs) -- sbo_mask_copy_info (option ``mask_delta``)."""
        b, r = C.c_int64(0), C.c_int32(0)
        self._check(self.lib.sbo_mask_copy_info(self._h, C.byref(b), C.byref(r)))
        return int(b.value), int(r.value)

    def order(self):
        """order[pos] = caller index of the observation stored at internal position pos."""
        out = np.zeros(self.n_obs, dtype=np.int32)
        self._check(self.lib.sbo_copy_order(self._h, out.ctypes.data_as(C.POINTER(C.c_int32))))
        return out

    def linv(self):
        n = self.n_obs
        out = np.zeros((n, n))
        self._check(self.lib.sbo_copy_linv(self._h, _dptr(out)))
        return out

    def alpha(self):
        out = np.zeros(self.n_obs)
        self._check(self.lib.sbo_copy_alpha(self._h, _dptr(out)))
        return out

    # -- the step -----------------------------------------------------------
    def step(self, beta, f_min, mask_out=None, incremental=False, want_mask=True):
        """One BO step; returns (StepResult, mask uint8[m_local] or None)."""
        if mask_out is None and want_mask:
            mask_out = np.empty(self.m_local, dtype=np.uint8)
        out = L.SboStepOut()
        fn = self.lib.sbo_step_incremental if incremental else self.lib.sbo_step
        self._check(fn(self._h, float(beta), float(f_min), _u8ptr(mask_out) if mask_out is not None else None,
                       C.byref(out)))
        return StepResult.from_c(out), mask_out

    def step_device(self, beta, f_min):
        out = L.SboStepOut()
        self._check(self.lib.sbo_step_device(self._h, float(beta), float(f_min), C.byref(out)))
        return StepResult.from_c(out)

    def compute_sets(self, mu, std, beta, f_min, want_q=True):
        mu = np.ascontiguousarray(mu, dtype=np.float64).reshape(-1)
        std = np.ascontiguousarray(std, dtype=np.float64).reshape(-1)
        if mu.size != std.size:
            raise ValueError("mu and std disagree on m")
        m = mu.size
        mask = np.empty(m, dtype=np.uint8)
        q0 = np.empty(m) if want_q else None
        q1 = np.empty(m) if want_q else None
        out = L.SboStepOut()
        self._check(self.lib.sbo_compute_sets(self._h, _dptr(mu), _dptr(std), m, float(beta), float(f_min), _u8ptr(mask),
                                              _dptr(q0) if want_q else None, _dptr(q1) if want_q else None, C.byref(out)))
        return StepResult.from_c(out), mask, q0, q1

    def select_subgoal(self, frontier, goal_xy) -> int:
        fr = np.ascontiguousarray(frontier, dtype=np.int32).reshape(-1)
        cell = C.c_int32(-1)
        self._check(self.lib.sbo_select_subgoal(self._h, fr.ctypes.data_as(C.POINTER(C.c_int32)), fr.size,
                                                float(goal_xy[0]), float(goal_xy[1]), C.byref(cell)))
        return int(cell.value)

    def subgoal_point(self, frontier, goal_xy, rings):
        """Reference :650-706: (kind, (x, y), cell, dist).  rings = [outer, hole, ...] as (k, 2) arrays."""
        fr = np.ascontiguousarray(frontier, dtype=np.int32).reshape(-1)
        xy, sizes = _pack_rings(rings)
        out = np.zeros(2)
        cell, kind, dist = C.c_int32(-1), C.c_int32(0), C.c_double(0.0)
        self._check(self.lib.sbo_subgoal_point(self._h, fr.ctypes.data_as(C.POINTER(C.c_int32)), fr.size,
                                               float(goal_xy[0]), float(goal_xy[1]), _dptr(xy),
                                               sizes.ctypes.data_as(C.POINTER(C.c_int32)), sizes.size, _dptr(out),
                                               C.byref(cell), C.byref(kind), C.byref(dist)))
        return kind.value, (float(out[0]), float(out[1])), int(cell.value), float(dist.value)

    def ingest(self, xy, values, uncertainty=None, stamps=None):
        """Feed SpatialMeasurement records (position x/y, value); returns (spatial_data_size, refitted)."""
        xy = np.asarray(xy, dtype=np.float64).reshape(-1, 2)
        values = np.asarray(values, dtype=np.float64).reshape(-1)
        recs = (L.SboMeasurement * len(values))()
        for i in range(len(values)):
            recs[i].x, recs[i].y, recs[i].z, recs[i].value = xy[i, 0], xy[i, 1], 0.0, values[i]
            recs[i].uncertainty = 0.0 if uncertainty is None else float(uncertainty[i])
            if stamps is not None:
                recs[i].stamp_sec, recs[i].stamp_nanosec = int(stamps[i][0]), int(stamps[i][1])
        size, refit = C.c_int32(0), C.c_int32(0)
        self._check(self.lib.sbo_ingest_measurements(self._h, recs, len(values), C.byref(size), C.byref(refit)))
        return int(size.value), bool(refit.value)

    # -- read-back ----------------------------------------------------------
    def tc_plan(self):
        """(cta_group, skipping, kept_fraction) of the tcgen05 posterior kernel on the current grid + observations."""
        cg, sk, kf = C.c_int32(0), C.c_int32(0), C.c_double(1.0)
        self._check(self.lib.sbo_tc_plan(self._h, C.byref(cg), C.byref(sk), C.byref(kf)))
        return cg.value, bool(sk.value), kf.value

    def tc_plan_detail(self):
        """dict: bits, ring_bits, kept_fraction, pass_fraction, linv_row_l1, strip_tiles (sbo_tc_plan_detail)."""
        out = np.zeros(6)
        self._check(self.lib.sbo_tc_plan_detail(self._h, _dptr(out)))
        return dict(zip(("bits", "ring_bits", "kept_fraction", "pass_fraction", "linv_row_l1", "strip_tiles"), out.tolist()))

    def tc_executed_flops(self) -> float:
        """Tensor flops the last full tcgen05 refresh executed (sbo_tc_executed_flops)."""
        v = C.c_double(0.0)
        self._check(self.lib.sbo_tc_executed_flops(self._h, C.byref(v)))
        return float(v.value)

    def frontier_image(self):
        """(img H x W uint8 0/255, col_of_px int32[W], row_of_py int32[H]): the image of reference :444-456 built on the
        GPU with the reference's pixel map, and the pixel -> cell tables that replace its hash map (:468-475)."""
        img = np.empty((self.H, self.W), dtype=np.uint8)
        col = np.empty(self.W, dtype=np.int32)
        row = np.empty(self.H, dtype=np.int32)
        i32 = C.POINTER(C.c_int32)
        self._check(self.lib.sbo_frontier_image(self._h, _u8ptr(img), col.ctypes.data_as(i32), row.ctypes.data_as(i32)))
        return img, col, row

    def compact_cells(self, border=False, cap=None):
        """Ascending global indices of the safe cells (border=True: only those with an unsafe cell or the grid frame
        in their 8-neighbourhood) of this shard, compacted on the GPU.  Returns (indices int32, total count)."""
        cap = self.m_local if cap is None else int(cap)
        idx = np.empty(max(cap, 1), dtype=np.int32)
        cnt = C.c_int64(0)
        self._check(self.lib.sbo_compact_cells(self._h, 1 if border else 0, idx.ctypes.data_as(C.POINTER(C.c_int32)), cap,
                                               C.byref(cnt)))
        return idx[:min(cap, cnt.value)], cnt.value

    def mu(self):
        out = np.empty(self.m_local)
        self._check(self.lib.sbo_copy_mu_f64(self._h, _dptr(out)))
        return out

    def sigma2(self):
        out = np.empty(self.m_local)
        self._check(self.lib.sbo_copy_sigma2_f64(self._h, _dptr(out)))
        return out

    def mask(self, m=None):
        out = np.empty(self.m_local if m is None else m, dtype=np.uint8)
        self._check(self.lib.sbo_copy_mask(self._h, _u8ptr(out)))
        return out

    # -- multi-GPU ------------------------------------------------------------
    def comm_unique_id(self) -> bytes:
        buf = (C.c_uint8 * 128)()
        st = self.lib.sbo_comm_unique_id(buf)
        if st != L.SBO_OK:
            raise L.SboError(st, "sbo_comm_unique_id failed (libnccl not loadable?)")
        return bytes(buf)

    def comm_init(self, uid: bytes, rank: int, world: int):
        buf = (C.c_uint8 * 128).from_buffer_copy(uid)
        self._check(self.lib.sbo_comm_init(self._h, buf, rank, world))


class SboGroup:
    """One object over several GPUs of the box, driven from one host thread (``sbo_group``): cost-balanced row bands, one
    H x W mask image, merged argmax records.  ``devices`` may repeat an ordinal (several bands on one GPU)."""

    def __init__(self, devices, *, sigma_f=1.0, length_scale=0.4, noise_var=0.1, prior_mean=0.0, precision="fp32",
                 kernel="auto", max_obs=1024, max_cells=1 << 20, include_noise_in_sigma=False):
        self.lib = L.load()
        cfg = L.SboConfig(sigma_f, length_scale, noise_var, prior_mean, int(include_noise_in_sigma),
                          {"fp32": L.SBO_FP32, "fp64": L.SBO_FP64}[precision],
                          {"auto": L.SBO_KERNEL_AUTO, "simt": L.SBO_KERNEL_SIMT, "tcgen05": L.SBO_KERNEL_TCGEN05}[kernel],
                          0, max_obs, 0, max_cells)
        dev = np.ascontiguousarray(devices, dtype=np.int32)
        self._h = C.c_void_p()
        st = self.lib.sbo_group_create(C.byref(cfg), dev.ctypes.data_as(C.POINTER(C.c_int32)), dev.size, C.byref(self._h))
        if st != L.SBO_OK:
            msg = self.lib.sbo_group_last_error(None).decode()
            self._h = None
            raise L.SboError(st, msg)
        self.size = dev.size
        self.W = self.H = 0

    def _check(self, st):
        if st != L.SBO_OK:
            raise L.SboError(st, self.lib.sbo_group_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            self.lib.sbo_group_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, name, value):
        self._check(self.lib.sbo_group_set_option(self._h, L.OPTIONS[name], int(value)))

    def set_grid(self, x0, y0, rx, ry, W, H):
        self._check(self.lib.sbo_group_set_grid(self._h, x0, y0, rx, ry, W, H))
        self.W, self.H = W, H

    def set_observations(self, xy, y):
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1, 2)
        y = np.ascontiguousarray(y, dtype=np.float64).reshape(-1)
        self._check(self.lib.sbo_group_set_observations(self._h, _dptr(xy), _dptr(y), xy.shape[0]))

    def append_observations(self, xy, y):
        xy = np.ascontiguousarray(xy, dtype=np.float64).reshape(-1, 2)
        y = np.ascontiguousarray(y, dtype=np.float64).reshape(-1)
        self._check(self.lib.sbo_group_append_observations(self._h, _dptr(xy), _dptr(y), xy.shape[0]))

    @property
    def n_obs(self):
        return int(self.lib.sbo_group_num_observations(self._h))

    def member_factor(self, i):
        """(L^-1, alpha, (fit ms, ranks that shared the fit)) of member i -- for checking the shared fit."""
        h = C.c_void_p(self.lib.sbo_group_member(self._h, int(i)))
        n = self.n_obs
        linv, alpha = np.zeros((n, n)), np.zeros(n)
        ms, w = C.c_float(0.0), C.c_int32(0)
        for st in (self.lib.sbo_copy_linv(h, _dptr(linv)), self.lib.sbo_copy_alpha(h, _dptr(alpha)),
                   self.lib.sbo_fit_info(h, C.byref(ms), C.byref(w))):
            if st != L.SBO_OK:
                raise L.SboError(st, self.lib.sbo_last_error(h).decode())
        return linv, alpha, (float(ms.value), int(w.value))

    def bands(self):
        cuts = np.zeros(self.size + 1, dtype=np.int32)
        self._check(self.lib.sbo_group_bands(self._h, cuts.ctypes.data_as(C.POINTER(C.c_int32))))
        return cuts.tolist()

    def step(self, beta, f_min, mask_out=None, incremental=False):
        if mask_out is None:
            mask_out = np.empty(self.W * self.H, dtype=np.uint8)
        out = L.SboStepOut()
        fn = self.lib.sbo_group_step_incremental if incremental else self.lib.sbo_group_step
        self._check(fn(self._h, float(beta), float(f_min), _u8ptr(mask_out), C.byref(out)))
        return StepResult.from_c(out), mask_out

    def select_subgoal(self, frontier, goal_xy) -> int:
        fr = np.ascontiguousarray(frontier, dtype=np.int32).reshape(-1)
        cell = C.c_int32(-1)
        self._check(self.lib.sbo_group_select_subgoal(self._h, fr.ctypes.data_as(C.POINTER(C.c_int32)), fr.size,
                                                      float(goal_xy[0]), float(goal_xy[1]), C.byref(cell)))
        return int(cell.value)

    def mu(self):
        out = np.empty(self.W * self.H)
        self._check(self.lib.sbo_group_copy_mu_f64(self._h, _dptr(out)))
        return out

    def sigma2(self):
        out = np.empty(self.W * self.H)
        self._check(self.lib.sbo_group_copy_sigma2_f64(self._h, _dptr(out)))
        return out


def shard_rows(H: int, rank: int, world: int):
    """Contiguous band of grid rows owned by `rank`."""
    base, rem = divmod(H, world)
    r0 = rank * base + min(rank, rem)
    return r0, r0 + base + (1 if rank < rem else 0)


def fit_column_owner(c: int, world: int) -> int:
    """Rank that owns block column ``c`` of a shared fit: the snake deal 0 1 .. W-1 W-1 .. 1 0 of ``csrc/factor_dmma.cuh:
    ColMap`` (column c costs ~ nblk - c tiles per step of the factorisation; the snake evens that out between ranks)."""
    t = c % (2 * world)
    return t if t < world else 2 * world - 1 - t


def fit_columns_of(rank: int, world: int, nblk: int):
    """Block columns of ``rank`` in increasing order (``ColMap::col`` for j = 0 .. ``ColMap::count_le(nblk - 1)`` - 1)."""
    return [c for c in range(nblk) if fit_column_owner(c, world) == rank]


def plan_row_bands(p, parts: int, max_rows: int = 0, precision=None, kernel="auto"):
    """Cost-balanced row bands for problem ``p`` (anything with xy, x0, y0, rx, ry, W, H and the hyper-parameters) over
    ``parts`` GPUs: (row_cuts list[parts + 1], cost_share list[parts]).  Host arithmetic (sbo_plan_row_bands)."""
    lib = L.load()
    cfg = L.SboConfig(p.sigma_f, p.length, p.noise_var, p.prior_mean, 0,
                      {"fp32": L.SBO_FP32, "fp64": L.SBO_FP64}[precision or p.precision],
                      {"auto": L.SBO_KERNEL_AUTO, "simt": L.SBO_KERNEL_SIMT, "tcgen05": L.SBO_KERNEL_TCGEN05}[kernel], 0,
                      max(int(p.xy.shape[0]), 1), 0, 1)
    xy = np.ascontiguousarray(p.xy, dtype=np.float64).reshape(-1, 2)
    cuts = np.zeros(parts + 1, dtype=np.int32)
    share = np.zeros(parts)
    st = lib.sbo_plan_row_bands(C.byref(cfg), _dptr(xy), xy.shape[0], p.x0, p.y0, p.rx, p.ry, p.W, p.H, parts, int(max_rows),
                                cuts.ctypes.data_as(C.POINTER(C.c_int32)), _dptr(share))
    if st != L.SBO_OK:
        raise L.SboError(st, "sbo_plan_row_bands")
    return cuts.tolist(), share.tolist()


def merge_results(parts):
    """Fold per-shard StepResults with the library's order (value desc, index asc)."""
    def better(a, b):
        return a[0] >= 0 and (b[0] < 0 or a[1] > b[1] or (a[1] == b[1] and a[0] < b[0]))
    ucb, wid = parts[0].argmax_ucb, parts[0].argmax_width
    for p in parts[1:]:
        if better(p.argmax_ucb, ucb):
            ucb = p.argmax_ucb
        if better(p.argmax_width, wid):
            wid = p.argmax_width
    return StepResult(sum(p.n_cells for p in parts), sum(p.safe_count for p in parts), ucb, wid,
                      sum(p.clamped_sigma2 for p in parts), parts[0].n_obs, parts[0].kernel_used,
                      max(p.ms_posterior for p in parts), max(p.ms_sets for p in parts),
                      max(p.ms_comm for p in parts), max(p.ms_total for p in parts),
                      sum(p.low_sigma2 for p in parts), max(p.refreshed for p in parts), max(p.drift for p in parts))
