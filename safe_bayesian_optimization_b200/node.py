"""``OptimizerCore`` -- the step logic of the reference's ``OptimizerNode`` with
the hot path behind libsbo.so.

Member and method names follow
/root/reference/src/safe_bayesian_optimization_node.cpp so that a reader of the
node finds the same things here:

=========================  =====================================================
reference                  here
=========================  =====================================================
``process_terrain_map``    :625-647 -- two forms: ``process_terrain_map`` takes
                           the service response arrays (mu/std computed
                           elsewhere, exactly the reference's boundary) and
                           ``process_observations`` computes the posterior too
``ComputeSets``            :399-416 -- GPU (``sbo_compute_sets`` / ``sbo_step``)
``FindSafetyContourIndices``  :418-497 -- host, OpenCV, incl. the reference's
                           int-truncated bounds and dropped last row/column
``GetNextSubgoal``         :499-550 -- contour on host, selection on the GPU
=========================  =====================================================
"""
from __future__ import annotations

import logging

import numpy as np

from .optimizer import SboContext

log = logging.getLogger("safe_bayesian_optimization_b200")


def find_safety_contour_indices(mask_img: np.ndarray, x0, y0, rx, ry):
    """Host stage that consumes the 0/255 mask image (H x W, cell = iy*W+ix).

    Reproduces :431-492: bounds truncated to int, pixel = int((c-min)/(max-min)*dim),
    cells whose pixel falls outside are dropped, a later cell overwrites an earlier
    one on the same pixel, external contours only, CHAIN_APPROX_NONE, pixel -> cell
    through the last writer.  Because the grid is regular the pixel maps are
    separable: one 1-D map for columns and one for rows.
    """
    import cv2
    H, W = mask_img.shape
    xs = x0 + np.arange(W, dtype=np.float64) * rx
    ys = y0 + np.arange(H, dtype=np.float64) * ry
    min_x, max_x, min_y, max_y = int(xs.min()), int(xs.max()), int(ys.min()), int(ys.max())
    if max_x == min_x or max_y == min_y:
        raise ZeroDivisionError("grid extent below one unit: the reference divides by zero here (:450-451)")
    px = ((xs - min_x) / (max_x - min_x) * W).astype(np.int64)
    py = ((ys - min_y) / (max_y - min_y) * H).astype(np.int64)
    okx = (px >= 0) & (px < W)
    oky = (py >= 0) & (py < H)
    # last writer per pixel column / row (cells are visited in increasing index order)
    col_of_px = -np.ones(W, dtype=np.int64)
    col_of_px[px[okx]] = np.arange(W)[okx]
    row_of_py = -np.ones(H, dtype=np.int64)
    row_of_py[py[oky]] = np.arange(H)[oky]
    img = np.zeros((H, W), dtype=np.uint8)
    vr, vc = row_of_py >= 0, col_of_px >= 0
    img[np.ix_(vr, vc)] = np.where(mask_img[np.ix_(row_of_py[vr], col_of_px[vc])] != 0, 255, 0)
    contours, _ = cv2.findContours(img, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_NONE)
    if not contours:
        return np.zeros(0, dtype=np.int32)
    pts = np.concatenate([c.reshape(-1, 2) for c in contours], axis=0)
    ix, iy = col_of_px[pts[:, 0]], row_of_py[pts[:, 1]]
    keep = (ix >= 0) & (iy >= 0)
    return (iy[keep] * W + ix[keep]).astype(np.int32)


def contour_indices_from_image(img: np.ndarray, col_of_px: np.ndarray, row_of_py: np.ndarray):
    """cv::findContours (:461-462) on the GPU-built safety image, contour pixels mapped back to cell indices through
    the two pixel -> cell tables (what :481-492 does with its hash map)."""
    import cv2
    H, W = img.shape
    contours, _ = cv2.findContours(img, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_NONE)
    if not contours:
        return np.zeros(0, dtype=np.int32)
    pts = np.concatenate([c.reshape(-1, 2) for c in contours], axis=0)
    ix, iy = col_of_px[pts[:, 0]].astype(np.int64), row_of_py[pts[:, 1]].astype(np.int64)
    keep = (ix >= 0) & (iy >= 0)
    return (iy[keep] * W + ix[keep]).astype(np.int32)


class OptimizerCore:
    """One safe-BO step at a time, single caller thread (reference :1503)."""

    def __init__(self, beta=2.0, f_min=500.0, *, ctx: SboContext | None = None, **ctx_kwargs):
        # opt.beta / opt.f_min defaults: config/safe_bayesian_optimization.yaml:4,6
        self.beta_ = float(beta)
        self.f_min_ = float(f_min)
        self.ctx = ctx if ctx is not None else SboContext(**ctx_kwargs)
        self.mu_ = self.std_ = self.S_ = self.Q_ = None
        self.grid = None
        self.current_goal_ = (0.0, 0.0)          # default-constructed PointStamped (:171)
        self.terrain_width_cells_ = self.terrain_height_cells_ = 0
        self.last = None
        self._mask_on_device = False             # True after a step that left its mask in the context (sbo_step)

    # -- goal_point_callback (:1324-1327) --------------------------------------
    def goal_point_callback(self, x, y):
        self.current_goal_ = (float(x), float(y))

    def _set_grid(self, x0, y0, rx, ry, W, H):
        self.grid = (float(x0), float(y0), float(rx), float(ry), int(W), int(H))
        self.terrain_width_cells_, self.terrain_height_cells_ = int(W), int(H)
        self.ctx.set_grid(x0, y0, rx, ry, W, H)

    # -- process_terrain_map (:625-647): response arrays in, sets out ------------
    def process_terrain_map(self, x_coords, y_coords, values, uncertainties, n_width_cells, n_height_cells):
        x = np.asarray(x_coords, dtype=np.float64); y = np.asarray(y_coords, dtype=np.float64)
        W, H = int(n_width_cells), int(n_height_cells)
        if x.size == 0 or x.size != W * H:
            log.warning("terrain map response is empty or inconsistent")   # :618-621 style soft failure
            return False
        rx = (x[1] - x[0]) if W > 1 else 1.0
        ry = (y[W] - y[0]) if H > 1 else 1.0
        self._set_grid(x[0], y[0], rx, ry, W, H)
        self.mu_ = np.asarray(values, dtype=np.float64)
        self.std_ = np.asarray(uncertainties, dtype=np.float64)
        self.ComputeSets()
        return True

    def ComputeSets(self):
        """:399-407 on the arrays held in mu_/std_ (GPU: sbo_compute_sets)."""
        res, mask, q0, q1 = self.ctx.compute_sets(self.mu_, self.std_, self.beta_, self.f_min_)
        self.S_ = mask
        self.Q_ = np.stack([q0, q1], axis=1)
        self.last = res
        self._mask_on_device = False
        return res

    # -- the integrated path: observations in, sets out ---------------------------
    def process_observations(self, xy, y, grid, incremental=False):
        """grid = (x0, y0, rx, ry, W, H).  Fits (or appends to) the GP and runs one step."""
        if self.grid != tuple(grid):
            self._set_grid(*grid)
            incremental = False
        if incremental:
            self.ctx.append_observations(xy, y)
        else:
            self.ctx.set_observations(xy, y)
        res, mask = self.ctx.step(self.beta_, self.f_min_, incremental=incremental)
        self.S_ = mask
        self.mu_ = self.std_ = self.Q_ = None      # stay on the device unless asked for
        self.last = res
        self._mask_on_device = True
        return res

    def FindSafetyContourIndices(self):
        if self.S_ is None or self.S_.size == 0:
            log.warning("D_ or S_ is empty")     # :419-422
            return np.zeros(0, dtype=np.int32)
        x0, y0, rx, ry, W, H = self.grid
        if getattr(self, "last", None) is not None and self.ctx.row0 == 0 and self.ctx.row1 == H and self._mask_on_device:
            img, col_of_px, row_of_py = self.ctx.frontier_image()     # image + tables built on the GPU (frontier.cu)
            return contour_indices_from_image(img, col_of_px, row_of_py)
        return find_safety_contour_indices(self.S_.reshape(H, W), x0, y0, rx, ry)

    def GetNextSubgoal(self) -> int:
        frontier = self.FindSafetyContourIndices()
        if frontier.size == 0:
            log.warning("No frontier points found")   # :503-506
            return -1
        return self.ctx.select_subgoal(frontier, self.current_goal_)

    # -- the tail of process_terrain_map (:650-706): goal or projected frontier subgoal ---------------
    def ComputeSubgoal(self, eroded_polygon_rings):
        """eroded_polygon_rings = [outer ring, holes...] of ``eroded_concave_polygon_`` (the alpha-shape / erosion
        stage stays with the host's CGAL/Boost).  Returns (x, y) or None ("No valid subgoal found", :700-703)."""
        from . import _lib as L
        frontier = np.zeros(0, dtype=np.int32)
        from .optimizer import point_within_polygon
        if not point_within_polygon(eroded_polygon_rings, self.current_goal_):
            frontier = self.FindSafetyContourIndices()      # only needed when the goal is not safe (:667)
        kind, xy, cell, dist = self.ctx.subgoal_point(frontier, self.current_goal_, eroded_polygon_rings)
        if kind == L.SBO_SUBGOAL_NONE:
            log.warning("No valid subgoal found")
            return None
        if kind == L.SBO_SUBGOAL_GOAL:
            log.info("Goal is within eroded safe region, using goal as subgoal: (%f, %f)", *xy)
        else:
            log.info("Projected subgoal point: (%f, %f), distance: %f", xy[0], xy[1], dist)
        return xy

    # -- SpatialMeasurement ingestion + the spatial_data_size trigger (:552-566) -----------------------
    def spatial_measurement_callback(self, x, y, value, uncertainty=0.0):
        size, _ = self.ctx.ingest([[x, y]], [value], [uncertainty])
        return self.spatial_data_size_callback(size)

    def spatial_data_size_callback(self, current_size) -> bool:
        """True when the size changed, i.e. when the reference requests a new terrain map (= runs a step)."""
        if current_size != getattr(self, "last_spatial_data_size_", 0):
            log.info("Spatial data size changed: %d -> %d", getattr(self, "last_spatial_data_size_", 0), current_size)
            self.last_spatial_data_size_ = current_size
            return True
        return False
