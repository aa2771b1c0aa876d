// hostgeom.cpp -- the two O(#vertices) host computations that sit between the chosen frontier cell and the
// published subgoal in the reference node.  Plain C++ (no CUDA, no Boost): they stay on the
// host in the reference too; restating them lets a host without Boost.Geometry finish the step.
//
//   sbo_point_within_polygon  = bg::within(goal_point, eroded_concave_polygon_)
//                               (reference src/safe_bayesian_optimization_node.cpp:657): winding strategy of
//                               Boost.Geometry for cartesian points -- boundary points are NOT within; a point
//                               inside or on a hole is not within.
//   sbo_polydist              = polydist(eroded_poly_for_projection, raw_subgoal_point)
//                               (src/libraries/polygeom_lib.cpp:400-474, called at node :677-691 after
//                               bg::correct on a counter-clockwise, closed polygon type, include/polygeom_lib.h:60).
//
// polydist as written in the reference does not project onto the edges: its "rolled" vertex list is filled with
// VertexListRolled[j] = VertexList[j] (:437), i.e. it is the vertex list itself, so the candidate for vertex i,
// (1 - w_i) V_i + w_i V_i, is V_i up to rounding and the function returns the NEAREST VERTEX (first minimum in
// corrected ring order).  The restatement keeps that arithmetic operation for operation -- results identical to
// the reference are the contract, not a better projection.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <vector>

#include "../../include/sbo.h"

namespace {

struct Pt { double x, y; };

// Boost's math::equals for doubles: |a - b| <= eps * max(1, |a|, |b|)
inline bool fp_equals(double a, double b) {
  if (a == b) return true;
  return std::fabs(a - b) <= DBL_EPSILON * std::max(1.0, std::max(std::fabs(a), std::fabs(b)));
}

// side_by_triangle: +1 point left of s1->s2, -1 right, 0 collinear (epsilon scaled by the operands)
inline int side_of(const Pt& s1, const Pt& s2, const Pt& p) {
  const double dx = s2.x - s1.x, dy = s2.y - s1.y, dpx = p.x - s1.x, dpy = p.y - s1.y;
  const double s = dx * dpy - dy * dpx;
  const double factor = std::max(1.0, std::max(std::max(std::fabs(dx), std::fabs(dy)), std::max(std::fabs(dpx), std::fabs(dpy))));
  if (std::fabs(s) <= DBL_EPSILON * factor) return 0;
  return s > 0 ? 1 : -1;
}

// winding of one ring around p: returns -1 outside, 0 on the boundary, +1 inside
int ring_winding(const Pt* r, int n, const Pt& p) {
  if (n < 3) return -1;
  int count = 0;
  const bool closed = (r[0].x == r[n - 1].x && r[0].y == r[n - 1].y);
  const int nseg = closed ? n - 1 : n;   // an open ring gets its closing segment
  for (int i = 0; i < nseg; ++i) {
    const Pt& s1 = r[i];
    const Pt& s2 = r[(i + 1) % n];
    const bool eq1 = fp_equals(s1.x, p.x), eq2 = fp_equals(s2.x, p.x);
    if (eq1 && eq2) {   // vertical segment through p.x: touches when p.y lies within it
      if ((s1.y <= p.y && s2.y >= p.y) || (s2.y <= p.y && s1.y >= p.y)) return 0;
      continue;
    }
    int c = 0;
    if (eq1) c = s2.x > p.x ? 1 : -1;
    else if (eq2) c = s1.x > p.x ? -1 : 1;
    else if (s1.x < p.x && s2.x > p.x) c = 2;
    else if (s2.x < p.x && s1.x > p.x) c = -2;
    if (c == 0) continue;
    const int side = side_of(s1, s2, p);   // +1 left of s1->s2, -1 right
    if (side == 0) return 0;               // on the segment: touches
    if (side * c > 0) count += c;
  }
  return count == 0 ? -1 : 1;
}

// twice the signed area (positive = counter-clockwise), shoelace over the closed ring
double ring_area2(const std::vector<Pt>& r) {
  double a = 0.0;
  for (size_t i = 0; i + 1 < r.size(); ++i) a += r[i].x * r[i + 1].y - r[i + 1].x * r[i].y;
  return a;
}

}  // namespace

extern "C" {

int32_t sbo_point_within_polygon(const double* rings_xy, const int32_t* ring_sizes, int32_t n_rings, double px,
                                 double py, int32_t* within) {
  if (!within || n_rings < 0 || (n_rings > 0 && (!rings_xy || !ring_sizes))) return SBO_ERR_INVALID_ARG;
  *within = 0;
  if (n_rings == 0) return SBO_OK;   // empty polygon: nothing is within
  const Pt p{px, py};
  const Pt* r = reinterpret_cast<const Pt*>(rings_xy);
  for (int32_t k = 0; k < n_rings; ++k) {
    if (ring_sizes[k] < 0) return SBO_ERR_INVALID_ARG;
    const int w = ring_winding(r, ring_sizes[k], p);
    if (k == 0) {
      if (w != 1) return SBO_OK;       // outside the outer ring or on it
    } else if (w != -1) {
      return SBO_OK;                   // inside a hole or on its border
    }
    r += ring_sizes[k];
  }
  *within = 1;
  return SBO_OK;
}

int32_t sbo_polydist(const double* ring_xy, int32_t nv, double px, double py, double* proj_xy, double* dist) {
  if (!proj_xy || !dist || nv < 0 || (nv > 0 && !ring_xy)) return SBO_ERR_INVALID_ARG;
  // "Distance to empty set is infinity" (:419-423; the reference then indexes the empty list -- undefined; here it returns)
  proj_xy[0] = 0.0; proj_xy[1] = 0.0; *dist = 100000000.0;
  if (nv == 0) return SBO_OK;
  std::vector<Pt> ring(reinterpret_cast<const Pt*>(ring_xy), reinterpret_cast<const Pt*>(ring_xy) + nv);
  // bg::correct (:683): close the ring (more than two points, end points differ), then make it counter-clockwise
  if (ring.size() > 2 && !(ring.front().x == ring.back().x && ring.front().y == ring.back().y)) ring.push_back(ring.front());
  if (ring_area2(ring) < 0.0) std::reverse(ring.begin(), ring.end());
  ring.pop_back();   // :427
  const size_t n = ring.size();
  if (n == 0) return SBO_OK;
  std::vector<Pt> d(n);
  std::vector<double> norm(n);
  for (size_t i = 0; i < n; ++i) {   // :433-443: d[j] = V[j] - V[j-1], the edge arriving at vertex j
    const size_t j = (i + 1) % n;
    d[j] = Pt{ring[j].x - ring[i].x, ring[j].y - ring[i].y};
    norm[j] = std::sqrt(d[j].x * d[j].x + d[j].y * d[j].y);
    if (norm[j] == 0.0) norm[j] = 1.0;
  }
  double best = 0.0;
  Pt best_pt{0.0, 0.0};
  for (size_t i = 0; i < n; ++i) {   // :449-459
    const double n2 = std::pow(norm[i], 2);
    const double wt = (px - ring[i].x) * (d[i].x / n2) + (py - ring[i].y) * (d[i].y / n2);
    const double w = std::max(std::min(wt, 1.0), 0.0);
    const Pt c{(1 - w) * ring[i].x + w * ring[i].x, (1 - w) * ring[i].y + w * ring[i].y};   // rolled list == list (:437)
    const double ex = px - c.x, ey = py - c.y;
    const double di = std::sqrt(ex * ex + ey * ey);
    if (i == 0 || di < best) { best = di; best_pt = c; }   // std::min_element: first minimum (:462-463)
  }
  proj_xy[0] = best_pt.x; proj_xy[1] = best_pt.y; *dist = best;
  return SBO_OK;
}

}  // extern "C"
