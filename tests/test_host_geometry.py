"""N3: goal-in-polygon and polydist -- host arithmetic exported by libsbo.so (no GPU involved), checked
against hand-derived answers and against the independent restatements in oracle/oracle.py.

Reference: src/safe_bayesian_optimization_node.cpp:650-706, src/libraries/polygeom_lib.cpp:400-474.  The reference
holds no tests for either function (parity unpinned); the known answers below follow from reading its code."""
import math

import numpy as np
import pytest

from safe_bayesian_optimization_b200.optimizer import point_within_polygon, polydist

SQUARE = [(0.0, 0.0), (4.0, 0.0), (4.0, 4.0), (0.0, 4.0), (0.0, 0.0)]          # counter-clockwise, closed
HOLE = [(1.0, 1.0), (1.0, 2.0), (2.0, 2.0), (2.0, 1.0), (1.0, 1.0)]
# an "L": concave at (2, 2)
ELL = [(0.0, 0.0), (4.0, 0.0), (4.0, 2.0), (2.0, 2.0), (2.0, 4.0), (0.0, 4.0), (0.0, 0.0)]


@pytest.mark.parametrize("rings,p,expect", [
    ([SQUARE], (2.0, 2.0), True),
    ([SQUARE], (4.0, 2.0), False),                 # on an edge: boundary is not within
    ([SQUARE], (0.0, 0.0), False),                 # on a vertex
    ([SQUARE], (2.0, 4.0), False),                 # on the top edge (horizontal)
    ([SQUARE], (5.0, 2.0), False),
    ([SQUARE], (-1e-9, 2.0), False),
    ([SQUARE], (1e-9, 2.0), True),
    ([SQUARE[::-1]], (2.0, 2.0), True),            # clockwise ring: same answer
    ([SQUARE[:-1]], (3.999, 0.001), True),         # open ring gets its closing segment
    ([SQUARE, HOLE], (1.5, 1.5), False),           # inside the hole
    ([SQUARE, HOLE], (1.0, 1.5), False),           # on the hole's border
    ([SQUARE, HOLE], (3.0, 3.0), True),
    ([ELL], (3.0, 3.0), False),                    # in the notch
    ([ELL], (1.0, 3.0), True),
    ([ELL], (3.0, 1.0), True),
    ([ELL], (2.0, 3.0), False),                    # on the notch's vertical edge
    ([ELL], (2.0, 1.0), True),                     # shares x with the concave vertex, below it: inside
    ([], (0.0, 0.0), False),                       # empty polygon
])
def test_point_within_polygon_known_answers(oracle, rings, p, expect):
    assert point_within_polygon(rings, p) is expect
    assert oracle.point_within_polygon(rings, p) is expect


def test_point_within_polygon_random_against_oracle(oracle):
    rng = np.random.default_rng(5)
    for trial in range(20):
        k = int(rng.integers(3, 12))
        ang = np.sort(rng.uniform(0, 2 * np.pi, k))
        rad = rng.uniform(0.5, 2.0, k)                         # star-shaped, possibly concave
        ring = np.stack([rad * np.cos(ang), rad * np.sin(ang)], axis=1)
        if trial % 2:
            ring = ring[::-1]
        hole = 0.2 * np.array([(1, 0), (0, 1), (-1, 0), (0, -1)], dtype=np.float64)
        rings = [ring, hole] if trial % 3 == 0 and rad.min() > 0.45 else [ring]
        pts = rng.uniform(-2.2, 2.2, (400, 2))
        pts[:k] = ring                                          # vertices themselves: never within
        got = [point_within_polygon(rings, q) for q in pts]
        ref = [oracle.point_within_polygon(rings, q) for q in pts]
        assert got == ref
        assert not any(got[:k])


@pytest.mark.parametrize("ring,p,proj,dist", [
    # nearest VERTEX, not the nearest boundary point (reference :437 copies the vertex list into the "rolled" list)
    (SQUARE, (2.0, -1.0), (0.0, 0.0), math.sqrt(5.0)),                        # tie (0,0)/(4,0): first in ring order
    (SQUARE, (3.0, -1.0), (4.0, 0.0), math.sqrt(2.0)),
    (SQUARE, (2.0, 2.0), (0.0, 0.0), math.sqrt(8.0)),                         # four-way tie inside
    (SQUARE[::-1], (2.0, -1.0), (0.0, 0.0), math.sqrt(5.0)),                  # clockwise input is reversed by bg::correct
    ([(4.0, 0.0), (0.0, 0.0), (0.0, 4.0), (4.0, 4.0), (4.0, 0.0)], (2.0, -1.0), (4.0, 0.0), math.sqrt(5.0)),
    ([(0.0, 0.0), (4.0, 0.0), (4.0, 4.0), (0.0, 4.0)], (5.0, 5.0), (4.0, 4.0), math.sqrt(2.0)),   # open ring is closed first
    (ELL, (3.0, 3.0), (4.0, 2.0), math.sqrt(2.0)),                            # three-way tie in the notch: first in ring order
    (ELL, (2.9, 2.9), (2.0, 2.0), math.hypot(0.9, 0.9)),
    ([], (1.0, 1.0), (0.0, 0.0), 100000000.0),                                # "distance to empty set is infinity"
])
def test_polydist_known_answers(oracle, ring, p, proj, dist):
    got_p, got_d = polydist(np.array(ring, dtype=np.float64).reshape(-1, 2), p)
    ref_p, ref_d = oracle.polydist(ring, p)
    assert got_p == pytest.approx(proj, abs=1e-15) and got_d == pytest.approx(dist, rel=1e-15)
    assert got_p == ref_p and got_d == ref_d                                  # bit-identical to the restatement


def test_polydist_random_bit_exact_against_oracle(oracle):
    rng = np.random.default_rng(9)
    for trial in range(50):
        k = int(rng.integers(3, 40))
        ang = np.sort(rng.uniform(0, 2 * np.pi, k))
        rad = rng.uniform(0.5, 2.0, k)
        ring = np.stack([3.0 + rad * np.cos(ang), -1.0 + rad * np.sin(ang)], axis=1)
        if trial % 2:
            ring = ring[::-1].copy()
        if trial % 3 == 0:
            ring = np.concatenate([ring, ring[:1]])
        if trial % 7 == 0:
            ring = np.concatenate([ring[:2], ring[1:2], ring[2:]])             # a repeated vertex: zero-length edge (:440-442)
        for q in rng.uniform(-2, 6, (20, 2)):
            got_p, got_d = polydist(ring, q)
            ref_p, ref_d = oracle.polydist(ring.tolist(), q.tolist())
            assert got_p == ref_p and got_d == ref_d
            # and it is a vertex of the ring up to the rounding of (1 - w) v + w v
            assert np.min(np.hypot(ring[:, 0] - got_p[0], ring[:, 1] - got_p[1])) < 1e-14
