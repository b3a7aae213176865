"""Shared generators for the parity tests (seeded; the same arrays feed the
oracle and the CUDA path)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


from ompl_b200.workloads import (m3_segments, random_quaternions, rng, se3_obstacles,  # noqa: E402,F401
                                 se3_robot, se3_states)


def uniform_points(seed, n, d, low=0.0, high=1.0):
    return rng(seed).uniform(low, high, size=(n, d))


def with_duplicates(pts, seed, frac=0.01):
    """Replace a fraction of rows by exact copies of other rows (forces ties)."""
    r = rng(seed)
    pts = pts.copy()
    m = max(1, int(len(pts) * frac))
    dst = r.choice(len(pts), size=m, replace=False)
    src = r.integers(0, len(pts), size=m)
    pts[dst] = pts[src]
    return pts


def se2_states(seed, n, low=0.0, high=1.0):
    r = rng(seed)
    return np.hstack([r.uniform(low, high, size=(n, 2)), r.uniform(-np.pi, np.pi, size=(n, 1))])


def hypercube_valid_states(seed, n, d, edge=0.1):
    """States on the narrow passage of demos/HypercubeBenchmark.cpp (valid by
    construction), plus uniform ones are produced by the caller."""
    r = rng(seed)
    out = np.empty((n, d))
    for i in range(n):
        k = r.integers(0, d)
        s = np.empty(d)
        # dims above k stay inside the edge, dim k roams, dims below k hug the far face
        s[:k] = r.uniform(1 - edge, 1, size=k)
        s[k] = r.uniform(0, 1)
        s[k + 1:] = r.uniform(0, edge, size=d - k - 1)
        out[i] = s
    return out


def load_circles():
    """tests/golden/circle_world.npz: the reference's 70-circle world and its
    100 start/goal queries (tests/resources/circle_obstacles.txt, circle_queries.txt)."""
    z = np.load(os.path.join(GOLDEN, "circle_world.npz"))
    return z


def ulp_diff(a, b):
    a = np.ascontiguousarray(a, dtype=np.float64).view(np.int64)
    b = np.ascontiguousarray(b, dtype=np.float64).view(np.int64)
    return np.abs(a - b)
