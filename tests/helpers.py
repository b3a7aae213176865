"""Shared generators for the parity tests (seeded; the same arrays feed the
oracle and the CUDA path)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


def rng(seed):
    return np.random.default_rng(seed)


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


def random_quaternions(seed, n):
    """Uniform unit quaternions (Shoemake), stored x,y,z,w like SO3StateSpace."""
    r = rng(seed)
    u1, u2, u3 = r.random(n), r.random(n), r.random(n)
    q = np.stack(
        [
            np.sqrt(1 - u1) * np.sin(2 * np.pi * u2),
            np.sqrt(1 - u1) * np.cos(2 * np.pi * u2),
            np.sqrt(u1) * np.sin(2 * np.pi * u3),
            np.sqrt(u1) * np.cos(2 * np.pi * u3),
        ],
        axis=1,
    )
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def se3_states(seed, n, low=-1.0, high=1.0):
    pos = rng(seed).uniform(low, high, size=(n, 3))
    return np.hstack([pos, random_quaternions(seed + 17, n)])


def se2_states(seed, n, low=0.0, high=1.0):
    r = rng(seed)
    return np.hstack([r.uniform(low, high, size=(n, 2)), r.uniform(-np.pi, np.pi, size=(n, 1))])


def se3_obstacles(seed=7, nsph=500, nbox=500, low=-1.0, high=1.0, rmin=0.02, rmax=0.08):
    """BASELINE config 4 / SURVEY 8d M3: spheres + axis-aligned boxes."""
    r = rng(seed)
    sph = np.hstack([r.uniform(low, high, size=(nsph, 3)), r.uniform(rmin, rmax, size=(nsph, 1))])
    box = np.hstack([r.uniform(low, high, size=(nbox, 3)), r.uniform(rmin, rmax, size=(nbox, 3))])
    return sph, box


def se3_robot():
    """4 body-fixed spheres (ox, oy, oz, r)."""
    return np.array(
        [[0.0, 0.0, 0.0, 0.03], [0.05, 0.0, 0.0, 0.02], [-0.05, 0.0, 0.0, 0.02], [0.0, 0.04, 0.02, 0.015]]
    )


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
