"""Seeded synthetic workloads of SURVEY.md 8(d), shared by bench.py and the tests
(tests/helpers.py re-exports these, so the parity tests and the bench use the
same arrays)."""
import numpy as np


def rng(seed):
    return np.random.default_rng(seed)


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


def m3_segments(n, seed=1):
    """M3: n SE3 segments, endpoints uniform in [-1,1]^3 x SO(3)."""
    return se3_states(seed, n), se3_states(seed + 1, n)
