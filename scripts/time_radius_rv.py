#!/usr/bin/env python
"""nearestR on a plain RealVector roadmap with and without the tensor-core filter (host-pointer API, as
NearestNeighborsCUDA::nearestRBatch calls it): python scripts/time_radius_rv.py [N] [Q] [d] [mean neighbours].
PRM*/BIT* use r ~ (log n / n)^(1/d); the radius here is chosen for the requested mean neighbour count."""
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ompl_b200 as ob  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
d = int(sys.argv[3]) if len(sys.argv) > 3 else 7
mean = float(sys.argv[4]) if len(sys.argv) > 4 else 10.0
r = np.random.default_rng(1)
pts, qs = r.random((n, d)), r.random((nq, d))
unit_ball = math.pi ** (d / 2) / math.gamma(d / 2 + 1)
rad = (mean / (n * unit_ball)) ** (1.0 / d)
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
res = {}
for mode in ("filter", "exact"):
    nn.set_option("force_exact", 1 if mode == "exact" else 0)
    m = nq if mode == "filter" else min(nq, 2048)
    best = 1e30
    for it in range(3 if mode == "filter" else 2):
        t0 = time.perf_counter()
        ids, dd, offs = nn.nearestRBatch(qs[:m], rad)
        best = min(best, time.perf_counter() - t0)
    res[mode] = (ids, dd, offs)
    print(f"{mode}: n={n} q={m} d={d} r={rad:.4f}: {best * 1e3:.1f} ms  {m / best / 1e6:.3f} M queries/s  "
          f"{offs[-1] / m:.1f} neighbours/query  filter {nn.stat('last_filter')}", flush=True)
m = min(nq, 2048)
a, b = res["filter"], res["exact"]
e = int(b[2][m])
print("offsets equal:", bool((a[2][:m + 1] == b[2][:m + 1]).all()), " ids equal:", bool((a[0][:e] == b[0][:e]).all()),
      " dists equal:", bool((a[1][:e] == b[1][:e]).all()))
