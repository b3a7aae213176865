#!/usr/bin/env python
"""One SE(3) kNN call for profiling: python scripts/run_se3_once.py N Q k"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
n, nq, k = (int(a) for a in sys.argv[1:4])
r = np.random.default_rng(1)
def st(m):
    q = r.normal(size=(m, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    return np.hstack([r.uniform(-1, 1, (m, 3)), q])
pts, qs = st(n), st(nq)
nn = ob.NearestNeighborsCUDA(ob.SE3StateSpace(-1.0, 1.0))
nn.add_rows(pts)
for _ in range(2):
    nn.nearestKBatch(qs, k)
print("filter", nn.stat("last_filter"), "redo", nn.stat("last_collect_redo"))
