#!/usr/bin/env python
"""kNN on compound / rotation spaces (exact path): python scripts/time_se3_knn.py [N] [Q] [k] [space: se3|se2|so3]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import ompl_b200 as ob
from oracle import oracle_py as orc

n, nq, k = (int(a) for a in (sys.argv[1:4] if len(sys.argv) > 3 else "100000 4096 10".split()))
space = sys.argv[4] if len(sys.argv) > 4 else "se3"
r = np.random.default_rng(1)


def rand_states(m):
    if space == "se3":
        q = r.normal(size=(m, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
        return np.hstack([r.uniform(-1, 1, (m, 3)), q])
    if space == "se2":
        return np.hstack([r.uniform(-1, 1, (m, 2)), r.uniform(-np.pi, np.pi, (m, 1))])
    q = r.normal(size=(m, 4)); return q / np.linalg.norm(q, axis=1, keepdims=True)


sp = {"se3": lambda: ob.SE3StateSpace(-1.0, 1.0), "se2": lambda: ob.SE2StateSpace(-1.0, 1.0), "so3": lambda: ob.SO3StateSpace()}[space]()
osp = {"se3": lambda: orc.space_se3(-1.0, 1.0), "se2": lambda: orc.space_se2(-1.0, 1.0), "so3": lambda: orc.space_so3()}[space]()
pts, qs = rand_states(n), rand_states(nq)
nn = ob.NearestNeighborsCUDA(sp)
nn.add_rows(pts)
for m in (1, 64, nq):
    nn.nearestKBatch(qs[:m], k)
    t0 = time.perf_counter()
    reps = 20 if m <= 64 else 3
    for _ in range(reps):
        ids, d = nn.nearestKBatch(qs[:m], k)
    t = (time.perf_counter() - t0) / reps
    print(f"{space} n={n} nq={m} k={k}: {t * 1e3:.3f} ms per call, {m / t / 1e3:.1f} K queries/s  path {nn.stat("last_path")} proven {nn.stat("last_proven")} redo {nn.stat("last_collect_redo")}", flush=True)
oi, od = orc.knn(osp, pts, qs[:64], k, nthreads=8)
print("ids equal:", bool((ids[:64] == oi).all()))
