import os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import ompl_b200 as ob
r = np.random.default_rng(1)
def st(m):
    q = r.normal(size=(m, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    return np.hstack([r.uniform(-1, 1, (m, 3)), q])
for n in (100000, 1000000):
    pts, qs = st(n), st(256)
    nn = ob.NearestNeighborsCUDA(ob.SE3StateSpace(-1.0, 1.0)); nn.add_rows(pts)
    for m in (1, 4, 8, 9, 12, 15, 16, 24, 32, 64, 128, 256):
        nn.nearestKBatch(qs[:m], 10)
        t0 = time.perf_counter()
        for _ in range(10): nn.nearestKBatch(qs[:m], 10)
        print(f"n={n} nq={m}: {(time.perf_counter()-t0)/10*1e3:.3f} ms  few {nn.stat('last_few')} filter {nn.stat('last_filter')}", flush=True)
