#!/usr/bin/env python
"""Latency of the literal drop-in call, nearestK of ONE query (PRM::addMilestone, prm/src/PRM.cpp:575), and of
small batches, through the host-pointer C-ABI call: python scripts/time_single.py [N] [d] [k]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ompl_b200 as ob  # noqa: E402
from oracle import oracle_py as orc  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 7
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
r = np.random.default_rng(1)
pts, qs = r.random((n, d)), r.random((4096, d))
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
nn.set_option("profile", 1)
for nq in (1, 8, 64, 128, 512):
    reps = 200
    nn.nearestKBatch(qs[:nq], k)
    t0 = time.perf_counter()
    for i in range(reps):
        ids, dd = nn.nearestKBatch(qs[i:i + nq], k)
    t = (time.perf_counter() - t0) / reps
    oi, od = orc.knn(orc.space_rv(d), pts, qs[reps - 1:reps - 1 + nq][:16], k, nthreads=8)
    ok = bool((ids[:16] == oi).all() and (dd[:16] == od).all())
    # the bare C-ABI call with preallocated buffers (no numpy work per call)
    import ctypes as C
    from ompl_b200._lib import check, lib
    L = lib()
    qb = np.ascontiguousarray(qs[:nq])
    ib, db, cb = np.empty((nq, k), np.uint64), np.empty((nq, k)), np.empty(nq, np.uint32)
    t0 = time.perf_counter()
    for i in range(reps):
        check(L.omplcu_nn_knn(nn.handle, qb.ctypes.data, nq, k, ib.ctypes.data, db.ctypes.data, cb.ctypes.data))
    tc = (time.perf_counter() - t0) / reps
    print(f"n={n} d={d} k={k} nq={nq}: {t * 1e6:.0f} us per call ({tc * 1e6:.0f} us in omplcu_nn_knn)  scan {nn.stat('last_scan_ns') / 1e3:.0f} us  "
          f"fallbacks {nn.stat('last_fallbacks')}  matches oracle {ok}", flush=True)
