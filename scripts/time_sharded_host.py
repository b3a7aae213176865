#!/usr/bin/env python
"""Phases of a large host-pointer kNN call on the sharded handle: python scripts/time_sharded_host.py [ngpu] [pinned 0/1]"""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
from ompl_b200._lib import check, lib
ng = int(sys.argv[1]) if len(sys.argv) > 1 else 8
pinned = int(sys.argv[2]) if len(sys.argv) > 2 else 1
n, nq, k, d = 1_000_000, 65536, 10, 7
r = np.random.default_rng(1)
pts, qs = r.random((n, d)), r.random((nq, d))
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d), devices=list(range(ng)))
nn.add_rows(pts)
nn.set_option("profile", 1)
ids, dd, cnt = np.empty((nq, k), np.uint64), np.empty((nq, k)), np.empty(nq, np.uint32)
L = lib()
if pinned:
    import torch
    for a in (qs, ids, dd):
        assert int(torch.cuda.cudart().cudaHostRegister(a.ctypes.data, a.nbytes, 0)) == 0
for it in range(6):
    t0 = time.perf_counter()
    check(L.omplcu_nn_knn(nn.handle, qs.ctypes.data, nq, k, ids.ctypes.data, dd.ctypes.data, cnt.ctypes.data))
    t = time.perf_counter() - t0
    print(f"call {t * 1e3:.3f} ms  scan phase {nn.stat('last_scan_phase_ns') / 1e6:.3f}  merge phase {nn.stat('last_merge_ns') / 1e6:.3f}  "
          f"slowest scan kernel {nn.stat('last_scan_ns') / 1e6:.3f} ms", flush=True)
