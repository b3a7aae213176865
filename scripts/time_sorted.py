#!/usr/bin/env python
"""Does spatial coherence help the tensor-core scan?  Same kNN problem with the queries and / or the points
pre-sorted along a Morton curve (lanes of a warp then hold nearby queries, point tiles hold nearby points):
python scripts/time_sorted.py [N] [Q] [k] [d]"""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
from ompl_b200._lib import check, lib

n, nq, k, d = (int(a) for a in (sys.argv[1:5] if len(sys.argv) > 4 else "1000000 65536 10 7".split()))
r = np.random.default_rng(1)
pts, qs = r.uniform(0, 1, (n, d)), r.uniform(0, 1, (nq, d))


def morton(x, bits=3):
    q = np.minimum((x * (1 << bits)).astype(np.uint64), (1 << bits) - 1)
    key = np.zeros(len(x), dtype=np.uint64)
    for b in range(bits - 1, -1, -1):
        for c in range(x.shape[1]):
            key = (key << np.uint64(1)) | ((q[:, c] >> np.uint64(b)) & np.uint64(1))
    return np.argsort(key, kind="stable")


L = lib()
BITS = [int(a) for a in os.environ.get("BITS", "0 3").split()]
for sq, sp in [(b, 0) for b in BITS] + ([(3, 1)] if os.environ.get("POINTS") else []):
    P = pts[morton(pts)] if sp else pts
    Q = qs[morton(qs, sq)] if sq else qs
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(P)
    nn.set_option("profile", 1)
    nn.set_option("tc_scan", 1)
    dq = torch.from_numpy(np.ascontiguousarray(Q)).cuda()
    ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dd = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream()
    sc = []
    for it in range(4):
        check(L.omplcu_nn_knn_device(nn.handle, dq.data_ptr(), nq, k, ids.data_ptr(), dd.data_ptr(), C.c_void_p(s.cuda_stream)))
        torch.cuda.synchronize()
        sc.append(nn.stat("last_scan_ns") / 1e6)
    print(f"queries sorted (Morton bits per dimension) {sq} points sorted {sp}: scan {min(sc[1:]):.3f} ms  fallbacks {nn.stat('last_fallbacks')}", flush=True)
