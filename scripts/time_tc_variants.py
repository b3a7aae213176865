#!/usr/bin/env python
"""Scan-kernel time of the tensor-core path under the diagnostic variants (results are NOT valid for variants >= 4)."""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
from ompl_b200._lib import check, lib
n, nq, k, d = (int(a) for a in os.environ.get("SHAPE", "100000 65536 10 100").split())
r = np.random.default_rng(1)
pts, qs = r.uniform(-np.pi, np.pi, (n, d)), r.uniform(-np.pi, np.pi, (nq, d))
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
nn.set_option("profile", 1)
nn.set_option("tc_scan", 1)
dq = torch.from_numpy(qs).cuda()
ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
dd = torch.empty((nq, k), dtype=torch.float64, device="cuda")
s = torch.cuda.current_stream()
for v in [int(a) for a in sys.argv[1:]] or [0, 2, 4]:
    nn.set_option("tc_variant", v % 100)   # diag bits: 2 no selection, 4 no insertion
    nn.set_option("tc_tiles", v // 100)    # hundreds digit: resident query tiles limit
    sc = []
    for it in range(3):
        check(lib().omplcu_nn_knn_device(nn.handle, dq.data_ptr(), nq, k, ids.data_ptr(), dd.data_ptr(), C.c_void_p(s.cuda_stream)))
        sc.append(nn.stat("last_scan_ns") / 1e6)
    print(f"variant {v}: scan {min(sc[1:]):.3f} ms, fallbacks {nn.stat('last_fallbacks')}", flush=True)
