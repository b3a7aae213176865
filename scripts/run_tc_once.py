#!/usr/bin/env python
"""One tensor-core kNN call for profiling: python scripts/run_tc_once.py N Q k d [reps]"""
import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
from ompl_b200._lib import check, lib
n, nq, k, d = (int(a) for a in sys.argv[1:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 2
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
for it in range(reps):
    check(lib().omplcu_nn_knn_device(nn.handle, dq.data_ptr(), nq, k, ids.data_ptr(), dd.data_ptr(), C.c_void_p(s.cuda_stream)))
    print(f"scan {nn.stat('last_scan_ns') / 1e6:.3f} ms, fallbacks {nn.stat('last_fallbacks')}, tc {nn.stat('last_tc')}", flush=True)
