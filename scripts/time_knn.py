#!/usr/bin/env python
"""Device-resident kNN timing for one index size (shard-size experiments): python scripts/time_knn.py N [Q] [k] [d]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ompl_b200 as ob  # noqa: E402
from ompl_b200._lib import check, lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
d = int(sys.argv[4]) if len(sys.argv) > 4 else 7
r = np.random.default_rng(1)
pts, qs = r.random((n, d)), r.random((nq, d))
L = lib()
for sample in (1, 0):
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    nn.set_option("profile", 1)
    nn.set_option("sample_thresholds", sample)
    dq = torch.from_numpy(qs).cuda()
    ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dd = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream()
    ts = []
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        check(L.omplcu_nn_knn_device(nn.handle, dq.data_ptr(), nq, k, ids.data_ptr(), dd.data_ptr(),
                                     C.c_void_p(s.cuda_stream)))
        e1.record(s)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    print(f"n={n} q={nq} k={k} sample_thr={sample}: call {min(ts[1:]):.3f} ms  scan {nn.stat('last_scan_ns') / 1e6:.3f} ms  "
          f"fallbacks {nn.stat('last_fallbacks')}  3d-flop rate {3 * d * n * nq / max(nn.stat("last_scan_ns"), 1) / 1e3:.1f} TFLOP/s; ideal at 63 TFLOP/s {3 * d * n * nq / 63e12 * 1e3:.3f} ms")
