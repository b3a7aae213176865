#!/usr/bin/env python
"""Tensor-core scan vs the FP32 register-tiled scan on the d >= 64 workload (M4 plain L2):
python scripts/time_tc.py [N] [Q] [k] [d].  Device-resident timing; both paths are exact, so their
ids and distances must be identical."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ompl_b200 as ob  # noqa: E402
from ompl_b200._lib import check, lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
d = int(sys.argv[4]) if len(sys.argv) > 4 else 100
r = np.random.default_rng(1)
pts, qs = r.uniform(-np.pi, np.pi, (n, d)), r.uniform(-np.pi, np.pi, (nq, d))
L = lib()
res = {}
for tc in (1, 0):
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    nn.set_option("profile", 1)
    nn.set_option("tc_scan", tc)
    dq = torch.from_numpy(qs).cuda()
    ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dd = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream()
    ts, sc = [], []
    for it in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        check(L.omplcu_nn_knn_device(nn.handle, dq.data_ptr(), nq, k, ids.data_ptr(), dd.data_ptr(),
                                     C.c_void_p(s.cuda_stream)))
        e1.record(s)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
        sc.append(nn.stat("last_scan_ns") / 1e6)
    res[tc] = (ids.cpu().numpy(), dd.cpu().numpy())
    print(f"tc={tc} (ran tc: {nn.stat('last_tc')}) n={n} q={nq} k={k} d={d}: call {min(ts[1:]):.3f} ms (first {ts[0]:.3f})  "
          f"scan {min(sc[1:]):.3f} ms  fallbacks {nn.stat('last_fallbacks')}  "
          f"{nq / min(ts[1:]) / 1e3:.2f} M queries/s  3d-flop rate {3 * d * n * nq / min(sc[1:]) / 1e9:.1f} TFLOP/s", flush=True)
print("ids equal:", bool((res[0][0] == res[1][0]).all()), " dists equal:", bool((res[0][1] == res[1][1]).all()))
