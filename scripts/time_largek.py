#!/usr/bin/env python
"""Large-k kNN (FMT*'s k ~ 2100 on M2: 100K hypercube samples, fmt/src/FMT.cpp:327-337) with and without the
tensor-core filter in front of the exact collect path: python scripts/time_largek.py [N] [Q] [k] [d].
Device-resident timing; the two paths must agree bit for bit."""
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
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
k = int(sys.argv[3]) if len(sys.argv) > 3 else 2100
d = int(sys.argv[4]) if len(sys.argv) > 4 else 8
r = np.random.default_rng(21)
pts = r.random((n, d))
qs = pts[:nq] if nq <= n else r.random((nq, d))
L = lib()
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
dq = torch.from_numpy(np.ascontiguousarray(qs)).cuda()
ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
dd = torch.empty((nq, k), dtype=torch.float64, device="cuda")
s = torch.cuda.current_stream()
res = {}
for mode in ("filter", "exact"):
    nn.set_option("force_exact", 1 if mode == "exact" else 0)
    m = nq if mode == "filter" else min(nq, 8192)
    ts = []
    for it in range(3 if mode == "filter" else 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        check(L.omplcu_nn_knn_device(nn.handle, dq.data_ptr(), m, k, ids.data_ptr(), dd.data_ptr(), C.c_void_p(s.cuda_stream)))
        e1.record(s)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    res[mode] = (ids[:min(nq, 8192)].cpu().numpy().copy(), dd[:min(nq, 8192)].cpu().numpy().copy())
    print(f"{mode}: n={n} q={m} k={k} d={d}: {min(ts):.1f} ms  {m / min(ts) / 1e3:.3f} M queries/s  path {nn.stat('last_path')} "
          f"filter {nn.stat('last_filter')} redo {nn.stat('last_collect_redo')}", flush=True)
print("ids equal:", bool((res["filter"][0] == res["exact"][0]).all()), " dists equal:", bool((res["filter"][1] == res["exact"][1]).all()))
