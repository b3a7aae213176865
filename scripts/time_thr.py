#!/usr/bin/env python
"""Threshold path (sample -> bound -> append scan -> exact epilogue) against the list scan:
python scripts/time_thr.py [N] [Q] [k] [d] [variants...].  Device-resident timing; both paths are exact, so
their ids and distances must be identical.  A variant is diag + 100 * tiles + 1000 * sample_step."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ompl_b200 as ob  # noqa: E402
from ompl_b200._lib import check, lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
d = int(sys.argv[4]) if len(sys.argv) > 4 else 7
variants = [int(a) for a in sys.argv[5:]] or [0]
r = np.random.default_rng(1)
pts, qs = r.random((n, d)), r.random((nq, d))
L = lib()
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
nn.set_option("profile", 1)
nn.set_option("tc_scan", 1)
dq = torch.from_numpy(qs).cuda()
ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
dd = torch.empty((nq, k), dtype=torch.float64, device="cuda")
s = torch.cuda.current_stream()


def run(label, reps=4):
    ts = []
    for it in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        check(L.omplcu_nn_knn_device(nn.handle, dq.data_ptr(), nq, k, ids.data_ptr(), dd.data_ptr(),
                                     C.c_void_p(s.cuda_stream)))
        e1.record(s)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    extra = ""
    if nn.stat("last_thr"):
        extra = (f" sample {nn.stat('last_sample_ns') / 1e6:.3f} main {nn.stat('last_main_ns') / 1e6:.3f} ms"
                 f" kept/query {nn.stat('last_kept') / nq:.1f} rank {nn.stat('last_rank')}")
    print(f"{label}: call {min(ts[1:]):.3f} ms (first {ts[0]:.3f}) scan {nn.stat('last_scan_ns') / 1e6:.3f} ms"
          f" fallbacks {nn.stat('last_fallbacks')} thr {nn.stat('last_thr')}{extra}"
          f"  {nq / min(ts[1:]) / 1e3:.2f} M queries/s", flush=True)
    return ids.cpu().numpy().copy(), dd.cpu().numpy().copy()


nn.set_option("thr_scan", 0)
ref = run(f"list scan n={n} q={nq} k={k} d={d}")
nn.set_option("thr_scan", 1)
for v in variants:
    nn.set_option("tc_variant", v % 100)
    nn.set_option("tc_tiles", (v // 100) % 10)
    nn.set_option("thr_sample_step", (v // 1000) or 16)
    got = run(f"thr variant {v}")
    if v % 100 in (0, 16):
        print("   ids equal:", bool((got[0] == ref[0]).all()), " dists equal:", bool((got[1] == ref[1]).all()), flush=True)
