#!/usr/bin/env python
"""Dense low-dimensional sets: share of queries the tensor-core filter cannot prove, and what the second pass costs."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
n, nq, k, d = (int(a) for a in (sys.argv[1:5] if len(sys.argv) > 4 else "1000000 8192 10 3".split()))
r = np.random.default_rng(1)
lo, hi = (float(a) for a in os.environ.get("RANGE", "0 1").split())
pts, qs = r.uniform(lo, hi, (n, d)), r.uniform(lo, hi, (nq, d))
for mode in [int(a) for a in os.environ.get("MODES", "-1 1 0").split()]:
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.set_option("tc_scan", mode)
    nn.add_rows(pts)
    nn.nearestKBatch(qs, k)
    print(f"  first call: tc {nn.stat('last_tc')} rescans {nn.stat('last_rescans')} fallbacks {nn.stat('last_fallbacks')} demoted {nn.stat('tc_demoted')}")
    t0 = time.perf_counter()
    ids, dist = nn.nearestKBatch(qs, k)
    dt = time.perf_counter() - t0
    print(f"tc_scan={mode}: {dt * 1e3:.2f} ms  tc {nn.stat('last_tc')} rescans {nn.stat('last_rescans')} fallbacks {nn.stat('last_fallbacks')}", flush=True)
