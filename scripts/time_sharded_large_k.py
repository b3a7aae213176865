"""Sharded handle, FMT*-sized k: the per-shard lists merged on the device (merge_sorted_kernel).
All shards on the GPUs present (several on one GPU when there is only one)."""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import ompl_b200.api as ob
from ompl_b200._lib import lib

c = C.c_int(0)
lib().omplcu_device_count(C.byref(c))
rng = np.random.default_rng(5)
d, n, nq, k = 8, 100000, 8192, 2100
pts = rng.random((n, d))
qs = rng.random((nq, d))
sp = ob.RealVectorStateSpace(d)
one = ob.NearestNeighborsCUDA(sp)
one.add_rows(pts)
for nsh in (4, 8):
    many = ob.NearestNeighborsCUDA(sp, devices=[i % max(c.value, 1) for i in range(nsh)])
    many.add_rows(pts)
    for nn, name in ((one, "one index"), (many, "%d shards on %d GPU(s)" % (nsh, c.value))):
        nn.nearestKBatch(qs[:64], k)
        t0 = time.perf_counter()
        ids, dist = nn.nearestKBatch(qs, k)
        t1 = time.perf_counter()
        print("%-24s %d queries, k = %d: %.1f ms" % (name, nq, k, (t1 - t0) * 1e3), flush=True)
        if nn is one:
            ref = (ids, dist)
        else:
            print("   equal to one index:", bool((ids == ref[0]).all() and (dist == ref[1]).all()), flush=True)
    del many
