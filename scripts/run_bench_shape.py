"""A few batched kNN calls on one shape (for ncu captures): python scripts/run_bench_shape.py N d [k] [nq]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
n, d = int(sys.argv[1]), int(sys.argv[2])
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
nq = int(sys.argv[4]) if len(sys.argv) > 4 else 65536
pts = np.random.default_rng(1).random((n, d))
qs = np.random.default_rng(1001).random((nq, d))
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
nn.set_option("profile", 1)
for i in range(4):
    nn.nearestKBatch(qs, k)
    print("scan ms", nn.stat("last_scan_ns") / 1e6, "fallbacks", nn.stat("last_fallbacks"), flush=True)
