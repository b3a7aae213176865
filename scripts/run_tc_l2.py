"""One batched kNN with / without the persisting L2 window (for ncu): python scripts/run_tc_l2.py N d window"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
n, d, w = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
r = np.random.default_rng(1)
pts, qs = r.random((n, d)), r.random((65536, d))
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
nn.set_option("l2_window", w)
nn.set_option("profile", 1)
for i in range(4):
    nn.nearestKBatch(qs, 10)
    print("scan ms", nn.stat("last_scan_ns") / 1e6, flush=True)
