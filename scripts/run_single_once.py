"""A few single-query nearestK calls (for the ncu launch list): python scripts/run_single_once.py N d"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob
n, d = int(sys.argv[1]), int(sys.argv[2])
r = np.random.default_rng(1)
pts, qs = r.random((n, d)), r.random((16, d))
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
nn.add_rows(pts)
for i in range(6):
    nn.nearestKBatch(qs[i:i + 1], 10)
print("done")
