#!/usr/bin/env python
"""nearestR timing on the chain metric: python scripts/time_radius.py [links] [nq]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ompl_b200 as ob
links = int(sys.argv[1]) if len(sys.argv) > 1 else 10
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
n = 100_000
r = np.random.default_rng(links)
s = np.vstack([r.uniform(-np.pi, np.pi, (n // 2, links)), r.normal(0, 0.15, (n // 2, links))])
nn = ob.NearestNeighborsCUDA(ob.KinematicChainSpace(links, 1.0 / links))
nn.add_rows(s)
q = s[:: n // nq][:nq]
for it in range(3):
    t0 = time.perf_counter()
    ids, d, offs = nn.nearestRBatch(q, 0.5 if links == 10 else 0.35)
    t = time.perf_counter() - t0
    print(f"links={links} nq={nq}: {t * 1e3:.1f} ms, {offs[-1]} neighbours, launches {nn.stat('launches')}")
