#!/usr/bin/env python
"""Times batched checkMotion on the M3 workload: SE3 in [-1,1]^3, 500 spheres + 500 boxes,
4-sphere robot, resolution 0.01.  Device-resident timing with CUDA events + parity on a sample."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
import ompl_b200 as ob  # noqa: E402
from ompl_b200._lib import check, lib  # noqa: E402
from oracle import oracle_py as orc  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
mode = sys.argv[2] if len(sys.argv) > 2 else "uniform"
sph, box = helpers.se3_obstacles()
robot = helpers.se3_robot()
a, b = helpers.se3_states(1, n), helpers.se3_states(2, n)
if mode == "short":  # PRM-like edges: neighbours at ~0.2 distance
    r = helpers.rng(3)
    b[:, :3] = np.clip(a[:, :3] + r.normal(0, 0.1, (n, 3)), -1, 1)
sp = ob.SE3StateSpace(-1.0, 1.0)
chk = ob.SE3ObstacleValidityChecker(robot, sph, box)
L = lib()
da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
dv = torch.empty(n, dtype=torch.uint8, device="cuda")
dt = torch.empty(n, dtype=torch.float64, device="cuda")
s = torch.cuda.current_stream()
for want_t in (False, True):
    ts = []
    for it in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        check(L.omplcu_check_motions_device(chk.handle, sp.handle, da.data_ptr(), db.data_ptr(), n, dv.data_ptr(),
                                            dt.data_ptr() if want_t else None, None, C.c_void_p(s.cuda_stream)))
        e1.record(s)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = min(ts[1:])
    print(f"n={n} mode={mode} last_valid={want_t}: {ms:.2f} ms  {n / ms * 1e3 / 1e6:.2f} M segments/s  "
          f"states in segments {chk.stat('last_states')}  evaluated {chk.stat('last_states_evaluated')}  "
          f"group {chk.stat('last_group')}  grid {chk.stat('grid_cells')} cells / {chk.stat('grid_items')} items  "
          f"valid {int(dv.sum())}")
m = min(n, 20000)
osp, ochk = orc.space_se3(-1.0, 1.0), orc.checker_se3(robot, sph, box)
t0 = time.perf_counter()
ov, ot = orc.check_motions(osp, ochk, a[:m], b[:m], last_valid=True, nthreads=os.cpu_count())
cpu = time.perf_counter() - t0
v, t = dv.cpu().numpy()[:m].astype(bool), dt.cpu().numpy()[:m]
print(f"parity on {m}: valid {np.array_equal(v, ov)}  last_t {np.array_equal(t[~ov], ot[~ov])};  "
      f"CPU oracle {m / cpu / 1e3:.1f} K segments/s on {os.cpu_count()} threads")
