#!/usr/bin/env python
"""Generate tests/golden/*.npz.  Runs in the build container only (needs
/root/reference for the circle-world fixture files and oracle/_ref, the
reference's own NearestNeighbors headers compiled unmodified).  The outputs are
committed; nothing on the GPU box reads /root/reference.

  circle_world.npz    inputs: tests/resources/circle_obstacles.txt (70 circles),
                      circle_queries.txt (100 start/goal pairs); expected
                      booleans / lastValid times from the oracle (the reference
                      stores none, SURVEY.md 4 consequence 3)
  knn_linear_*.npz    neighbour ids returned by the REAL NearestNeighborsLinear
                      (and checked equal to NearestNeighborsGNAT) on seeded inputs
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle_py as orc  # noqa: E402
import helpers  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def circle_world():
    circ = []
    with open(os.path.join(REF, "tests/resources/circle_obstacles.txt")) as f:
        lines = f.read().splitlines()[2:]  # circles2D.h:74-78 skips two header lines
    for ln in lines:
        p = ln.split()
        if len(p) >= 4:
            circ.append([float(p[1]), float(p[2]), float(p[3])])
    circ = np.array(circ)
    q = []
    with open(os.path.join(REF, "tests/resources/circle_queries.txt")) as f:
        toks = f.read().split()
    # "<id> start: x y" / "<id> goal: x y"  (circles2D.h:118-132)
    for i in range(0, len(toks) - 7, 8):
        q.append([float(toks[i + 2]), float(toks[i + 3]), float(toks[i + 6]), float(toks[i + 7])])
    q = np.array(q)
    low = [float((circ[:, 0] - circ[:, 2]).min()), float((circ[:, 1] - circ[:, 2]).min())]
    high = [float((circ[:, 0] + circ[:, 2]).max()), float((circ[:, 1] + circ[:, 2]).max())]
    sp = orc.space_rv(2, low, high, 0.01)
    chk = orc.checker_circles(circ)
    s1, s2 = q[:, :2].copy(), q[:, 2:].copy()
    v_start = orc.check_states(chk, s1)
    v_goal = orc.check_states(chk, s2)
    mv = orc.check_motions(sp, chk, s1, s2)
    mv2, t = orc.check_motions(sp, chk, s1, s2, last_valid=True)
    assert (mv == mv2).all()
    nd = orc.segment_count(sp, s1, s2)
    np.savez(os.path.join(OUT, "circle_world.npz"), circles=circ, queries=q, low=np.array(low), high=np.array(high),
             start_valid=v_start, goal_valid=v_goal, motion_valid=mv, last_t=t, nd=nd)
    print("circle_world: circles", circ.shape, "queries", q.shape, "valid motions", int(mv.sum()))


def knn_golden():
    # RealVector(7), the BASELINE synthetic config at a size Linear finishes in seconds
    sp = orc.space_rv(7)
    pts = helpers.uniform_points(1, 5000, 7)
    qs = helpers.uniform_points(1001, 64, 7)
    lin = orc.RefNN(orc.RefNN.LINEAR, sp)
    lin.add(pts)
    ids, _ = lin.knn(qs, 10)
    gn = orc.RefNN(orc.RefNN.GNAT, sp)
    gn.add(pts)
    gids, _ = gn.knn(qs, 10)
    assert (ids == gids).all(), "GNAT and Linear disagree"
    oi, od = orc.knn(sp, pts, qs, 10)
    assert (oi == ids).all(), "oracle and reference Linear disagree"
    np.savez(os.path.join(OUT, "knn_linear_rv7.npz"), points=pts, queries=qs, k=10, ids=ids, dists=od)
    print("knn_linear_rv7:", ids.shape)

    # SE3 with bounds [0,1]^3, n = 200, k = 10: tests/datastructures/nearestneighbors.cpp:59-85
    sp3 = orc.space_se3(0.0, 1.0)
    st = helpers.se3_states(5, 200, 0.0, 1.0)
    lin = orc.RefNN(orc.RefNN.LINEAR, sp3)
    lin.add(st)
    ids, _ = lin.knn(st, 10)
    gn = orc.RefNN(orc.RefNN.GNAT_SMALL, sp3)
    gn.add(st)
    gids, _ = gn.knn(st, 10)
    assert (ids == gids).all()
    oi, od = orc.knn(sp3, st, st, 10)
    assert (oi == ids).all()
    np.savez(os.path.join(OUT, "knn_linear_se3.npz"), points=st, queries=st, k=10, ids=ids, dists=od)
    print("knn_linear_se3:", ids.shape)


def ref_spaces_dmv():
    """Outputs of the reference's OWN compiled StateSpace / spaces / DiscreteMotionValidator code
    (oracle/_ref/libompl_refbase.so): distance, interpolate(0.37), validSegmentCount, checkMotion
    booleans and lastValid.second.  Validity predicates are the oracle's demo-checker restatements."""
    out = {}
    r = helpers.rng(21)
    n = 400
    # SE3 in [-1,1]^3 vs 500 spheres + 500 boxes (BASELINE config 4 shape), resolution 0.01
    rb = orc.RefBase("se3", 0, -1.0, 1.0)
    sph, box = helpers.se3_obstacles()
    robot = helpers.se3_robot()
    chk = orc.checker_se3(robot, sph, box)
    a, b = helpers.se3_states(31, n), helpers.se3_states(32, n)
    b[::2, :3] = a[::2, :3] + r.normal(0, 0.02, (n // 2, 3))
    b[::2, 3:] = a[::2, 3:]
    v, t, _ = rb.check_motions(chk, a, b, True, True)
    out.update(se3_a=a, se3_b=b, se3_dist=rb.distance(a, b), se3_interp=rb.interpolate(a, b, 0.37),
               se3_nd=rb.segment_count(a, b), se3_valid=v, se3_last_t=t, se3_sph=sph, se3_box=box, se3_robot=robot)
    # SE2 [0,1]^2 vs three circles
    rb = orc.RefBase("se2", 0, 0.0, 1.0)
    circ = np.array([[0.5, 0.5, 0.2], [0.2, 0.8, 0.1], [0.8, 0.2, 0.15]])
    chk = orc.checker_circles(circ)
    a, b = helpers.se2_states(33, n), helpers.se2_states(34, n)
    v, t, _ = rb.check_motions(chk, a, b, True, True)
    out.update(se2_a=a, se2_b=b, se2_dist=rb.distance(a, b), se2_interp=rb.interpolate(a, b, 0.37),
               se2_nd=rb.segment_count(a, b), se2_valid=v, se2_last_t=t, se2_circ=circ)
    # hypercube d = 8, resolution 0.001 (demos/HypercubeBenchmark.cpp:93-97)
    rb = orc.RefBase("rv", 8, 0.0, 1.0, fraction=0.001)
    chk = orc.checker_hypercube(8)
    a = helpers.hypercube_valid_states(35, n, 8)
    b = np.clip(a + r.normal(0, 0.03, a.shape), 0, 1)
    b[::3] = helpers.hypercube_valid_states(36, len(b[::3]), 8)
    v, t, _ = rb.check_motions(chk, a, b, True, True)
    out.update(hc_a=a, hc_b=b, hc_dist=rb.distance(a, b), hc_nd=rb.segment_count(a, b), hc_valid=v, hc_last_t=t)
    # SO3 alone (arc length + slerp)
    rb = orc.RefBase("so3")
    a, b = helpers.random_quaternions(37, n), helpers.random_quaternions(38, n)
    out.update(so3_a=a, so3_b=b, so3_dist=rb.distance(a, b), so3_interp=rb.interpolate(a, b, 0.37),
               so3_nd=rb.segment_count(a, b))
    np.savez_compressed(os.path.join(OUT, "ref_spaces_dmv.npz"), **out)
    print("ref_spaces_dmv: se3 valid", int(out["se3_valid"].sum()), "se2 valid", int(out["se2_valid"].sum()),
          "hypercube valid", int(out["hc_valid"].sum()))


if __name__ == "__main__":
    circle_world()
    knn_golden()
    ref_spaces_dmv()
