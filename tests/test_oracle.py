"""Pins the parity oracle (CPU only).

  (a) the reference's known-answer values, tests/base/state_spaces.cpp:120-166,197-240,266-305
      and the generic properties of tests/base/StateSpaceTest.h:71-117;
  (b) the committed fixtures under tests/golden/ (neighbour ids returned by the
      REAL NearestNeighborsLinear / GNAT, circle-world booleans);
  (c) when oracle/_ref was built in this container, a live differential run
      against the reference's own NearestNeighborsLinear / NearestNeighborsGNAT,
      mirroring tests/datastructures/nearestneighbors.cpp:117-287.
"""
import math
import os

import numpy as np
import pytest

import helpers
from oracle import oracle_py as orc

PI = math.pi


# ---------------------------------------------------------------- (a) known answers
def test_so2_known_answers():
    sp = orc.space_so2()
    assert sp.comp[0].lvs == PI * 0.01  # getMaximumExtent() == PI, state_spaces.cpp:130
    d = orc.distance(sp, [[-PI + 0.1]], [[PI - 0.1]])[0]
    assert abs(d - 0.2) < 1e-3
    assert orc.distance(sp, [[PI - 0.1]], [[PI - 0.1]])[0] == 0.0
    s3 = orc.interpolate(sp, [[PI - 0.08]], [[-PI + 0.1]], 0.5)[0, 0]
    assert abs(s3 - (-PI + 0.01)) < 1e-3
    assert abs(orc.distance(sp, [[0.1]], [[PI - 0.1]])[0] - (PI - 0.2)) < 1e-3
    assert abs(orc.interpolate(sp, [[PI - 0.1]], [[0.1]], 0.5)[0, 0] - PI / 2) < 1e-3
    assert orc.interpolate(sp, [[0.5]], [[1.5]], 0.0)[0, 0] == 0.5
    assert abs(orc.interpolate(sp, [[0.5]], [[1.5]], 1.0)[0, 0] - 1.5) < 1e-3


def test_realvector_known_answers():
    sp = orc.space_rv(3)
    z = np.zeros((1, 3))
    assert orc.distance(sp, z, z)[0] == 0.0
    one = np.array([[0.0, 0.0, 1.0]])
    assert orc.interpolate(sp, z, one, 0.5)[0, 2] == 0.5
    assert (orc.interpolate(sp, z, z, 0.6) == z).all()
    sp1 = orc.space_rv(1, 1.0, 2.0)
    assert abs(sp1.comp[0].lvs / 0.01 - 1.0) < 1e-3  # getMaximumExtent, state_spaces.cpp:301-303


def test_so3_interpolated_states_stay_unit():
    sp = orc.space_so3()
    assert sp.comp[0].lvs == 0.5 * PI * 0.01  # state_spaces.cpp:208
    q = helpers.random_quaternions(3, 2)
    for j in range(102):
        s = orc.interpolate(sp, q[0:1], q[1:2], j / 101.0)[0]
        assert abs(np.linalg.norm(s) - 1.0) < 1e-15 * 8  # 1e-15 in the reference; allow sqrt rounding
    assert orc.distance(sp, q[0:1], q[0:1])[0] == 0.0


@pytest.mark.parametrize("name", ["rv", "so2", "so3", "se2", "se3"])
def test_generic_space_properties(name):
    """StateSpaceTest.h:71-117: symmetry, positivity, interpolation endpoints, midpoint."""
    n = 300
    if name == "rv":
        sp, a, b = orc.space_rv(5), helpers.uniform_points(1, n, 5), helpers.uniform_points(2, n, 5)
    elif name == "so2":
        sp = orc.space_so2()
        a, b = helpers.rng(1).uniform(-PI, PI, (n, 1)), helpers.rng(2).uniform(-PI, PI, (n, 1))
    elif name == "so3":
        sp, a, b = orc.space_so3(), helpers.random_quaternions(1, n), helpers.random_quaternions(2, n)
    elif name == "se2":
        sp, a, b = orc.space_se2(), helpers.se2_states(1, n), helpers.se2_states(2, n)
    else:
        sp, a, b = orc.space_se3(), helpers.se3_states(1, n, 0, 1), helpers.se3_states(2, n, 0, 1)
    dab, dba = orc.distance(sp, a, b), orc.distance(sp, b, a)
    assert (dab > 0).all() and np.allclose(dab, dba, atol=1e-15)
    assert (orc.distance(sp, a, a) == 0).all()
    i0, i1 = orc.interpolate(sp, a, b, 0.0), orc.interpolate(sp, a, b, 1.0)
    assert (orc.distance(sp, a, i0) < 1e-12).all()
    assert (orc.distance(sp, b, i1) < 1e-7).all()  # SO3 slerp at t=1 is exact up to sin rounding
    mid = orc.interpolate(sp, a, b, 0.5)
    assert np.allclose(orc.distance(sp, a, mid) + orc.distance(sp, mid, b), dab, atol=1e-7)


def test_valid_segment_count_rules():
    # StateSpace.cpp:851-854 and the compound max rule :1085-1097
    sp = orc.space_rv(2, 0.0, 1.0, 0.01)
    lvs = math.sqrt(2.0) * 0.01
    a, b = np.array([[0.0, 0.0]]), np.array([[0.3, 0.4]])
    assert orc.segment_count(sp, a, b)[0] == math.ceil(0.5 / lvs)
    assert orc.segment_count(sp, a, a)[0] == 0
    se3 = orc.space_se3(-1.0, 1.0)
    s = helpers.se3_states(4, 50)
    t = helpers.se3_states(5, 50)
    rv3 = orc.space_rv(3, -1.0, 1.0)
    so3 = orc.space_so3()
    want = np.maximum(orc.segment_count(rv3, s[:, :3], t[:, :3]), orc.segment_count(so3, s[:, 3:], t[:, 3:]))
    assert (orc.segment_count(se3, s, t) == want).all()


# ---------------------------------------------------------------- checkers / DiscreteMotionValidator
def test_hypercube_checker_cases():
    # demos/HypercubeBenchmark.cpp:57-72, edgeWidth 0.1
    chk = orc.checker_hypercube(4)
    s = np.array([
        [0.05, 0.05, 0.05, 0.05],  # nothing above the edge: valid
        [0.95, 0.95, 0.95, 0.5],   # last dim leaves the edge, all lower dims >= 0.9: valid
        [0.5, 0.95, 0.95, 0.5],    # dim 0 below 0.9 after a found: invalid
        [0.05, 0.05, 0.5, 0.05],   # found at dim 2, dims 1,0 < 0.9: invalid
        [0.9, 0.9, 0.5, 0.05],     # boundary: s >= 0.9 passes (strict <)
        [0.1, 0.1, 0.1, 0.1],      # boundary: s > 0.1 is strict, not found
    ])
    assert orc.check_states(chk, s).tolist() == [True, True, False, False, True, True]


def test_motion_validator_matches_pointwise_definition():
    """checkMotion == (s2 valid and every j/nd valid); lastValid == first failing j."""
    sp = orc.space_rv(2, 0.0, 1.0, 0.01)
    circ = np.array([[0.5, 0.5, 0.2], [0.2, 0.8, 0.1]])
    chk = orc.checker_circles(circ)
    a, b = helpers.uniform_points(1, 400, 2), helpers.uniform_points(2, 400, 2)
    v = orc.check_motions(sp, chk, a, b)
    v2, t = orc.check_motions(sp, chk, a, b, last_valid=True)
    assert (v == v2).all() and 0 < v.sum() < len(v)
    nd = orc.segment_count(sp, a, b)
    for i in range(len(a)):
        n = int(nd[i])
        first_bad = None
        for j in range(1, n):
            st = orc.interpolate(sp, a[i:i + 1], b[i:i + 1], j / n)
            if not orc.check_states(chk, st)[0]:
                first_bad = j
                break
        s2ok = orc.check_states(chk, b[i:i + 1])[0]
        assert v[i] == (first_bad is None and s2ok)
        if first_bad is not None:
            assert t[i] == (first_bad - 1) / n
        elif not s2ok:
            assert t[i] == (n - 1) / n


def test_circle_world_fixture():
    z = helpers.load_circles()
    sp = orc.space_rv(2, z["low"], z["high"], 0.01)
    chk = orc.checker_circles(z["circles"])
    q = z["queries"]
    assert z["circles"].shape == (70, 3) and q.shape == (100, 4)
    assert (orc.check_states(chk, q[:, :2]) == z["start_valid"]).all()
    assert (orc.check_states(chk, q[:, 2:]) == z["goal_valid"]).all()
    # every start/goal of the reference's query file is collision free (the planners need that)
    assert z["start_valid"].all() and z["goal_valid"].all()
    v, t = orc.check_motions(sp, chk, q[:, :2], q[:, 2:], last_valid=True)
    assert (v == z["motion_valid"]).all() and (t[~v] == z["last_t"][~v]).all()
    assert (orc.segment_count(sp, q[:, :2], q[:, 2:]) == z["nd"]).all()


# ---------------------------------------------------------------- (b) kNN fixtures from the real Linear
@pytest.mark.parametrize("fx,space", [("knn_linear_rv7.npz", "rv7"), ("knn_linear_se3.npz", "se3")])
def test_knn_matches_reference_linear_fixture(fx, space):
    z = np.load(os.path.join(helpers.GOLDEN, fx))
    sp = orc.space_rv(7) if space == "rv7" else orc.space_se3(0.0, 1.0)
    ids, d = orc.knn(sp, z["points"], z["queries"], int(z["k"]), nthreads=4)
    assert (ids == z["ids"]).all()
    assert (d == z["dists"]).all()


def test_knn_tie_break_is_insertion_index():
    sp = orc.space_rv(3)
    pts = np.zeros((20, 3))
    pts[10:] = 0.5
    ids, d = orc.knn(sp, pts, np.zeros((1, 3)), 15)
    assert ids[0].tolist() == list(range(15))
    ridx, rd = orc.radius(sp, pts, np.zeros(3), 0.0)
    assert ridx.tolist() == list(range(10))  # inclusive radius, NearestNeighborsLinear.h:139


# ---------------------------------------------------------------- (c) live differential vs the reference headers
needs_ref = pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built (reference tree absent)")


@needs_ref
@pytest.mark.parametrize("kind", [orc.RefNN.LINEAR, orc.RefNN.GNAT, orc.RefNN.GNAT_SMALL])
@pytest.mark.parametrize("space", ["rv7", "se3"])
def test_oracle_vs_reference_nn(kind, space):
    """stateSpaceTest, tests/datastructures/nearestneighbors.cpp:117-206 (n=200, k=10)."""
    n, k = 200, 10
    if space == "rv7":
        sp, st = orc.space_rv(7), helpers.uniform_points(11, n, 7)
    else:
        sp, st = orc.space_se3(0.0, 1.0), helpers.se3_states(12, n, 0.0, 1.0)
    nn = orc.RefNN(kind, sp)
    nn.add(st)
    assert nn.size() == n and len(nn.list()) == n
    rid, _ = nn.knn(st, k)
    oid, od = orc.knn(sp, st, st, k)
    assert (rid == oid).all()  # continuous data: no ties, identity must agree
    assert (oid[:, 0] == np.arange(n)).all() and (od[:, 0] == 0).all()
    # nearestR(inf) returns everything sorted with self first (:163-173)
    r = nn.radius(st[5], float("inf"))
    oi, _ = orc.radius(sp, st, st[5], float("inf"))
    assert len(r) == n and r[0] == 5 and (r == oi).all()
    # nearestK(2n) returns n (:175-184)
    r2, _ = nn.knn(st[:3], 2 * n)
    o2, _ = orc.knn(sp, st, st[:3], 2 * n)
    assert r2.shape == (3, n) and (r2 == o2).all()
    # nearest() == element 0 of nearestK
    assert nn.nearest(st[7]) == 7


@needs_ref
def test_reference_nearest_on_empty_throws():
    nn = orc.RefNN(orc.RefNN.LINEAR, orc.space_rv(2))
    with pytest.raises(RuntimeError, match="No elements found"):
        nn.nearest(np.zeros(2))


@needs_ref
def test_reference_remove_semantics():
    sp = orc.space_rv(2)
    pts = helpers.uniform_points(3, 50, 2)
    nn = orc.RefNN(orc.RefNN.LINEAR, sp)
    nn.add(pts)
    assert nn.remove(7) and not nn.remove(7) and nn.size() == 49
    lst = nn.list().tolist()
    assert lst == [i for i in range(50) if i != 7]  # insertion order kept, :86-96
    rid, _ = nn.knn(pts[7:8], 3)
    keep = np.array(lst)
    oid, _ = orc.knn(sp, pts[keep], pts[7:8], 3)
    assert (keep[oid[0].astype(int)] == rid[0]).all()


@needs_ref
def test_oracle_equals_linear_when_no_tie_at_boundary():
    """Duplicate points force ties; Linear (unstable partial_sort) must still
    return the same DISTANCES rank by rank, and the same ids wherever the
    distance is unique."""
    sp = orc.space_rv(4)
    pts = helpers.with_duplicates(helpers.uniform_points(21, 600, 4), 22, frac=0.2)
    qs = pts[:100]
    lin = orc.RefNN(orc.RefNN.LINEAR, sp)
    lin.add(pts)
    rid, _ = lin.knn(qs, 12)
    oid, od = orc.knn(sp, pts, qs, 12)
    rd = np.stack([orc.distance(sp, pts[rid[i].astype(int)], np.repeat(qs[i:i + 1], 12, 0)) for i in range(len(qs))])
    assert (rd == od).all()
    uniq = np.ones_like(od, dtype=bool)
    uniq[:, 1:] &= od[:, 1:] != od[:, :-1]
    uniq[:, :-1] &= od[:, :-1] != od[:, 1:]
    uniq[:, -1] = False  # boundary membership may differ under a tie with rank k+1
    assert (rid[uniq] == oid[uniq]).all()


# ---------------------------------------------------------------- (d) the reference's own state spaces + DiscreteMotionValidator
needs_refbase = pytest.mark.skipif(not orc.have_refbase(), reason="oracle/_ref/libompl_refbase.so not built")


def _ref_and_oracle(kind):
    if kind == "rv":
        return orc.RefBase("rv", 7), orc.space_rv(7), lambda s, n: helpers.uniform_points(s, n, 7)
    if kind == "so2":
        return orc.RefBase("so2"), orc.space_so2(), lambda s, n: helpers.rng(s).uniform(-PI, PI, (n, 1))
    if kind == "so3":
        return orc.RefBase("so3"), orc.space_so3(), helpers.random_quaternions
    if kind == "se2":
        return orc.RefBase("se2", 0, 0.0, 1.0), orc.space_se2(), helpers.se2_states
    return orc.RefBase("se3", 0, -1.0, 1.0), orc.space_se3(-1.0, 1.0), helpers.se3_states


@needs_refbase
@pytest.mark.parametrize("kind", ["rv", "so2", "so3", "se2", "se3"])
def test_oracle_space_math_equals_reference_code(kind):
    """distance / interpolate / validSegmentCount of the restatement are bit-identical to the
    reference's compiled StateSpace.cpp + spaces/src/*.cpp (same libm, same flags)."""
    rb, osp, gen = _ref_and_oracle(kind)
    n = 4000
    a, b = gen(1, n), gen(2, n)
    a[:50] = b[:50]  # zero distances, nd = 0
    assert rb.stride == osp.stride
    assert (rb.distance(a, b) == orc.distance(osp, a, b)).all()
    for t in (0.0, 0.125, 0.37, 0.5, 1.0):
        assert (rb.interpolate(a, b, t) == orc.interpolate(osp, a, b, t)).all()
    assert (rb.segment_count(a, b) == orc.segment_count(osp, a, b)).all()
    # longest valid segment: extent * 0.01 per (sub)space, StateSpace.cpp:237-248
    if not osp.compound:
        assert rb.longest_valid_segment() == osp.comp[0].lvs


@needs_refbase
def test_oracle_segment_count_factor_and_resolution():
    rb = orc.RefBase("rv", 8, 0.0, 1.0, fraction=0.001, factor=3)
    osp = orc.space_rv(8, 0.0, 1.0, 0.001)
    osp.comp[0].factor = 3
    a, b = helpers.uniform_points(3, 2000, 8), helpers.uniform_points(4, 2000, 8)
    assert (rb.segment_count(a, b) == orc.segment_count(osp, a, b)).all()


@needs_refbase
@pytest.mark.parametrize("case", ["circles_rv2", "hypercube_rv8", "se3_obstacles", "se2_circles"])
def test_oracle_motion_validator_equals_reference_code(case):
    """Both checkMotion overloads of the restated DiscreteMotionValidator against the reference's
    compiled DiscreteMotionValidator.cpp (booleans, lastValid.second, lastValid.first, counters)."""
    r = helpers.rng(11)
    if case == "circles_rv2":
        z = helpers.load_circles()
        rb, osp = orc.RefBase("rv", 2, z["low"], z["high"]), orc.space_rv(2, z["low"], z["high"])
        chk = orc.checker_circles(z["circles"])
        a = np.vstack([z["queries"][:, :2], r.uniform(z["low"], z["high"], (1500, 2))])
        b = np.vstack([z["queries"][:, 2:], r.uniform(z["low"], z["high"], (1500, 2))])
    elif case == "hypercube_rv8":
        rb, osp = orc.RefBase("rv", 8, 0.0, 1.0, fraction=0.001), orc.space_rv(8, 0.0, 1.0, 0.001)
        chk = orc.checker_hypercube(8)
        a = helpers.hypercube_valid_states(1, 600, 8)
        b = np.clip(a + r.normal(0, 0.03, a.shape), 0, 1)
        b[::3] = helpers.hypercube_valid_states(3, len(b[::3]), 8)
    elif case == "se3_obstacles":
        rb, osp = orc.RefBase("se3", 0, -1.0, 1.0), orc.space_se3(-1.0, 1.0)
        sph, box = helpers.se3_obstacles()
        chk = orc.checker_se3(helpers.se3_robot(), sph, box)
        a, b = helpers.se3_states(1, 1500), helpers.se3_states(2, 1500)
        b[::2, :3] = a[::2, :3] + r.normal(0, 0.02, (750, 3))
        b[::2, 3:] = a[::2, 3:]
    else:
        rb, osp = orc.RefBase("se2", 0, 0.0, 1.0), orc.space_se2()
        chk = orc.checker_circles(np.array([[0.5, 0.5, 0.2], [0.2, 0.8, 0.1], [0.8, 0.2, 0.15]]))
        a, b = helpers.se2_states(1, 1500), helpers.se2_states(2, 1500)
    rv = rb.check_motions(chk, a, b)
    assert tuple(rb.counts) == (int(rv.sum()), int((~rv).sum()))  # one count per segment, :139-142
    rv2, rt, rst = rb.check_motions(chk, a, b, last_valid=True, last_state=True)
    ov, ot = orc.check_motions(osp, chk, a, b, last_valid=True)
    assert (rv == ov).all() and (rv2 == ov).all() and 0 < ov.sum() < len(ov)
    assert (rt[~ov] == ot[~ov]).all()
    for i in np.where(~ov)[0][:100]:
        ok, t, st = orc.check_motion_last(osp, chk, a[i], b[i])
        assert not ok and t == rt[i] and (st == rst[i]).all()


def test_oracle_matches_reference_spaces_dmv_fixture():
    """tests/golden/ref_spaces_dmv.npz was produced by the reference's compiled code
    (scripts/gen_golden.py::ref_spaces_dmv); it pins the restatement where _ref cannot travel."""
    z = np.load(os.path.join(helpers.GOLDEN, "ref_spaces_dmv.npz"))
    se3, se2, so3 = orc.space_se3(-1.0, 1.0), orc.space_se2(), orc.space_so3()
    hc = orc.space_rv(8, 0.0, 1.0, 0.001)
    for name, sp in (("se3", se3), ("se2", se2), ("so3", so3)):
        a, b = z[name + "_a"], z[name + "_b"]
        assert (orc.distance(sp, a, b) == z[name + "_dist"]).all()
        assert (orc.interpolate(sp, a, b, 0.37) == z[name + "_interp"]).all()
        assert (orc.segment_count(sp, a, b) == z[name + "_nd"]).all()
    assert (orc.distance(hc, z["hc_a"], z["hc_b"]) == z["hc_dist"]).all()
    assert (orc.segment_count(hc, z["hc_a"], z["hc_b"]) == z["hc_nd"]).all()
    cases = (
        ("se3", se3, orc.checker_se3(z["se3_robot"], z["se3_sph"], z["se3_box"])),
        ("se2", se2, orc.checker_circles(z["se2_circ"])),
        ("hc", hc, orc.checker_hypercube(8)),
    )
    for name, sp, chk in cases:
        v, t = orc.check_motions(sp, chk, z[name + "_a"], z[name + "_b"], last_valid=True)
        assert (v == z[name + "_valid"]).all() and (t[~v] == z[name + "_last_t"][~v]).all()
