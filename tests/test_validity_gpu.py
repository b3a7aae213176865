"""GPU parity tests for batched state validity and batched motion validity
(through the C ABI) against the oracle's restatement of the demo checkers and
DiscreteMotionValidator (base/src/DiscreteMotionValidator.cpp:48-145).

Bar: booleans and lastValid times bit-identical.
"""
import numpy as np
import pytest

import helpers
import ompl_b200 as ob
from oracle import oracle_py as orc

pytestmark = pytest.mark.gpu
PI = np.pi


def _spaces(name):
    if name == "rv8":
        return ob.RealVectorStateSpace(8), orc.space_rv(8)
    if name == "so2":
        return ob.SO2StateSpace(), orc.space_so2()
    if name == "so3":
        return ob.SO3StateSpace(), orc.space_so3()
    if name == "se2":
        return ob.SE2StateSpace(), orc.space_se2()
    if name == "se3":
        return ob.SE3StateSpace(-1.0, 1.0), orc.space_se3(-1.0, 1.0)
    if name == "chain10":
        return ob.KinematicChainSpace(10, 0.1), orc.space_chain(10, 0.1)
    raise KeyError(name)


def _states(name, seed, n):
    if name == "rv8":
        return helpers.uniform_points(seed, n, 8)
    if name == "so2":
        return helpers.rng(seed).uniform(-PI, PI, (n, 1))
    if name == "so3":
        return helpers.random_quaternions(seed, n)
    if name == "se2":
        return helpers.se2_states(seed, n)
    if name == "se3":
        return helpers.se3_states(seed, n)
    return helpers.rng(seed).uniform(-PI, PI, (n, 10))


# ---------------------------------------------------------------- space math
@pytest.mark.parametrize("name", ["rv8", "so2", "se2"])
def test_space_math_bit_exact(gpu, name):
    sp, osp = _spaces(name)
    a, b = _states(name, 1, 3000), _states(name, 2, 3000)
    assert (sp.distance(a, b) == orc.distance(osp, a, b)).all()
    for t in (0.0, 0.25, 0.5, 1.0):
        assert (sp.interpolate(a, b, t) == orc.interpolate(osp, a, b, t)).all()
    assert (sp.validSegmentCount(a, b) == orc.segment_count(osp, a, b)).all()


@pytest.mark.parametrize("name", ["so3", "se3", "chain10"])
def test_space_math_transcendental(gpu, name):
    """acos / sin / cos spaces: at most a last-bit difference from glibc."""
    sp, osp = _spaces(name)
    a, b = _states(name, 1, 3000), _states(name, 2, 3000)
    # the chain metric sums `links` terms built from running cos/sin sums: a few ulp
    tol = 2 if name != "chain10" else 64
    assert (helpers.ulp_diff(sp.distance(a, b), orc.distance(osp, a, b)) <= tol).all()
    got, want = sp.interpolate(a, b, 0.3), orc.interpolate(osp, a, b, 0.3)
    assert np.allclose(got, want, rtol=0, atol=2e-15)
    assert (sp.validSegmentCount(a, b) == orc.segment_count(osp, a, b)).all()


def test_so2_wraparound_known_answers(gpu):
    sp = ob.SO2StateSpace()
    assert abs(sp.distance([[-PI + 0.1]], [[PI - 0.1]])[0] - 0.2) < 1e-3
    assert abs(sp.interpolate([[PI - 0.08]], [[-PI + 0.1]], 0.5)[0, 0] - (-PI + 0.01)) < 1e-3
    assert abs(sp.interpolate([[PI - 0.1]], [[0.1]], 0.5)[0, 0] - PI / 2) < 1e-3


# ---------------------------------------------------------------- state validity
def test_hypercube_states(gpu):
    for d in (2, 8, 25):
        chk, ochk = ob.HypercubeValidityChecker(d), orc.checker_hypercube(d)
        s = np.vstack([helpers.uniform_points(d, 20000, d), helpers.hypercube_valid_states(d, 2000, d)])
        # boundary values of both strict comparisons
        s[:50, 0] = 0.9
        s[50:100, -1] = 0.1
        v = chk.isValidBatch(s)
        assert (v == orc.check_states(ochk, s)).all() and v[-2000:].all() and not v.all()


def test_circles_states_and_fixture(gpu):
    z = helpers.load_circles()
    chk = ob.Circles2DValidityChecker(z["circles"])
    q = z["queries"]
    assert (chk.isValidBatch(q[:, :2]) == z["start_valid"]).all()
    assert (chk.isValidBatch(q[:, 2:]) == z["goal_valid"]).all()
    lo, hi = z["low"], z["high"]
    s = helpers.rng(1).uniform(lo, hi, (50000, 2))
    assert (chk.isValidBatch(s) == orc.check_states(orc.checker_circles(z["circles"]), s, nthreads=8)).all()


@pytest.mark.parametrize("links", [10, 100])
def test_chain_states(gpu, links):
    L = 1.0 / links
    env = orc.horn_environment(links, np.log(links) / links)
    assert len(env) == 2 * (links - 1)
    chk, ochk = ob.KinematicChainValidityChecker(links, L, env), orc.checker_chain(links, L, env)
    n = 6000 if links == 10 else 600
    r = helpers.rng(links)
    s = np.vstack([r.uniform(-PI, PI, (n, links)), r.normal(0, 0.15, (n, links))])  # the second half is mostly valid
    v = chk.isValidBatch(s)
    ov = orc.check_states(ochk, s, nthreads=8)
    assert (v == ov).all() and 0 < v.sum() < len(v)


def test_se3_states(gpu):
    sph, box = helpers.se3_obstacles()
    robot = helpers.se3_robot()
    chk, ochk = ob.SE3ObstacleValidityChecker(robot, sph, box), orc.checker_se3(robot, sph, box)
    s = helpers.se3_states(3, 40000)
    v = chk.isValidBatch(s)
    assert (v == orc.check_states(ochk, s, nthreads=8)).all() and 0 < v.sum() < len(v)
    # only spheres / only boxes / no obstacles
    for sp_, bx_ in ((sph, box[:0]), (sph[:0], box), (sph[:0], box[:0])):
        c2, o2 = ob.SE3ObstacleValidityChecker(robot, sp_, bx_), orc.checker_se3(robot, sp_, bx_)
        assert (c2.isValidBatch(s[:5000]) == orc.check_states(o2, s[:5000], nthreads=8)).all()


def _grazing_states(robot, sph, box, seed, n):
    """States whose robot spheres sit within a few ulps .. 1e-4 of an obstacle surface:
    the FP32 filter must hand every one of them to the exact test."""
    r = helpers.rng(seed)
    q = helpers.random_quaternions(seed + 1, n)
    x, y, z, w = q.T
    out = np.empty((n, 7))
    for i in range(n):
        k = r.integers(0, len(robot))
        o, rr = robot[k, :3], robot[k, 3]
        qv = np.array([x[i], y[i], z[i]])
        t = 2.0 * np.cross(qv, o)
        rot = o + w[i] * t + np.cross(qv, t)
        d = r.normal(size=3)
        d /= np.linalg.norm(d)
        # down to a few ulps of the FP64 expressions, and across the FP32 narrow phase's band (about 4e-5 A wide)
        eps = r.choice([0.0, 1e-15, -1e-15, 1e-12, -1e-12, 1e-9, -1e-9, 1e-8, -1e-8, 1e-7, -1e-7, 1e-6, -1e-6, 1e-5,
                        -1e-5, 4e-5, -4e-5, 1e-4, -1e-4])
        if r.random() < 0.5 and len(sph):
            j = r.integers(0, len(sph))
            c = sph[j, :3] + d * (sph[j, 3] + rr + eps)
        else:
            j = r.integers(0, len(box))
            face = r.integers(0, 3)
            c = box[j, :3] + r.uniform(-1, 1, 3) * box[j, 3:]
            c[face] = box[j, face] + r.choice([-1, 1]) * (box[j, 3 + face] + rr + eps)
        out[i, :3] = c - rot
        out[i, 3:] = q[i]
    return out


def test_se3_filter_grazing_contacts(gpu):
    sph, box = helpers.se3_obstacles()
    robot = helpers.se3_robot()
    chk, ochk = ob.SE3ObstacleValidityChecker(robot, sph, box), orc.checker_se3(robot, sph, box)
    s = _grazing_states(robot, sph, box, 11, 20000)
    v = chk.isValidBatch(s)
    assert (v == orc.check_states(ochk, s, nthreads=8)).all()
    assert 0.1 < v.mean() < 0.9  # the sample really straddles the surfaces


@pytest.mark.parametrize("offset,scale", [(1000.0, 1.0), (0.0, 1e-3), (0.0, 250.0), (-3e4, 10.0)])
def test_se3_filter_far_from_origin_and_scaled(gpu, offset, scale):
    """FP32 resolution shrinks with |coordinate|: the margin must follow (more survivors, same booleans)."""
    sph, box = helpers.se3_obstacles(nsph=200, nbox=200)
    robot = helpers.se3_robot()
    sph, box, robot = sph * scale, box * scale, robot * scale
    sph[:, :3] += offset
    box[:, :3] += offset
    chk, ochk = ob.SE3ObstacleValidityChecker(robot, sph, box), orc.checker_se3(robot, sph, box)
    s = helpers.se3_states(5, 20000)
    s[:, :3] = s[:, :3] * scale + offset
    g = _grazing_states(robot, sph, box, 12, 5000)
    s = np.vstack([s, g])
    v = chk.isValidBatch(s)
    assert (v == orc.check_states(ochk, s, nthreads=8)).all() and 0 < v.sum() < len(v)


def test_se3_filter_non_unit_quaternions_and_clutter(gpu):
    robot = helpers.se3_robot()
    # 300 obstacles piled on one spot: the per-lane survivor list (12 entries) overflows
    r = helpers.rng(21)
    sph = np.hstack([r.normal(0, 0.01, (150, 3)) + 0.3, r.uniform(0.02, 0.05, (150, 1))])
    box = np.hstack([r.normal(0, 0.01, (150, 3)) + 0.3, r.uniform(0.02, 0.05, (150, 3))])
    chk, ochk = ob.SE3ObstacleValidityChecker(robot, sph, box), orc.checker_se3(robot, sph, box)
    s = helpers.se3_states(6, 20000)
    s[:, :3] = 0.3 + helpers.rng(22).normal(0, 0.08, (20000, 3))  # around the pile, in and out
    s[::3, 3:] *= helpers.rng(23).uniform(0.2, 3.0, (len(s[::3]), 1))  # |q| != 1: the filter steps aside
    s[5, 3:] = 0.0
    v = chk.isValidBatch(s)
    assert (v == orc.check_states(ochk, s, nthreads=8)).all() and 0 < v.sum() < len(v)
    # the same world through the motion validator (short segments around the pile)
    sp, osp = _spaces("se3")
    a = s[1::3][:3000].copy()
    b = a.copy()
    b[:, :3] += helpers.rng(24).normal(0, 0.05, (len(a), 3))
    mv = ob.BatchedDiscreteMotionValidator(sp, chk)
    gv, gt = mv.checkMotionBatch(a, b, last_valid=True)[:2]
    ov, ot = orc.check_motions(osp, ochk, a, b, last_valid=True, nthreads=8)
    assert (gv == ov).all() and (gt[~ov] == ot[~ov]).all() and 0 < ov.sum() < len(ov)


@pytest.mark.parametrize("step", [0.0005, 0.003, 0.01, 0.05])
def test_motion_lane_group_sizes(gpu, step):
    """Mean states per segment ~1.5 .. 130: every lane-group width (4, 8, 16, 32) gives the same answers."""
    sp, osp = ob.RealVectorStateSpace(2), orc.space_rv(2)
    circ = np.array([[0.5, 0.5, 0.2], [0.2, 0.8, 0.1], [0.8, 0.2, 0.15]])
    chk, ochk = ob.Circles2DValidityChecker(circ), orc.checker_circles(circ)
    r = helpers.rng(31)
    a = r.uniform(0, 1, (6000, 2))
    b = a + r.normal(0, step * 10, a.shape)
    b[::7] = r.uniform(0, 1, (len(b[::7]), 2))  # a few long ones in every batch
    _motion_parity(sp, osp, chk, ochk, a, b)
    assert chk.stat("last_group") in (4, 8, 16, 32)


def test_allvalid_and_single_state(gpu):
    chk = ob.AllValidStateValidityChecker()
    assert chk.isValidBatch(np.zeros((5, 3))).all()
    assert ob.HypercubeValidityChecker(3).isValid([0.95, 0.95, 0.5])
    assert not ob.HypercubeValidityChecker(3).isValid([0.5, 0.95, 0.95])


# ---------------------------------------------------------------- motion validity
def _motion_parity(sp, osp, chk, ochk, a, b, last_state=True):
    mv = ob.BatchedDiscreteMotionValidator(sp, chk)
    v = mv.checkMotionBatch(a, b)
    ov, ot = orc.check_motions(osp, ochk, a, b, last_valid=True, nthreads=8)
    assert (v == ov).all(), f"{(v != ov).sum()} of {len(v)} motion booleans differ"
    v2, t, st = mv.checkMotionBatch(a, b, last_valid=True, last_state=last_state)
    assert (v2 == ov).all()
    assert (t[~ov] == ot[~ov]).all()
    assert mv.getValidMotionCount() == 2 * int(ov.sum()) and mv.getCheckedMotionCount() == 2 * len(a)
    return v, t, st


def test_motion_hypercube_resolution_0001(gpu):
    """HypercubeBenchmark: d = 8, resolution 0.001 (demos/HypercubeBenchmark.cpp:93-97)."""
    d = 8
    sp, osp = ob.RealVectorStateSpace(d), orc.space_rv(d, 0.0, 1.0, 0.001)
    sp.setLongestValidSegmentFraction(0.001)
    chk, ochk = ob.HypercubeValidityChecker(d), orc.checker_hypercube(d)
    a = helpers.hypercube_valid_states(1, 4000, d)
    r = helpers.rng(2)
    b = np.clip(a + r.normal(0, 0.03, a.shape), 0, 1)
    b[::3] = helpers.hypercube_valid_states(3, len(b[::3]), d)
    v, t, st = _motion_parity(sp, osp, chk, ochk, a, b)
    assert 0 < v.sum() < len(v)
    # lastValid.first is the interpolated state at lastValid.second
    bad = np.where(~v)[0][:200]
    want = orc.interpolate(osp, a[bad], b[bad], t[bad])
    assert (st[bad] == want).all()


def test_motion_circle_world_fixture(gpu):
    z = helpers.load_circles()
    sp = ob.RealVectorStateSpace(2, z["low"], z["high"])
    mv = ob.BatchedDiscreteMotionValidator(sp, ob.Circles2DValidityChecker(z["circles"]))
    q = z["queries"]
    v, t, _ = mv.checkMotionBatch(q[:, :2], q[:, 2:], last_valid=True, last_state=True)
    assert (v == z["motion_valid"]).all() and (t[~v] == z["last_t"][~v]).all()
    assert (sp.validSegmentCount(q[:, :2], q[:, 2:]) == z["nd"]).all()
    # single-segment overloads, as SpaceInformation::checkMotion calls them
    for i in range(0, 100, 9):
        assert mv.checkMotion(q[i, :2], q[i, 2:]) == bool(z["motion_valid"][i])
        lv = [np.zeros(2), -1.0]
        ok = mv.checkMotion(q[i, :2], q[i, 2:], lv)
        assert ok == bool(z["motion_valid"][i])
        if not ok:
            assert lv[1] == z["last_t"][i]


def test_motion_se2_circles(gpu):
    sp, osp = _spaces("se2")
    circ = np.array([[0.5, 0.5, 0.2], [0.2, 0.8, 0.1], [0.8, 0.2, 0.15]])
    a, b = helpers.se2_states(1, 5000), helpers.se2_states(2, 5000)
    _motion_parity(sp, osp, ob.Circles2DValidityChecker(circ), orc.checker_circles(circ), a, b)


def test_motion_se3_obstacles(gpu):
    sp, osp = _spaces("se3")
    sph, box = helpers.se3_obstacles()
    robot = helpers.se3_robot()
    a, b = helpers.se3_states(1, 20000), helpers.se3_states(2, 20000)
    # a share of short segments so that some motions are valid
    b[::2, :3] = a[::2, :3] + helpers.rng(3).normal(0, 0.02, (10000, 3))
    b[::2, 3:] = a[::2, 3:]
    v, _, _ = _motion_parity(sp, osp, ob.SE3ObstacleValidityChecker(robot, sph, box),
                             orc.checker_se3(robot, sph, box), a, b, last_state=False)
    assert 0 < v.sum() < len(v)


def test_motion_chain(gpu):
    links = 10
    sp, osp = _spaces("chain10")
    env = orc.horn_environment(links, np.log(links) / links)
    chk, ochk = ob.KinematicChainValidityChecker(links, 0.1, env), orc.checker_chain(links, 0.1, env)
    r = helpers.rng(5)
    a = r.uniform(-PI, PI, (3000, links))
    b = a + r.normal(0, 0.03, a.shape)
    v, _, _ = _motion_parity(sp, osp, chk, ochk, a, b, last_state=False)
    assert 0 < v.sum() < len(v)


def test_motion_degenerate_segments(gpu):
    """nd = 0 (identical states) and nd = 1 check only s2 (DiscreteMotionValidator.cpp:103-109)."""
    sp, osp = ob.RealVectorStateSpace(2), orc.space_rv(2)
    circ = np.array([[0.5, 0.5, 0.2]])
    chk, ochk = ob.Circles2DValidityChecker(circ), orc.checker_circles(circ)
    a = np.array([[0.1, 0.1], [0.5, 0.5], [0.1, 0.1], [0.5, 0.5], [0.1, 0.1]])
    b = np.array([[0.1, 0.1], [0.5, 0.5], [0.1, 0.105], [0.5, 0.505], [0.9, 0.9]])
    v, t, _ = _motion_parity(sp, osp, chk, ochk, a, b)
    assert v.tolist() == [True, False, True, False, False]


def test_motion_count_factor_and_empty(gpu):
    sp, osp = ob.RealVectorStateSpace(3), orc.space_rv(3)
    sp.setValidSegmentCountFactor(3)
    osp.comp[0].factor = 3
    a, b = helpers.uniform_points(1, 500, 3), helpers.uniform_points(2, 500, 3)
    assert (sp.validSegmentCount(a, b) == orc.segment_count(osp, a, b)).all()
    chk = ob.HypercubeValidityChecker(3)
    _motion_parity(sp, osp, chk, orc.checker_hypercube(3), a, b)
    mv = ob.BatchedDiscreteMotionValidator(sp, chk)
    assert len(mv.checkMotionBatch(np.zeros((0, 3)), np.zeros((0, 3)))) == 0


# ---------------------------------------------------------------- vs the reference's own compiled code
def test_gpu_matches_reference_spaces_dmv_fixture(gpu):
    """tests/golden/ref_spaces_dmv.npz: outputs of the reference's compiled StateSpace / spaces /
    DiscreteMotionValidator (scripts/gen_golden.py).  RealVector / SE2 results are bit-exact;
    SO3 / SE3 go through acos / sin and may differ from glibc in the last bit."""
    import os

    z = np.load(os.path.join(helpers.GOLDEN, "ref_spaces_dmv.npz"))
    se2 = ob.SE2StateSpace()
    a, b = z["se2_a"], z["se2_b"]
    assert (se2.distance(a, b) == z["se2_dist"]).all()
    assert (se2.interpolate(a, b, 0.37) == z["se2_interp"]).all()
    assert (se2.validSegmentCount(a, b) == z["se2_nd"]).all()
    mv = ob.BatchedDiscreteMotionValidator(se2, ob.Circles2DValidityChecker(z["se2_circ"]))
    v, t, _ = mv.checkMotionBatch(a, b, last_valid=True)
    assert (v == z["se2_valid"]).all() and (t[~v] == z["se2_last_t"][~v]).all()

    hc = ob.RealVectorStateSpace(8)
    hc.setLongestValidSegmentFraction(0.001)
    a, b = z["hc_a"], z["hc_b"]
    assert (hc.distance(a, b) == z["hc_dist"]).all() and (hc.validSegmentCount(a, b) == z["hc_nd"]).all()
    mv = ob.BatchedDiscreteMotionValidator(hc, ob.HypercubeValidityChecker(8))
    v, t, _ = mv.checkMotionBatch(a, b, last_valid=True)
    assert (v == z["hc_valid"]).all() and (t[~v] == z["hc_last_t"][~v]).all()

    se3 = ob.SE3StateSpace(-1.0, 1.0)
    a, b = z["se3_a"], z["se3_b"]
    assert (helpers.ulp_diff(se3.distance(a, b), z["se3_dist"]) <= 2).all()
    assert np.allclose(se3.interpolate(a, b, 0.37), z["se3_interp"], rtol=0, atol=2e-15)
    assert (se3.validSegmentCount(a, b) == z["se3_nd"]).all()
    mv = ob.BatchedDiscreteMotionValidator(se3, ob.SE3ObstacleValidityChecker(z["se3_robot"], z["se3_sph"], z["se3_box"]))
    v, t, _ = mv.checkMotionBatch(a, b, last_valid=True)
    assert (v == z["se3_valid"]).all() and (t[~v] == z["se3_last_t"][~v]).all()


@pytest.mark.skipif(not orc.have_refbase(), reason="oracle/_ref/libompl_refbase.so did not travel")
def test_gpu_vs_live_reference_motion_validator(gpu):
    """Larger live differential run against the reference's compiled DiscreteMotionValidator."""
    rb = orc.RefBase("se2", 0, 0.0, 1.0)
    circ = np.array([[0.5, 0.5, 0.2], [0.2, 0.8, 0.1], [0.8, 0.2, 0.15], [0.3, 0.3, 0.05]])
    a, b = helpers.se2_states(41, 20000), helpers.se2_states(42, 20000)
    rv, rt, rst = rb.check_motions(orc.checker_circles(circ), a, b, last_valid=True, last_state=True)
    mv = ob.BatchedDiscreteMotionValidator(ob.SE2StateSpace(), ob.Circles2DValidityChecker(circ))
    v, t, st = mv.checkMotionBatch(a, b, last_valid=True, last_state=True)
    assert (v == rv).all() and (t[~v] == rt[~v]).all() and (st[~v] == rst[~v]).all()
