"""kNN / nearestR on SO(2), SO(3), SE(2), SE(3) roadmaps large enough for the filtered exact path (>= 32768
states): the tensor-core threshold scan runs on a Euclidean lower-bound embedding of the states (nn.cu,
embed_state), the kept slots get the space's own exact distance, and the answers must stay those of
NearestNeighborsLinear: same ids (ties by insertion index), distances as in the unfiltered path.  The adversarial
cases aim at the places where the embedding's bound is tight or its assumptions end."""
import numpy as np
import pytest

import helpers
import ompl_b200 as ob
from oracle import oracle_py as orc

pytestmark = pytest.mark.gpu


def _space(name):
    if name == "so2":
        return ob.SO2StateSpace(), orc.space_so2()
    if name == "so3":
        return ob.SO3StateSpace(), orc.space_so3()
    if name == "se2":
        return ob.SE2StateSpace(), orc.space_se2()
    return ob.SE3StateSpace(), orc.space_se3()


def _states(name, seed, n):
    if name == "so2":
        return helpers.rng(seed).uniform(-np.pi, np.pi, (n, 1))
    if name == "so3":
        return helpers.random_quaternions(seed, n)
    if name == "se2":
        return helpers.se2_states(seed, n)
    return helpers.se3_states(seed, n, 0, 1)


def _same(nn, osp, pts, qs, k, exact_dist):
    ids, d = nn.nearestKBatch(qs, k)
    oi, od = orc.knn(osp, pts, qs, k, nthreads=16)
    assert (ids == oi).all(), f"{(ids != oi).any(axis=1).sum()} of {len(qs)} queries differ"
    if exact_dist:
        assert (d == od).all()
    else:
        assert np.allclose(d, od, rtol=4e-16, atol=1e-15)  # acos: DESIGN.md 5
    return ids, d


@pytest.mark.parametrize("name", ["so2", "so3", "se2", "se3"])
@pytest.mark.parametrize("k", [1, 10, 60])
def test_filtered_knn_on_rotation_and_compound_spaces(gpu, name, k):
    n, nq = 50000, 257
    sp, osp = _space(name)
    pts = _states(name, 11, n)
    qs = np.vstack([_states(name, 12, nq - 32), pts[:32]])  # some queries ARE elements
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    _same(nn, osp, pts, qs, k, name in ("so2", "se2"))
    assert nn.stat("last_path") == 2 and nn.stat("last_filter") == 1
    if k <= 40:
        # one query tile cut into 31 segments: the list scan on the embedding proves (nearly) every query by itself
        assert nn.stat("last_proven") > 0
    # the unfiltered path gives the same bits
    ids, d = nn.nearestKBatch(qs, k)
    nn.set_option("tc_scan", 0)
    ids0, d0 = nn.nearestKBatch(qs, k)
    assert nn.stat("last_filter") == 0
    assert (ids == ids0).all() and (d == d0).all()


def test_filtered_knn_se3_adversarial_rotations(gpu):
    """Rotations where the bound is tight or the reference's distance collapses: pairs closer than the 1e-9
    cut-off of SO3StateSpace::arcLength (distance 0 although the quaternions differ), antipodal quaternions
    (q and -q are the same rotation), rotations by exactly pi / 2 from the query (|dot| ~ 0, the largest distance),
    and exact duplicates (ties by insertion index)."""
    n = 40000
    sp, osp = _space("se3")
    pts = _states("se3", 21, n)
    r = helpers.rng(22)
    base = pts[:2000].copy()
    # 1. near-copies within 3e-5 rad and 1e-6 in position
    near = base.copy()
    dq = r.normal(size=(2000, 4)) * 1.5e-5
    near[:, 3:] = base[:, 3:] + dq
    near[:, 3:] /= np.linalg.norm(near[:, 3:], axis=1, keepdims=True)
    near[:, :3] += r.normal(size=(2000, 3)) * 1e-6
    pts[2000:4000] = near
    # 2. antipodal copies
    anti = base.copy()
    anti[:, 3:] = -anti[:, 3:]
    pts[4000:6000] = anti
    # 3. exact duplicates
    pts[6000:8000] = base
    qs = np.vstack([base[:200], near[:100], anti[:100], _states("se3", 23, 100)])
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    ids, d = _same(nn, osp, pts, qs, 12, False)
    assert nn.stat("last_filter") == 1
    assert (d[:200, :3] == 0.0).all() or (d[:200, 0] == 0.0).all()


def test_filtered_knn_so2_wraparound(gpu):
    n = 40000
    sp, osp = _space("so2")
    pts = _states("so2", 31, n)
    pts[:500, 0] = np.pi
    pts[500:1000, 0] = -np.pi
    pts[1000:1500, 0] = np.nextafter(np.pi, 0)
    qs = np.array([[np.pi], [-np.pi], [3.1415], [-3.1415], [0.0], [1e-300]])
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    _same(nn, osp, pts, np.vstack([qs, _states("so2", 32, 200)]), 700, True)
    assert nn.stat("last_filter") == 1


def test_states_outside_the_embeddings_assumptions_switch_the_filter_off(gpu):
    """A stored quaternion that is not unit to 1e-9 (the reference does not normalise either: its distance is
    whatever acos of the raw dot product gives) switches the filter off for the structure; a query like that is
    redone unfiltered.  Either way the answers equal the oracle's."""
    n = 40000
    sp, osp = _space("se3")
    pts = _states("se3", 41, n)
    qs = _states("se3", 42, 64)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    bad_q = qs.copy()
    bad_q[::4, 3:] *= 1.0 + 1e-6
    _same(nn, osp, pts, bad_q, 10, False)
    assert nn.stat("last_filter") == 1  # the good queries of the batch were filtered
    bad = pts[:1].copy()
    bad[0, 3:] *= 1.0 - 1e-5
    nn.add_rows(bad)
    allp = np.vstack([pts, bad])
    _same(nn, osp, allp, qs, 10, False)
    assert nn.stat("last_filter") == 0
    assert nn.remove_id(n)  # the offender goes; compaction rebuilds the filter copy
    for i in range(0, n, 2):
        nn.remove_id(i)
    live = list(range(1, n, 2))
    ids, d = nn.nearestKBatch(qs, 10)
    oi, od = orc.knn(osp, pts[live], qs, 10, nthreads=16)
    assert (ids == np.asarray(live, dtype=np.uint64)[oi.astype(np.int64)]).all()


@pytest.mark.parametrize("name", ["se2", "se3"])
def test_filtered_radius_on_compound_spaces(gpu, name):
    n = 50000
    sp, osp = _space(name)
    pts = _states(name, 51, n)
    qs = np.vstack([_states(name, 52, 40), pts[:8]])
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    r = 0.12 if name == "se2" else 0.35
    gi, gd, offs = nn.nearestRBatch(qs, r)
    assert nn.stat("last_filter") == 1
    total = 0
    for i, q in enumerate(qs):
        oi, od = orc.radius(osp, pts, q, r)
        assert (gi[offs[i]:offs[i + 1]] == oi).all()
        assert np.allclose(gd[offs[i]:offs[i + 1]], od, rtol=4e-16, atol=1e-15)
        total += len(oi)
    assert total > 100


def test_filtered_knn_large_batch_needs_the_threshold_scan(gpu):
    """16384 queries on 200K SE(3) states: few segments per query tile, so the lists alone prove little and the
    threshold scan on the embedding decides; same answers as the unfiltered path and (sampled) the oracle."""
    n, nq = 200000, 16384
    sp, osp = _space("se3")
    pts, qs = _states("se3", 81, n), _states("se3", 82, nq)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    ids, d = nn.nearestKBatch(qs, 10)
    assert nn.stat("last_filter") == 1 and nn.stat("last_proven") < nq
    sub = helpers.rng(83).choice(nq, 128, replace=False)
    oi, od = orc.knn(osp, pts, qs[sub], 10, nthreads=16)
    assert (ids[sub] == oi).all() and np.allclose(d[sub], od, rtol=4e-16, atol=1e-15)
    assert (np.diff(d, axis=1) >= 0).all()


@pytest.mark.parametrize("name", ["se2", "se3"])
def test_filtered_knn_on_a_small_roadmap(gpu, name):
    """From 4096 states on the embedded path needs no sample (its bound comes from the list scan), so small SE(2) / SE(3)
    roadmaps are filtered too; removals included."""
    n, nq = 6000, 300
    sp, osp = _space(name)
    pts, qs = _states(name, 91, n), _states(name, 92, nq)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    _same(nn, osp, pts, qs, 10, name == "se2")
    assert nn.stat("last_filter") == 1
    live = [i for i in range(n) if i % 5]
    assert nn.remove_ids([i for i in range(n) if i % 5 == 0]) == n - len(live)
    ids, d = nn.nearestKBatch(qs, 10)
    oi, od = orc.knn(osp, pts[live], qs, 10, nthreads=8)
    assert (ids == np.asarray(live, dtype=np.uint64)[oi.astype(np.int64)]).all()
