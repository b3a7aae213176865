"""A roadmap sharded over several GPUs of one process (omplcu_nn_create_sharded)
answers exactly like one index holding all the states: ids and FP64 distances `==`.

On a one-GPU box the shards all live on device 0 (a device may be named more than once), which
exercises everything but the NVLink hop; with more GPUs visible the test spreads them out.
"""
import ctypes as C

import numpy as np
import pytest

import helpers
import ompl_b200 as ob
from ompl_b200._lib import lib
from oracle import oracle_py as orc

pytestmark = pytest.mark.gpu


def _devices(n):
    c = C.c_int(0)
    lib().omplcu_device_count(C.byref(c))
    return [i % max(c.value, 1) for i in range(n)]


@pytest.mark.parametrize("nshards", [2, 3])
def test_sharded_knn_equals_single_index_and_the_oracle(gpu, nshards):
    d, n, nq, k = 7, 60000, 1500, 10
    pts = helpers.with_duplicates(helpers.uniform_points(3, n, d), 4, frac=0.05)
    qs = np.vstack([pts[:300], helpers.uniform_points(4, nq - 300, d)])
    one = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    many = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d), devices=_devices(nshards))
    assert many.stat("shards") == nshards
    for nn in (one, many):
        nn.add_rows(pts[:50000])   # bulk: cut into contiguous blocks
        nn.add_rows(pts[50000:50100])  # small: least-loaded shard
        nn.add_rows(pts[50100:])
    assert one.size() == many.size() == n
    i1, d1 = one.nearestKBatch(qs, k)
    i2, d2 = many.nearestKBatch(qs, k)
    assert (i1 == i2).all() and (d1 == d2).all()
    oi, od = orc.knn(orc.space_rv(d), pts, qs[:200], k, nthreads=8)
    assert (i2[:200] == oi.astype(np.uint64)).all() and (d2[:200] == od).all()
    ties = d2[:, 1:] == d2[:, :-1]
    assert ties.any() and (i2[:, 1:][ties] > i2[:, :-1][ties]).all()  # ties by insertion index across shards
    # removal (single and bulk), list, get
    gone = list(range(0, 30000, 3))
    for nn in (one, many):
        assert nn.remove_id(5)
        assert not nn.remove_id(5)
        assert nn.remove_ids(gone) == len(gone)
    assert (one.list_ids() == many.list_ids()).all()
    assert (one.get_row(31) == many.get_row(31)).all() and (many.get_row(31) == pts[31]).all()
    i1, d1 = one.nearestKBatch(qs, k)
    i2, d2 = many.nearestKBatch(qs, k)
    assert (i1 == i2).all() and (d1 == d2).all()
    assert not np.isin(i2, np.array(gone, dtype=np.uint64)).any()
    # k larger than what some shard holds; k larger than everything
    small = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d), devices=_devices(nshards))
    small.add_rows(pts[:5])
    ref = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    ref.add_rows(pts[:5])
    a, b = small.nearestKBatch(qs[:10], 8), ref.nearestKBatch(qs[:10], 8)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
    many.clear()
    assert many.size() == 0
    first = many.add_rows(pts[:1000])
    assert first == n  # ids keep increasing across clear()


def test_sharded_radius_and_large_k(gpu):
    d, n = 3, 40000
    pts = helpers.with_duplicates(helpers.uniform_points(8, n, d), 3, frac=0.1)
    qs = helpers.uniform_points(9, 64, d)
    one = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    many = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d), devices=_devices(4))
    one.add_rows(pts)
    many.add_rows(pts)
    r = np.full(len(qs), 0.08)
    r[::7] = 0.0
    a, b = one.nearestRBatch(qs, r), many.nearestRBatch(qs, r)
    assert (a[2] == b[2]).all() and a[2][-1] > 1000
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
    # nparts * k beyond the device merge's shared-memory budget (FMT*'s k ~ 2100): ranks by binary search in the
    # sorted per-shard lists; duplicates make runs of equal distances that cross shards
    k = 2100
    i1, d1 = one.nearestKBatch(qs[:16], k)
    i2, d2 = many.nearestKBatch(qs[:16], k)
    assert (i1 == i2).all() and (d1 == d2).all()
    # the same with a gather buffer that only holds 5 queries' lists at a time
    many.set_option("gather_entries", 5 * 4 * k)
    i3, d3 = many.nearestKBatch(qs[:16], k)
    assert (i1 == i3).all() and (d1 == d3).all()
    many.set_option("gather_entries", 1 << 27)
    # more neighbours asked for than one shard holds: absent entries at the end of every list
    k = 12000
    i1, d1 = one.nearestKBatch(qs[:3], k)
    i2, d2 = many.nearestKBatch(qs[:3], k)
    assert (i1 == i2).all() and (d1 == d2).all()


def test_sharded_se3_metric(gpu):
    z = np.load(helpers.GOLDEN + "/knn_linear_se3.npz")
    nn = ob.NearestNeighborsCUDA(ob.SE3StateSpace(0.0, 1.0), devices=_devices(2))
    nn.add_rows(z["points"])
    ids, d = nn.nearestKBatch(z["queries"], int(z["k"]))
    assert (ids == z["ids"]).all() and (helpers.ulp_diff(d, z["dists"]) <= 2).all()


def test_motion_batches_split_over_checkers(gpu):
    r = helpers.rng(21)
    n = 30001
    q = r.normal(size=(2 * n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    s = np.hstack([r.uniform(-1, 1, size=(2 * n, 3)), q])
    sph = np.hstack([r.uniform(-1, 1, size=(40, 3)), r.uniform(0.02, 0.08, size=(40, 1))])
    box = np.hstack([r.uniform(-1, 1, size=(40, 3)), r.uniform(0.02, 0.08, size=(40, 3))])
    robot = np.array([[0.0, 0.0, 0.0, 0.03], [0.05, 0.0, 0.0, 0.02]])
    se3 = ob.SE3StateSpace(-1.0, 1.0)
    devs = _devices(3)
    chks = [ob.SE3ObstacleValidityChecker(robot, sph, box, device=dv) for dv in devs]
    one = ob.BatchedDiscreteMotionValidator(se3, chks[0])
    many = ob.BatchedDiscreteMotionValidator(se3, None, checkers=chks)
    a, b = s[:n].copy(), s[n:].copy()
    b[: n // 2] = a[: n // 2] + 0.01 * r.normal(size=(n // 2, 7))  # short edges: many valid
    v1, t1, s1 = one.checkMotionBatch(a, b, last_valid=True, last_state=True)
    v2, t2, s2 = many.checkMotionBatch(a, b, last_valid=True, last_state=True)
    assert 0 < v1.sum() < n
    assert (v1 == v2).all() and (t1 == t2).all() and (s1 == s2).all()
    assert many.getCheckedMotionCount() == n


@pytest.mark.parametrize("nshards,nq", [(2, 4096), (3, 5001), (8, 9000)])
def test_sharded_large_host_batches_merge_per_query_block(gpu, nshards, nq):
    """Host-pointer batches of 4096+ queries: every shard fetches the queries itself, answers them, merges ONE block of
    the queries from all shards' lists and copies that block out (ShardedNN::knn_host_distributed).  Ragged block
    sizes, duplicates (ties across shards), k larger than a shard's share."""
    d, n = 7, 40000
    pts = helpers.with_duplicates(helpers.uniform_points(5, n, d), 6, frac=0.05)
    qs = np.vstack([pts[:500], helpers.uniform_points(6, nq - 500, d)])
    one = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    many = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d), devices=_devices(nshards))
    for nn in (one, many):
        nn.add_rows(pts)
    for k in (1, 10):
        i1, d1 = one.nearestKBatch(qs, k)
        i2, d2 = many.nearestKBatch(qs, k)
        assert (i1 == i2).all() and (d1 == d2).all()
    tiny = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d), devices=_devices(nshards))
    tiny.add_rows(pts[:7])
    ref = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    ref.add_rows(pts[:7])
    a, b = tiny.nearestKBatch(qs, 12), ref.nearestKBatch(qs, 12)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
