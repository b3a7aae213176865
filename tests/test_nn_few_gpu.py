"""The few-query scan (nn.cu, knn_few_kernel): what a literal nearestK / nearest call of a planner runs
(PRM::addMilestone, prm/src/PRM.cpp:575; RRT::solve's nn_->nearest, rrt/src/RRT.cpp:170).  In the default mode
calls of up to 8 queries on a plain RealVector space take it; results must equal NearestNeighborsLinear's
(ids and FP64 distances bit for bit, ties by insertion order), through adds, removes and duplicates."""
import numpy as np
import pytest

import helpers
import ompl_b200 as ob
from oracle import oracle_py as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _auto_mode(monkeypatch):
    for v in ("OMPLCU_TC_SCAN", "OMPLCU_THR_SCAN"):
        monkeypatch.delenv(v, raising=False)


def _check(nn, d, pts_live, live_ids, qs, k):
    ids, dist = nn.nearestKBatch(qs, k)
    oi, od = orc.knn(orc.space_rv(d), pts_live, qs, k, nthreads=8)
    want = np.asarray(live_ids, dtype=np.uint64)[oi.astype(np.int64)]
    assert (ids == want).all() and (dist == od).all()
    return ids, dist


@pytest.mark.parametrize("d", [1, 2, 7, 12, 33, 100, 128])
@pytest.mark.parametrize("nq", [1, 3, 4, 8])
def test_few_queries_match_the_oracle(gpu, d, nq):
    n = 30000
    pts, qs = helpers.uniform_points(100 + d, n, d), helpers.uniform_points(200 + d, nq, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    _check(nn, d, pts, np.arange(n), qs, 10)
    assert nn.stat("last_path") == 1 and nn.stat("last_few") == 1 and nn.stat("last_tc") == 0
    _check(nn, d, pts, np.arange(n), qs, 1)   # nearest()
    _check(nn, d, pts, np.arange(n), qs, 36)  # the longest list of the fast path
    # nine queries are a batch again
    q9 = helpers.uniform_points(300 + d, 9, d)
    _check(nn, d, pts, np.arange(n), q9, 10)
    assert nn.stat("last_few") == 0


@pytest.mark.parametrize("n", [1, 5, 31, 33, 127, 129, 4097])
def test_few_queries_on_small_and_ragged_sets(gpu, n):
    d = 7
    pts, qs = helpers.uniform_points(7, n, d), helpers.uniform_points(8, 2, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    ids, dist = nn.nearestKBatch(qs, 10)
    oi, od = orc.knn(orc.space_rv(d), pts, qs, min(10, n), nthreads=2)
    assert (ids[:, :min(10, n)] == oi).all() and (dist[:, :min(10, n)] == od).all()


def test_few_queries_follow_add_remove_and_duplicates(gpu):
    """Interleaved addMilestone-style use: one add, one nearestK; removed elements never come back, exact
    duplicates are ordered by insertion index."""
    d, r = 7, helpers.rng(9)
    pts = helpers.with_duplicates(helpers.uniform_points(9, 6000, d), 10, frac=0.05)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts[:3000])
    live = list(range(3000))
    for i in range(3000, 3040):
        nn.add_rows(pts[i:i + 1])
        live.append(i)
        _check(nn, d, pts[live], live, pts[i:i + 1] + 1e-9, 10)
        assert nn.stat("last_few") == 1
    for i in r.choice(3000, 1500, replace=False):
        assert nn.remove_id(int(i))
        live.remove(int(i))
    _check(nn, d, pts[live], live, helpers.uniform_points(10, 5, d), 10)
    q = pts[live[17]][None, :]  # a query that IS an element (and maybe one of its duplicates)
    ids, dist = _check(nn, d, pts[live], live, q, 10)
    assert dist[0, 0] == 0.0


def test_few_queries_dense_low_dimensional_set(gpu):
    """1M states in the unit square: the FP32 evaluation's band is narrow enough that a lone query is still
    proven by the filter (the BF16 split's is not, tests/test_nn_large_gpu.py)."""
    d, n = 2, 1_000_000
    pts, qs = helpers.uniform_points(21, n, d), helpers.uniform_points(22, 4, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    _check(nn, d, pts, np.arange(n), qs, 10)
    assert nn.stat("last_few") == 1 and nn.stat("last_fallbacks") <= 1


# ---------------------------------------------------------------- any space: the exact few-query pass
def _space_and_states(name, seed, n):
    if name == "so2":
        return ob.SO2StateSpace(), orc.space_so2(), helpers.rng(seed).uniform(-np.pi, np.pi, (n, 1))
    if name == "so3":
        return ob.SO3StateSpace(), orc.space_so3(), helpers.random_quaternions(seed, n)
    if name == "se2":
        return ob.SE2StateSpace(), orc.space_se2(), helpers.se2_states(seed, n)
    if name == "se3":
        return ob.SE3StateSpace(), orc.space_se3(), helpers.se3_states(seed, n, 0, 1)
    return ob.KinematicChainSpace(10, 0.1), orc.space_chain(10, 0.1), helpers.rng(seed).uniform(-np.pi, np.pi, (n, 10))


@pytest.mark.parametrize("name", ["so2", "so3", "se2", "se3", "chain"])
@pytest.mark.parametrize("nq,k", [(1, 1), (1, 10), (5, 32), (8, 10)])
def test_exact_few_queries_on_every_space(gpu, name, nq, k):
    """nn_->nearest / nearestK of RRT and PRM on rotation, compound and chain spaces (rrt/src/RRT.cpp:170): one pass
    with the space's exact distance, per-warp (distance, slot) lists, no sort of all slots.  Duplicates test the
    insertion-order tie-break across warps."""
    n = 20000
    sp, osp, pts = _space_and_states(name, 60, n)
    pts = helpers.with_duplicates(pts, 61, frac=0.05)
    qs = np.vstack([pts[7:8], _space_and_states(name, 62, 8)[2]])[:nq]
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    ids, d = nn.nearestKBatch(qs, k)
    assert nn.stat("last_path") == 2 and nn.stat("last_few") == 2
    oi, od = orc.knn(osp, pts, qs, k, nthreads=8)
    assert (ids == oi).all()
    if name in ("so2", "se2"):
        assert (d == od).all()
    else:
        assert np.allclose(d, od, rtol=4e-16 if name != "chain" else 1e-13, atol=1e-15)
    # the sort path gives the same bits
    nn.set_option("few_scan", 0)
    ids0, d0 = nn.nearestKBatch(qs, k)
    assert nn.stat("last_few") == 0
    assert (ids == ids0).all() and (d == d0).all()


def test_exact_few_queries_follow_removals_and_small_sets(gpu):
    sp, osp, pts = _space_and_states("se3", 70, 3000)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    live = list(range(3000))
    for i in helpers.rng(71).choice(3000, 1200, replace=False):
        assert nn.remove_id(int(i))
        live.remove(int(i))
    qs = _space_and_states("se3", 72, 3)[2]
    ids, d = nn.nearestKBatch(qs, 20)
    oi, od = orc.knn(osp, pts[live], qs, 20, nthreads=4)
    assert (ids == np.asarray(live, dtype=np.uint64)[oi.astype(np.int64)]).all()
    assert np.allclose(d, od, rtol=4e-16, atol=1e-15)
