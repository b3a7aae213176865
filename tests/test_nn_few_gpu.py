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
