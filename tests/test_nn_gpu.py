"""GPU parity tests for NearestNeighborsCUDA (through the C ABI) against the
oracle (NearestNeighborsLinear semantics, ties by insertion index) and the
fixtures produced by the reference's real NearestNeighborsLinear.

Bar: neighbour ids AND exact FP64 distances bit-identical.
Mirrors tests/datastructures/nearestneighbors.cpp:117-287.
"""
import os

import numpy as np
import pytest

import helpers
import ompl_b200 as ob
from oracle import oracle_py as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["fp32-scan", "tensor-core-scan"])
def scan_kernel(request, monkeypatch):
    """Every test of this module runs twice: once with the filter scan on the FP32 pipe
    (knn_scan.cuh / knn_scan_hd.cuh) and once on the tensor cores (knn_scan_tc.cu, forced even for
    single queries).  OMPLCU_TC_SCAN is the default of the `tc_scan` option for new handles."""
    monkeypatch.setenv("OMPLCU_TC_SCAN", "0" if request.param == "fp32-scan" else "1")
    return request.param


def _check_knn(nn, osp, pts_live, live_ids, qs, k, nthreads=8):
    ids, d = nn.nearestKBatch(qs, k)
    oi, od = orc.knn(osp, pts_live, qs, k, nthreads=nthreads)
    want = np.asarray(live_ids, dtype=np.uint64)[oi.astype(np.int64)]
    assert ids.shape == want.shape
    assert (ids == want).all(), f"ids differ in {(ids != want).any(axis=1).sum()} of {len(qs)} queries"
    assert (d == od).all()
    return ids, d


# ---------------------------------------------------------------- fixtures from the real reference
@pytest.mark.parametrize("fx,space", [("knn_linear_rv7.npz", "rv7"), ("knn_linear_se3.npz", "se3")])
def test_golden_reference_linear(gpu, fx, space):
    z = np.load(os.path.join(helpers.GOLDEN, fx))
    sp = ob.RealVectorStateSpace(7) if space == "rv7" else ob.SE3StateSpace(0.0, 1.0)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(z["points"])
    # SE3 distances go through acos: last-bit agreement with glibc is measured, not proven
    tol = 0 if space == "rv7" else 2
    ids, d = nn.nearestKBatch(z["queries"], int(z["k"]))
    assert (ids == z["ids"]).all() and (helpers.ulp_diff(d, z["dists"]) <= tol).all()
    nn.set_option("force_exact", 1)
    ids2, d2 = nn.nearestKBatch(z["queries"], int(z["k"]))
    assert (ids2 == z["ids"]).all() and (helpers.ulp_diff(d2, z["dists"]) <= tol).all()


# ---------------------------------------------------------------- fast scan path
@pytest.mark.parametrize("d", [1, 2, 3, 7, 8, 16])
@pytest.mark.parametrize("k", [1, 10, 14, 36])
def test_scan_path_dims_and_k(gpu, d, k):
    n, nq = 30000, 700
    pts, qs = helpers.uniform_points(d, n, d), helpers.uniform_points(100 + d, nq, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    _check_knn(nn, orc.space_rv(d), pts, np.arange(n), qs, k)
    assert nn.stat("last_path") == 1


@pytest.mark.parametrize("d", [17, 31, 64, 100, 128])
@pytest.mark.parametrize("k", [1, 10, 36])
def test_high_dimensional_scan_path(gpu, d, k):
    """17 <= d <= 128: the register-tiled scan of knn_scan_hd.cuh feeds the same exact re-rank."""
    n, nq = 20000 + d, 333  # neither a multiple of the tile sizes
    pts, qs = helpers.uniform_points(d, n, d), helpers.uniform_points(200 + d, nq, d)
    qs[:40] = pts[:40]  # exact hits
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts[:15000])
    nn.add_rows(pts[15000:])  # two appends: filter rows are extended, not rebuilt
    _check_knn(nn, orc.space_rv(d), pts, np.arange(n), qs, k)
    assert nn.stat("last_path") == 1
    assert nn.stat("last_fallbacks") <= nq // 10


def test_high_dimensional_scan_ties_removal_and_scales(gpu):
    d = 100
    r = helpers.rng(12)
    pts = helpers.with_duplicates(r.random((12000, d)), 3, frac=0.2)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    qs = np.vstack([pts[:200], r.random((200, d))])
    ids, dd = _check_knn(nn, orc.space_rv(d), pts, np.arange(len(pts)), qs, 10)
    ties = dd[:, 1:] == dd[:, :-1]
    assert ties.any() and (ids[:, 1:][ties] > ids[:, :-1][ties]).all()
    # tombstones: removed points never come back
    for i in range(0, 2000, 2):
        assert nn.remove_id(i)
    live = np.setdiff1d(np.arange(len(pts)), np.arange(0, 2000, 2))
    _check_knn(nn, orc.space_rv(d), pts[live], live, qs[:128], 10)
    # far from the origin with a tiny spread: the band check must send what it cannot prove to the exact path
    far = 1e5 + r.random((8000, d)) * 1e-4
    nn2 = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn2.add_rows(far)
    _check_knn(nn2, orc.space_rv(d), far, np.arange(len(far)), far[:64] + 1e-9, 10)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_scan_path_rv7_seeds(gpu, seed):
    n, nq = 200000, 2048
    pts, qs = helpers.uniform_points(seed, n, 7), helpers.uniform_points(seed + 1000, nq, 7)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(7))
    nn.add_rows(pts)
    _check_knn(nn, orc.space_rv(7), pts, np.arange(n), qs, 10)
    assert nn.stat("last_path") == 1
    assert nn.stat("last_fallbacks") <= nq // 50  # the filter proves nearly every query


def test_scan_path_duplicates_force_ties(gpu):
    n = 50000
    pts = helpers.with_duplicates(helpers.uniform_points(5, n, 7), 6, frac=0.3)
    qs = np.vstack([pts[:400], helpers.uniform_points(7, 400, 7)])
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(7))
    nn.add_rows(pts)
    ids, d = _check_knn(nn, orc.space_rv(7), pts, np.arange(n), qs, 10)
    assert (d[:400, 0] == 0).all()
    ties = (d[:, 1:] == d[:, :-1])
    assert ties.any()
    assert (ids[:, 1:][ties] > ids[:, :-1][ties]).all()  # ties ordered by insertion index


def test_scan_path_many_identical_points(gpu):
    """More exact ties than k: membership at the k-th boundary is by insertion index."""
    pts = np.vstack([np.full((500, 3), 0.25), helpers.uniform_points(1, 3000, 3)])
    qs = np.vstack([np.full((1, 3), 0.25), helpers.uniform_points(2, 63, 3)])
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(3))
    nn.add_rows(pts)
    ids, _ = _check_knn(nn, orc.space_rv(3), pts, np.arange(len(pts)), qs, 10)
    assert ids[0].tolist() == list(range(10))


def test_scan_path_adversarial_scales(gpu):
    """Large offsets / tiny spreads stress the FP32 filter bound: results stay exact
    (the band check sends unprovable queries to the exact path)."""
    r = helpers.rng(9)
    cases = [
        1e6 + r.random((20000, 5)) * 1e-3,          # far from the origin, tiny spread
        r.normal(size=(20000, 5)) * 1e-8,           # denormal-ish FP32 squares
        np.round(r.random((20000, 5)) * 8) / 8,     # grid: massive ties
        r.random((20000, 5)) * np.array([1e4, 1, 1e-4, 1, 1]),  # anisotropic
    ]
    for pts in cases:
        nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(5))
        nn.add_rows(pts)
        qs = np.vstack([pts[:100], pts[100:200] + 1e-9])
        _check_knn(nn, orc.space_rv(5), pts, np.arange(len(pts)), qs, 10)


def test_full_size_sampled_against_oracle(gpu):
    """BASELINE config: 1M x 7-D, 64K queries, k=10; 512 sampled queries vs the oracle,
    all queries through size-independent properties."""
    n, nq, k = 1_000_000, 65536, 10
    pts, qs = helpers.uniform_points(1, n, 7), helpers.uniform_points(1001, nq, 7)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(7))
    nn.add_rows(pts)
    ids, d = nn.nearestKBatch(qs, k)
    assert nn.stat("last_path") == 1
    assert ids.shape == (nq, k)
    assert (np.diff(d, axis=1) >= 0).all()                     # sorted
    assert (ids < n).all()
    s = np.sort(ids, axis=1)
    assert (s[:, 1:] != s[:, :-1]).all()                       # no duplicates
    sel = helpers.rng(3).choice(nq, 4096, replace=False)       # returned distances are the exact metric
    diff = pts[ids[sel].astype(np.int64)] - qs[sel][:, None, :]
    acc = np.zeros(diff.shape[:2])
    for c in range(7):
        acc = acc + diff[:, :, c] * diff[:, :, c]
    assert (np.sqrt(acc) == d[sel]).all()
    sub = helpers.rng(4).choice(nq, 512, replace=False)
    oi, od = orc.knn(orc.space_rv(7), pts, qs[sub], k, nthreads=8)
    assert (ids[sub] == oi).all() and (d[sub] == od).all()
    # self-queries: every point is its own nearest neighbour at distance 0
    ids2, d2 = nn.nearestKBatch(pts[:4096], 1)
    assert (ids2[:, 0] == np.arange(4096)).all() and (d2 == 0).all()


# ---------------------------------------------------------------- exact path: every space, any k
@pytest.mark.parametrize("space", ["so2", "so3", "se2", "se3", "chain", "rv40_forced"])
def test_exact_path_spaces(gpu, space):
    n, nq, k = 1500, 200, 12
    if space == "so2":
        sp, osp = ob.SO2StateSpace(), orc.space_so2()
        pts = helpers.rng(1).uniform(-np.pi, np.pi, (n, 1))
    elif space == "so3":
        sp, osp, pts = ob.SO3StateSpace(), orc.space_so3(), helpers.random_quaternions(1, n)
    elif space == "se2":
        sp, osp, pts = ob.SE2StateSpace(), orc.space_se2(), helpers.se2_states(1, n)
    elif space == "se3":
        sp, osp, pts = ob.SE3StateSpace(), orc.space_se3(), helpers.se3_states(1, n, 0, 1)
    elif space == "chain":
        sp, osp = ob.KinematicChainSpace(10, 0.1), orc.space_chain(10, 0.1)
        pts = helpers.rng(1).uniform(-np.pi, np.pi, (n, 10))
    else:
        # every RealVector dimension up to the stride limit has a scan kernel now: force the sort path
        sp, osp, pts = ob.RealVectorStateSpace(40), orc.space_rv(40), helpers.uniform_points(1, n, 40)
    qs = np.vstack([pts[: nq // 2], pts[::-1][: nq // 2] * 0.999])
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    if space == "rv40_forced":
        nn.set_option("force_exact", 1)
    ids, d = nn.nearestKBatch(qs, k)
    oi, od = orc.knn(osp, pts, qs, k, nthreads=8)
    if space in ("so3", "se3", "chain"):
        # transcendental metrics: ids must agree wherever the oracle's ordering is not a
        # last-bit near-tie; distances within 1 ulp ("transcendentals")
        assert np.allclose(d, od, rtol=4e-16 if space != "chain" else 1e-13, atol=1e-15)
        mism = (ids != oi).any(axis=1)
        assert mism.sum() == 0, f"{mism.sum()} queries differ"
    else:
        assert (ids == oi).all() and (d == od).all()
    assert nn.stat("last_path") == 2


def test_exact_collect_path_large_index(gpu):
    """>= 32K slots: the exact path bounds each query from a sample, collects and sorts only what passed."""
    n = 40000
    # SE3 (acos metric), k = 10
    sp, osp, pts = ob.SE3StateSpace(), orc.space_se3(), helpers.se3_states(3, n, 0, 1)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(pts)
    qs = np.vstack([pts[:100], helpers.se3_states(4, 100, 0, 1)])
    ids, d = nn.nearestKBatch(qs, 10)
    oi, od = orc.knn(osp, pts, qs, 10, nthreads=8)
    assert nn.stat("last_path") == 2 and (ids == oi).all() and np.allclose(d, od, rtol=4e-16, atol=1e-15)
    nn.set_option("collect", 0)
    ids2, d2 = nn.nearestKBatch(qs, 10)
    assert (ids2 == ids).all() and (d2 == d).all()  # identical to the full sort, bit for bit
    # RealVector(8), FMT*-sized k, 30% duplicates (ties across the bound), tombstones
    pts = helpers.with_duplicates(helpers.uniform_points(5, n, 8), 6, frac=0.3)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(8))
    nn.add_rows(pts)
    for i in range(0, 3000, 3):
        nn.remove_id(i)
    live = np.setdiff1d(np.arange(n), np.arange(0, 3000, 3))
    qs = np.vstack([pts[3000:3064], helpers.uniform_points(7, 64, 8)])
    ids, d = nn.nearestKBatch(qs, 2100)
    oi, od = orc.knn(orc.space_rv(8), pts[live], qs, 2100, nthreads=8)
    assert nn.stat("last_path") == 2
    assert (ids == live.astype(np.uint64)[oi.astype(np.int64)]).all() and (d == od).all()
    ties = d[:, 1:] == d[:, :-1]
    assert ties.any() and (ids[:, 1:][ties] > ids[:, :-1][ties]).all()


def test_exact_collect_redo_when_the_sample_misleads(gpu):
    """A dense cluster the strided sample never sees: the bound catches far more than `cap` slots for queries
    inside it, and those queries must come back through the full sort with the same answers."""
    n = 65536
    r = helpers.rng(17)
    pts = np.column_stack([r.random((n, 2)) * 100.0, r.uniform(-3.0, 3.0, n)])  # SE2: x, y, theta
    step = n // 1024
    hidden = np.array([i for i in range(n) if i % step != 0][:30000])
    # 30000 states in a 1e-3 box, none of them on the sample stride
    pts[hidden] = np.column_stack([50.0 + r.random((len(hidden), 2)) * 1e-3, r.random(len(hidden)) * 1e-3])
    nn = ob.NearestNeighborsCUDA(ob.SE2StateSpace(0.0, 100.0))
    nn.add_rows(pts)
    qs = np.vstack([pts[hidden[:32]], np.column_stack([r.random((32, 2)) * 100.0, r.uniform(-3.0, 3.0, 32)])])
    ids, dd = nn.nearestKBatch(qs, 50)
    oi, od = orc.knn(orc.space_se2(0.0, 100.0), pts, qs, 50, nthreads=8)
    assert nn.stat("last_path") == 2 and nn.stat("last_collect_redo") >= 1
    assert (ids == oi).all() and (dd == od).all()


def test_large_k_uses_exact_path(gpu):
    n = 20000
    pts = helpers.uniform_points(2, n, 8)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(8))
    nn.add_rows(pts)
    qs = helpers.uniform_points(3, 50, 8)
    for k in (41, 300, 2100):
        ids, d = nn.nearestKBatch(qs, k)
        oi, od = orc.knn(orc.space_rv(8), pts, qs, k, nthreads=8)
        assert (ids == oi).all() and (d == od).all()


# ---------------------------------------------------------------- reference unit-test mirror
def test_state_space_test_mirror(gpu):
    """nearestneighbors.cpp:117-206 on SE3 [0,1]^3, n = 200, k = 10."""
    n, k = 200, 10
    st = helpers.se3_states(5, n, 0.0, 1.0)
    sp, osp = ob.SE3StateSpace(0.0, 1.0), orc.space_se3(0.0, 1.0)
    elems = [("state", i) for i in range(n)]
    nn = ob.NearestNeighborsCUDA(sp, state_accessor=lambda e: st[e[1]])
    assert nn.reportsSortedResults()
    nn.add(elems)
    assert nn.size() == n and len(nn.list()) == n
    for i in range(0, n, 7):
        nbh = nn.nearestK(elems[i], k)
        assert len(nbh) == k and nbh[0] == elems[i]
        oi, _ = orc.knn(osp, st, st[i:i + 1], k)
        assert [e[1] for e in nbh] == oi[0].tolist()
        assert nn.nearest(elems[i]) == elems[i]
    allr = nn.nearestR(elems[0], float("inf"))
    assert len(allr) == n and allr[0] == elems[0]
    dd = orc.distance(osp, st[[e[1] for e in allr]], np.repeat(st[0:1], n, 0))
    assert (np.diff(dd) >= 0).all()
    assert len(nn.nearestK(elems[0], 2 * n)) == n
    for i in range(n):
        assert nn.remove(elems[i])
        assert nn.size() == n - i - 1 and len(nn.list()) == n - i - 1
    assert not nn.remove(elems[0])
    with pytest.raises(ob.Exception_, match="No elements found"):
        nn.nearest(elems[0])
    assert nn.nearestK(elems[0], 3) == [] and nn.nearestR(elems[0], 1.0) == []


def test_random_access_pattern_mirror(gpu):
    """nearestneighbors.cpp:208-287: interleaved add / nearestK / nearestR / remove."""
    r = helpers.rng(77)
    d = 7
    sp, osp = ob.RealVectorStateSpace(d), orc.space_rv(d)
    nn = ob.NearestNeighborsCUDA(sp)
    live_ids, live_pts = [], []
    for step in range(400):
        p = r.random(d)
        i = nn.add_rows(p.reshape(1, -1))
        live_ids.append(i)
        live_pts.append(p)
        pts = np.array(live_pts)
        q = r.random((1, d))
        k = int(r.integers(1, 31))
        ids, dd = nn.nearestKBatch(q, k)
        oi, od = orc.knn(osp, pts, q, k)
        assert (ids[0] == np.array(live_ids, dtype=np.uint64)[oi[0].astype(int)]).all() and (dd == od).all()
        rad = float(r.uniform(0, 3)) if step % 3 else float(r.uniform(0, 0.5))
        rids, rd, offs = nn.nearestRBatch(q, [rad])
        ri, rdd = orc.radius(osp, pts, q[0], rad)
        assert (rids == np.array(live_ids, dtype=np.uint64)[ri.astype(int)]).all() and (rd == rdd).all()
        if r.random() < 0.5 and live_ids:
            j = int(r.integers(0, len(live_ids)))
            assert nn.remove_id(live_ids[j])
            live_ids.pop(j)
            live_pts.pop(j)
        assert nn.size() == len(live_ids)
        assert nn.list_ids().tolist() == live_ids


def test_remove_tombstones_and_compaction(gpu):
    n = 6000
    pts = helpers.uniform_points(8, n, 7)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(7))
    nn.add_rows(pts)
    r = helpers.rng(1)
    dead = r.choice(n, 4500, replace=False)  # > half: triggers compaction
    for i in dead:
        assert nn.remove_id(int(i))
    assert not nn.remove_id(int(dead[0]))
    live = np.setdiff1d(np.arange(n), dead)
    assert nn.size() == len(live) and (nn.list_ids() == live).all()
    assert nn.stat("slots") < n
    qs = helpers.uniform_points(9, 300, 7)
    _check_knn(nn, orc.space_rv(7), pts[live], live, qs, 10)
    # ids keep growing after compaction and after clear()
    first = nn.add_rows(helpers.uniform_points(10, 5, 7))
    assert first == n
    nn.clear()
    assert nn.size() == 0
    first = nn.add_rows(pts[:100])
    assert first == n + 5
    _check_knn(nn, orc.space_rv(7), pts[:100], np.arange(100) + first, qs[:20], 10)


def test_k_larger_than_size_and_empty(gpu):
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(3))
    ids, d = nn.nearestKBatch(np.zeros((4, 3)), 5)
    assert ids.shape == (4, 0)
    pts = helpers.uniform_points(1, 7, 3)
    nn.add_rows(pts)
    ids, d = nn.nearestKBatch(pts, 50)
    oi, od = orc.knn(orc.space_rv(3), pts, pts, 50)
    assert ids.shape == (7, 7) and (ids == oi).all() and (d == od).all()
    ids, d, offs = nn.nearestRBatch(np.zeros((0, 3)), [])
    assert len(ids) == 0 and offs.tolist() == [0]


def test_radius_batch(gpu):
    n = 20000
    pts = helpers.uniform_points(4, n, 7)
    qs = helpers.uniform_points(5, 40, 7)
    radii = helpers.rng(6).uniform(0.0, 0.45, 40)
    radii[0] = 0.0
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(7))
    nn.add_rows(pts)
    ids, d, offs = nn.nearestRBatch(qs, radii)
    osp = orc.space_rv(7)
    for i in range(len(qs)):
        oi, od = orc.radius(osp, pts, qs[i], radii[i])
        assert (ids[offs[i]:offs[i + 1]] == oi).all() and (d[offs[i]:offs[i + 1]] == od).all()
    # inclusive bound: radius equal to an exact neighbour distance keeps it
    ids10, d10 = nn.nearestKBatch(qs[:1], 10)
    ids_r, d_r, _ = nn.nearestRBatch(qs[:1], [d10[0, -1]])
    assert (ids_r == ids10[0]).all()


def test_merge_topk_equals_single_index(gpu):
    """Sharded roadmap: per-shard top-k merged by (distance, id) == one index."""
    import ctypes as C

    import torch

    from ompl_b200._lib import check, lib

    n, nq, k, parts = 40000, 1000, 10, 4
    pts = helpers.with_duplicates(helpers.uniform_points(3, n, 7), 4, 0.1)
    qs = np.vstack([pts[:500], helpers.uniform_points(5, nq - 500, 7)])
    sp = ob.RealVectorStateSpace(7)
    whole = ob.NearestNeighborsCUDA(sp)
    whole.add_rows(pts)
    wi, wd = whole.nearestKBatch(qs, k)
    dq = torch.from_numpy(qs).cuda()
    pd = torch.empty((parts, nq, k), dtype=torch.float64, device="cuda")
    pi = torch.empty((parts, nq, k), dtype=torch.int64, device="cuda")
    shards = []
    bounds = np.linspace(0, n, parts + 1).astype(int)
    for p in range(parts):
        s = ob.NearestNeighborsCUDA(sp)
        s.set_id_base(int(bounds[p]))
        s.add_rows(pts[bounds[p]:bounds[p + 1]])
        shards.append(s)
        check(lib().omplcu_nn_knn_device(s.handle, dq.data_ptr(), nq, k, pi[p].data_ptr(), pd[p].data_ptr(), None))
    od = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    oi = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    check(lib().omplcu_merge_topk_device(0, pd.data_ptr(), pi.data_ptr(), parts, nq, k, od.data_ptr(), oi.data_ptr(),
                                         None))
    assert (oi.cpu().numpy().astype(np.uint64) == wi).all() and (od.cpu().numpy() == wd).all()
    # the packed layout bench.py all-gathers: [part][ids | dists][nq][k], merged in place with a stride
    packed = torch.empty((parts, 2, nq, k), dtype=torch.int64, device="cuda")
    packed[:, 0] = pi
    packed[:, 1] = pd.view(torch.int64)
    od.zero_()
    oi.zero_()
    check(lib().omplcu_merge_topk_strided_device(0, packed[0, 1].data_ptr(), packed[0, 0].data_ptr(), parts,
                                                 2 * nq * k, nq, k, od.data_ptr(), oi.data_ptr(), None))
    assert (oi.cpu().numpy().astype(np.uint64) == wi).all() and (od.cpu().numpy() == wd).all()


def test_merge_topk_more_candidates_than_shared_memory_holds(gpu):
    """nparts * k > 3200: merge_sorted_kernel (binary searches in the sorted lists) == one index, duplicates included."""
    import torch

    from ompl_b200._lib import check, lib

    n, nq, k, parts = 12000, 300, 900, 5
    pts = helpers.with_duplicates(helpers.uniform_points(13, n, 3), 5, 0.2)
    qs = np.vstack([pts[:100], helpers.uniform_points(15, nq - 100, 3)])
    sp = ob.RealVectorStateSpace(3)
    whole = ob.NearestNeighborsCUDA(sp)
    whole.add_rows(pts)
    wi, wd = whole.nearestKBatch(qs, k)
    dq = torch.from_numpy(qs).cuda()
    pd = torch.empty((parts, nq, k), dtype=torch.float64, device="cuda")
    pi = torch.empty((parts, nq, k), dtype=torch.int64, device="cuda")
    # uneven shards; the last one holds fewer than k states, so its lists end in absent entries
    bounds = np.array([0, 4000, 7000, 9500, 11500, n])
    shards = []
    for p in range(parts):
        s = ob.NearestNeighborsCUDA(sp)
        s.set_id_base(int(bounds[p]))
        s.add_rows(pts[bounds[p]:bounds[p + 1]])
        shards.append(s)
        check(lib().omplcu_nn_knn_device(s.handle, dq.data_ptr(), nq, k, pi[p].data_ptr(), pd[p].data_ptr(), None))
    od = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    oi = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    check(lib().omplcu_merge_topk_device(0, pd.data_ptr(), pi.data_ptr(), parts, nq, k, od.data_ptr(), oi.data_ptr(),
                                         None))
    assert (oi.cpu().numpy().astype(np.uint64) == wi).all() and (od.cpu().numpy() == wd).all()


# ---------------------------------------------------------------- tensor-core scan (knn_scan_tc.cu)
def _tc_err_units(d):
    k16 = (3 * d + 3 + 15) // 16  # tcgen05.mma instructions (K = 16) per pair of tiles
    return 1.02 * (3.0 + 768.0 + 32.0 * k16)


@pytest.mark.parametrize("d", [1, 7, 9, 10, 20, 21, 33, 52, 53, 64, 100, 128])
def test_tensor_core_scan_values_within_the_stated_error_bound(gpu, scan_kernel, d):
    """Raw scan values (tcgen05.mma on BF16 split operands) against an FP64 evaluation of
    s = |pt|^2 - 2 qt.pt on the FP32-rounded operands: the band check of the re-rank assumes
    |s_tc - s| <= tc_err_units(d) * 2^-24 * (pmax2 + 2 |qt| pmax) (knn_scan_tc.cuh).  The measured
    error must stay below half of that bound; every operand-image shape is covered
    (one or several chunks, odd and even MMA counts, 1 / 2 / 4 resident query tiles)."""
    if scan_kernel != "tensor-core-scan":
        pytest.skip("tensor-core scan only")
    import torch

    n, nq, k = 1500, 300, 10
    r = np.random.default_rng(d)
    pts, qs = r.uniform(-np.pi, np.pi, (n, d)), r.uniform(-np.pi, np.pi, (nq, d))
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    units, qpad = (n + 127) // 128, (nq + 511) // 512 * 512
    buf = torch.full((qpad, units * 128), float("nan"), dtype=torch.float32, device="cuda")
    nn.set_option("tc_scan_values_out", buf.data_ptr())
    ids, dist = nn.nearestKBatch(qs, k)
    nn.set_option("tc_scan_values_out", 0)
    torch.cuda.synchronize()
    assert nn.stat("last_tc") == 1
    got = buf.cpu().numpy().astype(np.float64)[:nq, :n]
    c = pts.mean(axis=0)
    pt = (pts - c).astype(np.float32).astype(np.float64)
    qt = (qs - c).astype(np.float32).astype(np.float64)
    pn = (pt * pt).sum(1).astype(np.float32).astype(np.float64)  # the norm as stored
    want = pn[None, :] - 2.0 * qt @ pt.T
    pmax2, qn = pn.max(), np.sqrt((qt * qt).sum(1))
    bound = _tc_err_units(d) * 2.0**-24 * (pmax2 + 2.0 * qn * np.sqrt(pmax2))
    err = np.abs(got - want)
    assert np.isfinite(got).all()
    assert (err <= 0.5 * bound[:, None]).all(), f"max error {np.max(err / bound[:, None]):.3f} of the bound"
    oi, od = orc.knn(orc.space_rv(d), pts, qs, k, nthreads=8)
    assert (ids == oi).all() and (dist == od).all()


def test_tensor_core_scan_is_the_default(gpu, monkeypatch):
    monkeypatch.delenv("OMPLCU_TC_SCAN", raising=False)
    d, n = 7, 20000
    pts = helpers.uniform_points(1, n, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    _check_knn(nn, orc.space_rv(d), pts, np.arange(n), helpers.uniform_points(2, 200, d), 10)
    assert nn.stat("last_path") == 1 and nn.stat("last_tc") == 1
    _check_knn(nn, orc.space_rv(d), pts, np.arange(n), helpers.uniform_points(3, 1, d), 10)
    assert nn.stat("last_path") == 1 and nn.stat("last_few") == 1  # up to 8 queries: the few-query scan (tests/test_nn_few_gpu.py)
    nn.set_option("few_scan", 0)
    _check_knn(nn, orc.space_rv(d), pts, np.arange(n), helpers.uniform_points(3, 1, d), 10)
    assert nn.stat("last_path") == 1 and nn.stat("last_tc") == 1  # without it: the tensor-core scan, even for one query
    nn.set_option("tc_scan", 0)
    _check_knn(nn, orc.space_rv(d), pts, np.arange(n), helpers.uniform_points(2, 200, d), 10)
    assert nn.stat("last_tc") == 0


@pytest.mark.parametrize("d", [7, 40, 100])
def test_tensor_core_images_follow_add_and_remove(gpu, scan_kernel, d):
    """The BF16 operand image is rebuilt lazily after every mutation: interleave add / remove /
    query and compare each answer with the oracle on the live set."""
    r = np.random.default_rng(7 + d)
    pts = r.uniform(-1.0, 1.0, (6000, d))
    qs = r.uniform(-1.0, 1.0, (130, d))
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts[:2500])
    live = list(range(2500))
    _check_knn(nn, orc.space_rv(d), pts[live], live, qs, 10)
    for i in r.choice(2500, 400, replace=False):
        assert nn.remove_id(int(i))
        live.remove(int(i))
    _check_knn(nn, orc.space_rv(d), pts[live], live, qs, 10)
    nn.add_rows(pts[2500:])
    live += list(range(2500, 6000))
    _check_knn(nn, orc.space_rv(d), pts[live], live, qs, 14)
    for i in r.choice(live, 3500, replace=False):  # crosses the compaction threshold
        assert nn.remove_id(int(i))
        live.remove(int(i))
    _check_knn(nn, orc.space_rv(d), pts[live], live, qs, 36)
    assert nn.stat("last_path") == 1


def test_sharded_radius_merge_equals_single_index(gpu):
    """SURVEY 8(e), radius row: two shards with global id bases, nearestRBatch on each, merged by
    (distance, id) (sharding.merge_radius_parts, the function all_gather_radius applies after the two
    exchanges) == the single index holding everything, including duplicates that straddle the shards."""
    from ompl_b200 import sharding

    d, n, nq = 7, 6000, 96
    pts = helpers.with_duplicates(helpers.uniform_points(11, n, d), 12, frac=0.05)
    qs = np.vstack([helpers.uniform_points(13, nq - 6, d), pts[:6]])
    radii = np.where(np.arange(nq) % 4 == 0, 0.0, 0.5)
    whole = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    whole.add_rows(pts)
    w_i, w_d, w_o = whole.nearestRBatch(qs, radii)
    parts = []
    for r in range(2):
        lo, hi = sharding.shard_range(n, r, 2)
        nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
        nn.set_id_base(lo)
        nn.add_rows(pts[lo:hi])
        i, x, o = nn.nearestRBatch(qs, radii)
        parts.append((i, x, np.diff(o)))
    m_i, m_d, m_o = sharding.merge_radius_parts(parts)
    assert (m_o == w_o).all() and (m_i == w_i).all() and (m_d == w_d).all() and m_o[-1] > 1000
