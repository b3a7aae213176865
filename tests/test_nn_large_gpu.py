"""The upper range of BASELINE config 3 (4M and 16M states) and the full-size secondary workloads:
M2's FMT* neighbourhood k = 2100 at n = 100K (fmt/src/FMT.cpp:327-337) and chain-metric near-ties.

Sizes the oracle cannot cover whole are checked on a sample against the oracle and, for ALL queries,
through a size-independent property: the tensor-core scan and the FP32-pipe scan are two filters with
different error bounds in front of the same exact FP64 re-rank, so every query -- including the ones that
fall back to the exact path -- must come out bit-identical from both.
"""
import numpy as np
import pytest

import helpers
import ompl_b200 as ob
from oracle import oracle_py as orc

pytestmark = pytest.mark.gpu


def _both_scans(pts, qs, k):
    out = {}
    for tc in (1, 0):
        nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(pts.shape[1]))
        nn.set_option("tc_scan", tc)
        nn.add_rows(pts)
        ids, d = nn.nearestKBatch(qs, k)
        assert nn.stat("last_path") == 1 and nn.stat("last_tc") == tc
        out[tc] = (ids, d, nn.stat("last_fallbacks"))
        del nn
    return out


def test_4m_states_sampled_against_the_oracle_both_scan_kernels(gpu):
    n, nq, k = 4_000_000, 65536, 10
    pts, qs = helpers.uniform_points(1, n, 7), helpers.uniform_points(1001, nq, 7)
    r = _both_scans(pts, qs, k)
    assert (r[0][0] == r[1][0]).all() and (r[0][1] == r[1][1]).all()
    sub = helpers.rng(4).choice(nq, 512, replace=False)
    oi, od = orc.knn(orc.space_rv(7), pts, qs[sub], k, nthreads=16)
    for tc in (0, 1):
        assert (r[tc][0][sub] == oi).all() and (r[tc][1][sub] == od).all()
    assert (np.diff(r[1][1], axis=1) >= 0).all()


def test_16m_states_fallback_queries_stay_exact(gpu):
    """16M states: the tensor-core filter cannot prove a handful of the 65536 queries (its band is wider than the
    gap between the k-th and the list's last value there) and hands them to the exact path; they must still be
    bit-identical to the FP32 scan's answers, and a sample must equal the oracle."""
    n, nq, k = 16_000_000, 65536, 10
    pts, qs = helpers.uniform_points(1, n, 7), helpers.uniform_points(1001, nq, 7)
    r = _both_scans(pts, qs, k)
    assert (r[0][0] == r[1][0]).all() and (r[0][1] == r[1][1]).all()
    assert r[1][2] + r[0][2] >= 1, "expected at least one query on the exact fallback at this size"
    assert r[1][2] <= nq // 100
    sub = helpers.rng(5).choice(nq, 64, replace=False)
    oi, od = orc.knn(orc.space_rv(7), pts, qs[sub], k, nthreads=16)
    assert (r[1][0][sub] == oi).all() and (r[1][1][sub] == od).all()


def test_m2_hypercube_fmt_neighbourhood_k2100_at_100k(gpu):
    """M2 at the config's own size: 100K samples in [0,1]^8, FMT*'s k ~ 2100 nearest per sample."""
    n, k, d = 100_000, 2100, 8
    pts = helpers.uniform_points(21, n, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    sel = helpers.rng(6).choice(n, 2048, replace=False)
    ids, dd = nn.nearestKBatch(pts[sel], k)
    assert ids.shape == (2048, k) and (ids[:, 0] == sel).all() and (dd[:, 0] == 0).all()
    assert (np.diff(dd, axis=1) >= 0).all()
    sub = np.arange(0, 2048, 32)
    oi, od = orc.knn(orc.space_rv(d), pts, pts[sel][sub], k, nthreads=16)
    assert (ids[sub] == oi).all() and (dd[sub] == od).all()


def test_chain_metric_near_ties_keep_a_valid_order(gpu):
    """The chain metric sums d libm calls per distance, so device and glibc distances may differ in the last
    bits: up to 64 ulp measured on distances of the order of the chain's length (tests/test_validity_gpu.py), and
    by the same ABSOLUTE amount on small distances, where cos(theta1) - cos(theta2) cancels.  Near-ties closer
    than that may legitimately swap.  What must hold: (1) the returned distances are sorted, (2) every returned
    distance is within the tolerance of the oracle's distance for the SAME id, (3) where the id lists differ,
    the distances at those positions tie within the tolerance, and any id only one side returned ties with the k-th."""
    links, L, n, k = 10, 0.1, 4000, 12
    tol = 64 * np.spacing(links * L)  # 64 ulp at the scale of the chain's length

    def close(a, b):
        return np.abs(np.asarray(a) - np.asarray(b)) <= np.maximum(tol, 64 * np.spacing(np.abs(np.asarray(b))))

    r = helpers.rng(31)
    base = r.uniform(-np.pi, np.pi, (n // 4, links))
    pts = np.vstack([base, base + r.normal(0, 1e-15, base.shape), base + r.normal(0, 3e-16, base.shape),
                     np.nextafter(base, np.inf)])
    qs = r.uniform(-np.pi, np.pi, (128, links))
    qs[:64] = base[:64] + 1e-3
    nn = ob.NearestNeighborsCUDA(ob.KinematicChainSpace(links, L))
    nn.add_rows(pts)
    ids, dd = nn.nearestKBatch(qs, k)
    osp = orc.space_chain(links, L)
    oi, od = orc.knn(osp, pts, qs, k, nthreads=8)
    assert (np.diff(dd, axis=1) >= 0).all()
    swaps = 0
    for q in range(len(qs)):
        dist_q = orc.distance(osp, pts[ids[q].astype(np.int64)], np.repeat(qs[q:q + 1], k, axis=0))
        assert close(dd[q], dist_q).all()
        if not (ids[q] == oi[q]).all():
            swaps += 1
            pos = np.nonzero(ids[q] != oi[q])[0]
            assert close(dd[q, pos], od[q, pos]).all()
            extra = np.setdiff1d(ids[q], oi[q])
            if len(extra):
                de = orc.distance(osp, pts[extra.astype(np.int64)], np.repeat(qs[q:q + 1], len(extra), axis=0))
                assert close(de, np.full(len(extra), od[q, -1])).all()
    assert swaps > 0, "the data was built to contain ties below the tolerance"
    print(f"chain near-ties: {swaps} of {len(qs)} queries differ from the glibc order, all inside the tolerance")


def test_dense_low_dimensional_set_second_pass_on_the_fp32_scan(gpu, monkeypatch):
    """A dense 2-D roadmap (PRM on R^2 at 1M milestones): the k-th neighbour distance is inside the error band of the
    BF16 split, so the tensor-core filter cannot prove the queries.  In its default ("auto") mode the structure hands
    them to the FP32-pipe scan as ONE batch (band ~100x narrower) instead of redoing each exactly, and after a call
    like that makes the FP32-pipe scan the first filter for this set.  Answers stay bit-identical to the oracle and
    to the FP32-only structure throughout."""
    monkeypatch.delenv("OMPLCU_TC_SCAN", raising=False)
    n, nq, k, d = 1_000_000, 4096, 10, 2
    pts, qs = helpers.uniform_points(11, n, d), helpers.uniform_points(12, nq, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    ids, dist = nn.nearestKBatch(qs, k)
    assert nn.stat("last_path") == 1 and nn.stat("last_tc") == 1
    assert nn.stat("last_rescans") >= nq // 4, "expected unproven queries at this density"
    assert nn.stat("last_fallbacks") <= nn.stat("last_rescans") // 10
    assert nn.stat("tc_demoted") == 1
    ids1, dist1 = nn.nearestKBatch(qs, k)  # now straight to the FP32-pipe scan
    assert nn.stat("last_tc") == 0 and nn.stat("last_rescans") == 0
    assert (ids == ids1).all() and (dist == dist1).all()
    nn.set_option("tc_scan", 0)
    ids0, dist0 = nn.nearestKBatch(qs, k)
    assert (ids == ids0).all() and (dist == dist0).all()
    sub = helpers.rng(6).choice(nq, 256, replace=False)
    oi, od = orc.knn(orc.space_rv(d), pts, qs[sub], k, nthreads=16)
    assert (ids[sub] == oi).all() and (dist[sub] == od).all()


@pytest.mark.parametrize("d,n,nq,opts", [
    (7, 300_000, 8192, ["sort_queries", "tc_align"]),
    (100, 60_000, 1024, ["tc_pair", "tc_ts"]),
    (64, 60_000, 1024, ["tc_pair"]),
])
def test_scan_schedule_options_do_not_change_results(gpu, monkeypatch, d, n, nq, opts):
    """Morton-ordered queries, aligned work pieces, CTA pairs and the TMEM-resident query tile only change how the
    filter scan is scheduled: with each of them switched off the ids and distances must be the same bits, and a
    sample must equal the oracle."""
    monkeypatch.delenv("OMPLCU_TC_SCAN", raising=False)
    pts, qs = helpers.uniform_points(31 + d, n, d), helpers.uniform_points(32 + d, nq, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    ids, dist = nn.nearestKBatch(qs, 10)
    assert nn.stat("last_tc") == 1
    sub = helpers.rng(33).choice(nq, 128, replace=False)
    oi, od = orc.knn(orc.space_rv(d), pts, qs[sub], 10, nthreads=16)
    assert (ids[sub] == oi).all() and (dist[sub] == od).all()
    for o in opts:
        nn.set_option(o, 0)
        i2, d2 = nn.nearestKBatch(qs, 10)
        assert (ids == i2).all() and (dist == d2).all(), o
        nn.set_option(o, 1)
