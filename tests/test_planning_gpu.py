"""Batch-submit call sites (ompl_b200/planning.py, include/ompl_b200/BatchPlanning.h) against the
sequential loops of the reference planners, run with the oracle:
  FMT::sampleFree / saveNeighborhood     geometric/planners/fmt/src/FMT.cpp:169-240
  PRM::addMilestone                      geometric/planners/prm/src/PRM.cpp:562-596
  SpaceInformation::checkMotion(states)  base/src/SpaceInformation.cpp:274-285, getMotionStates :201-272
  PathGeometric::check                   geometric/src/PathGeometric.cpp:159-180
Hypercube world of demos/HypercubeBenchmark.cpp (d = 8, resolution 0.001)."""
import numpy as np
import pytest

import helpers
import ompl_b200 as ob
from ompl_b200 import planning
from oracle import oracle_py as orc

pytestmark = pytest.mark.gpu
D = 8


def _world():
    sp = ob.RealVectorStateSpace(D)
    sp.setLongestValidSegmentFraction(0.001)
    osp = orc.space_rv(D, 0.0, 1.0, 0.001)
    return sp, osp, ob.HypercubeValidityChecker(D), orc.checker_hypercube(D)


def _stream(seed):
    r = helpers.rng(seed)

    def draw(n):
        # corridor-biased proposals so that a useful share is valid
        base = helpers.hypercube_valid_states(int(r.integers(1 << 30)), n, D)
        return np.clip(base + r.normal(0, 0.04, base.shape), 0, 1)

    return draw


def test_sample_free_batch_equals_sequential_loop(gpu):
    sp, osp, chk, ochk = _world()
    rows, attempts = planning.sample_free_batch(chk, _stream(3), 3000, chunk=1024)
    # the sequential loop on the same stream (chunks of the stream concatenated)
    draw = _stream(3)
    seq, tries = [], 0
    while len(seq) < 3000:
        cand = draw(1024)
        ok = orc.check_states(ochk, cand, nthreads=8)
        for c, v in zip(cand, ok):
            if len(seq) == 3000:
                break
            tries += 1
            if v:
                seq.append(c)
    assert attempts == tries and attempts > 3000
    assert (rows == np.array(seq)).all()


def test_fmt_neighbourhoods_batch(gpu):
    sp, osp, chk, ochk = _world()
    rows, _ = planning.sample_free_batch(chk, _stream(4), 4000, chunk=4096)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(rows)
    nbh = planning.save_neighborhoods_batch(nn, rows, k=36)
    oi, _ = orc.knn(osp, rows, rows[:300], 36, nthreads=8)
    for i in range(300):
        assert (nbh[i] == oi[i, 1:]).all()  # FMT.cpp:181-183 drops the node itself
    nbr = planning.save_neighborhoods_batch(nn, rows[:64], r=0.12)
    for i in range(64):
        want, _ = orc.radius(osp, rows, rows[i], 0.12)
        assert (nbr[i] == want[1:]).all()


def test_prm_connect_batch_equals_sequential_semantics(gpu):
    sp, osp, chk, ochk = _world()
    rows, _ = planning.sample_free_batch(chk, _stream(5), 3000, chunk=4096)
    roadmap, batch = rows[:2000], rows[2000:]
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(roadmap)
    mv = ob.BatchedDiscreteMotionValidator(sp, chk)
    k = 10
    frm, to, valid = planning.connect_milestones_batch(nn, mv, batch, k, state_rows=lambda ids: rows[ids.astype(np.int64)])
    assert nn.size() == len(rows)
    # sequential: add the batch, then for every m: nearestK(m, k+1) minus m, checkMotion(n, m)
    oi, _ = orc.knn(osp, rows, batch, k + 1, nthreads=8)
    wf, wt = [], []
    for i, row in enumerate(oi):
        own = 2000 + i
        row = row[row != own][:k]
        wf.extend(row.tolist())
        wt.extend([own] * len(row))
    assert frm.tolist() == wf and to.tolist() == wt
    ov, _ = orc.check_motions(osp, ochk, rows[np.array(wf)], rows[np.array(wt)], last_valid=True, nthreads=8)
    assert (valid == ov).all() and 0 < valid.sum() < len(valid)
    assert mv.getCheckedMotionCount() == len(wf)


def test_state_sequences_and_paths(gpu):
    sp, osp, chk, ochk = _world()
    good = helpers.hypercube_valid_states(6, 50, D)
    seq = good.copy()
    seq[23] = 0.5  # the centre of the cube is in collision
    assert planning.check_states_batch(chk, good) == (True, None)
    assert planning.check_states_batch(chk, seq) == (False, 23)
    # getMotionStates: count interior states at j / (count + 1)
    a, b = good[0], good[1]
    ms = planning.get_motion_states(sp, a, b, 9, endpoints=True)
    assert ms.shape == (11, D) and (ms[0] == a).all() and (ms[-1] == b).all()
    t = np.arange(1, 10) / 10.0
    want = orc.interpolate(osp, np.repeat(a[None], 9, 0), np.repeat(b[None], 9, 0), t)
    assert (ms[1:-1] == want).all()
    assert planning.get_motion_states(sp, a, b, 0, endpoints=True).shape == (2, D)
    assert planning.get_motion_states(sp, a, b, 3, endpoints=False).shape == (3, D)
    # PathGeometric::check
    mv = ob.BatchedDiscreteMotionValidator(sp, chk)
    edge = np.linspace(0.0, 1.0, 40)[:, None] * np.eye(D)[D - 1][None, :]  # along the last axis: valid corridor
    ov, _ = orc.check_motions(osp, ochk, edge[:-1], edge[1:], last_valid=True)
    assert planning.check_path_batch(chk, mv, edge) == bool(ov.all() and orc.check_states(ochk, edge[:1])[0])
    bad = np.vstack([edge[:5], np.full((1, D), 0.5), edge[5:]])
    assert not planning.check_path_batch(chk, mv, bad)
    assert planning.check_path_batch(chk, mv, edge[:0]) and planning.check_path_batch(chk, mv, edge[:1])


def test_bitstar_sample_batches_neighbourhoods_and_bulk_prune(gpu):
    """BIT* call sites: ImplicitGraph::updateSamples (ImplicitGraph.cpp:965-996),
    nearestSamples (:303-321) and the removals of pruneSamples, against the sequential loops."""
    sp, osp, chk, ochk = _world()
    nn = ob.NearestNeighborsCUDA(sp)
    have, required, factor = 25, 2500, 100
    rows, ids, tries = planning.update_samples_batch(chk, nn, _stream(9), have, required, factor, chunk=512)
    draw = _stream(9)
    seq, seq_tries, n = [], 0, have
    while seq_tries < factor * required and n < required:
        cand = draw(512)
        ok = orc.check_states(ochk, cand, nthreads=8)
        for c, v in zip(cand, ok):
            if not (seq_tries < factor * required and n < required):
                break
            seq_tries += 1
            if v:
                seq.append(c)
                n += 1
    assert tries == seq_tries and len(rows) == required - have and (rows == np.array(seq)).all()
    assert nn.size() == len(rows) and (ids == np.arange(len(rows), dtype=np.uint64)).all()
    # tries exhausted before enough samples are found: the loop stops at the limit
    nn2 = ob.NearestNeighborsCUDA(sp)
    r2, _, t2 = planning.update_samples_batch(ob.HypercubeValidityChecker(D), nn2, lambda m: np.full((m, D), 0.5),
                                              0, 10, 3, chunk=7)
    assert t2 == 30 and len(r2) == 0 and nn2.size() == 0

    verts = rows[:200]
    nbk = planning.nearest_samples_batch(nn, verts, k=12)
    oi, _ = orc.knn(osp, rows, verts, 12, nthreads=8)
    assert all((nbk[i] == oi[i]).all() for i in range(len(verts)))
    nbr = planning.nearest_samples_batch(nn, verts[:40], r=0.1)
    for i in range(40):
        want, _ = orc.radius(osp, rows, verts[i], 0.1)
        assert (nbr[i] == want).all()

    pruned = ids[::3]
    assert planning.prune_samples_batch(nn, pruned) == len(pruned)
    assert planning.prune_samples_batch(nn, pruned) == 0  # already gone
    live = np.setdiff1d(ids, pruned)
    assert nn.size() == len(live)
    nbk = planning.nearest_samples_batch(nn, verts, k=12)
    oi, _ = orc.knn(osp, rows[live.astype(np.int64)], verts, 12, nthreads=8)
    assert all((nbk[i] == live[oi[i].astype(np.int64)]).all() for i in range(len(verts)))
    # prune past the compaction threshold, then add again: ids keep increasing
    more = live[: len(live) * 3 // 4]
    assert planning.prune_samples_batch(nn, more) == len(more)
    live = np.setdiff1d(live, more)
    first = nn.add_rows(rows[:50])
    assert first == len(rows)
    allrows = np.vstack([rows, rows[:50]])
    live = np.concatenate([live, first + np.arange(50, dtype=np.uint64)])
    nbk = planning.nearest_samples_batch(nn, verts[:64], k=5)
    oi, _ = orc.knn(osp, allrows[live.astype(np.int64)], verts[:64], 5, nthreads=8)
    assert all((nbk[i] == live[oi[i].astype(np.int64)]).all() for i in range(64))
