"""Batch-submit call sites for sampling-based planners -- the Python mirror of
include/ompl_b200/BatchPlanning.h (FMT*, PRM, BIT*).  Elements are roadmap ids
(insertion indices); states are rows in copyToReals order.

Each function is a loop of the reference with its per-element calls replaced by one batch
call; the docstring names the loop (paths relative to /root/reference/src/ompl)."""
import numpy as np


def sample_free_batch(checker, draw, wanted, chunk=65536, max_attempts=None):
    """FMT::sampleFree, geometric/planners/fmt/src/FMT.cpp:212-240.

    draw(n) returns the next n samples of the random stream as rows.  Valid samples are kept in
    drawing order until `wanted` are found.  Returns (rows, attempts) with attempts = draws up to and
    including the last accepted sample, i.e. the reference loop's sampleAttempts."""
    kept, attempts, have = [], 0, 0
    while have < wanted and (max_attempts is None or attempts < max_attempts):
        n = chunk if max_attempts is None else min(chunk, max_attempts - attempts)
        cand = np.ascontiguousarray(draw(n), dtype=np.float64)
        ok = checker.isValidBatch(cand)
        idx = np.flatnonzero(ok)
        if have + len(idx) >= wanted:
            idx = idx[: wanted - have]
            attempts += int(idx[-1]) + 1
            kept.append(cand[idx])
            have = wanted
            break
        attempts += len(cand)
        have += len(idx)
        kept.append(cand[idx])
    rows = np.vstack(kept) if kept else np.zeros((0, 0))
    return rows, attempts


def save_neighborhoods_batch(nn, rows, k=None, r=None):
    """FMT::saveNeighborhood for every node at once, FMT.cpp:169-191: nearestK(m, k) (or nearestR(m, r))
    without its first element, which is m itself.  Returns a list of id arrays."""
    if k:
        ids, _ = nn.nearestKBatch(rows, k)
        return [row[1:] for row in ids]
    ids, _, offs = nn.nearestRBatch(rows, r)
    return [ids[offs[i] + 1:offs[i + 1]] if offs[i + 1] > offs[i] else ids[0:0] for i in range(len(rows))]


def connect_milestones_batch(nn, mv, batch_rows, k, state_rows=None):
    """PRM::addMilestone for a batch, geometric/planners/prm/src/PRM.cpp:562-596, with the batch semantics
    of BatchPlanning.h: the whole batch is added first, every member is then connected to its k nearest
    among roadmap + batch (itself excluded); edge (n, m) is kept iff checkMotion(n, m).

    state_rows(ids) -> rows returns the coordinates of roadmap ids (defaults to a lookup in an array the
    caller keeps: pass a function).  Returns (from_ids, to_ids, valid)."""
    batch_rows = np.ascontiguousarray(batch_rows, dtype=np.float64)
    if len(batch_rows) == 0 or k == 0:
        z = np.zeros(0, dtype=np.uint64)
        return z, z, np.zeros(0, dtype=bool)
    first = nn.add_rows(batch_rows)
    own = first + np.arange(len(batch_rows), dtype=np.uint64)
    ids, _ = nn.nearestKBatch(batch_rows, k + 1)
    frm, to = [], []
    for i, row in enumerate(ids):
        row = row[row != own[i]][:k]
        frm.append(row)
        to.append(np.full(len(row), own[i], dtype=np.uint64))
    frm, to = np.concatenate(frm), np.concatenate(to)
    if state_rows is None:
        raise ValueError("connect_milestones_batch needs state_rows(ids) to fetch neighbour coordinates")
    valid = mv.checkMotionBatch(state_rows(frm), state_rows(to))
    return frm, to, valid


def check_edges_batch(mv, rows_a, rows_b):
    """The lazy edge checks of FMT::expandTreeFromNode (FMT.cpp:559-573) / BITstar::checkEdge
    (informedtrees/src/BITstar.cpp:799-817) for a whole wavefront."""
    return mv.checkMotionBatch(rows_a, rows_b)


def check_states_batch(checker, rows):
    """SpaceInformation::checkMotion(states, count, firstInvalidStateIndex), base/src/SpaceInformation.cpp:274-285.
    Returns (all_valid, first_invalid_index or None)."""
    ok = checker.isValidBatch(rows)
    bad = np.flatnonzero(~ok)
    return (len(bad) == 0), (int(bad[0]) if len(bad) else None)


def get_motion_states(space, s1, s2, count, endpoints=True):
    """SpaceInformation::getMotionStates, base/src/SpaceInformation.cpp:201-272 (alloc = true): `count`
    states strictly between s1 and s2 at t = j / (count + 1), plus the end points if asked."""
    s1 = np.asarray(s1, dtype=np.float64).reshape(1, -1)
    s2 = np.asarray(s2, dtype=np.float64).reshape(1, -1)
    segs = count + 1
    if segs < 2:
        return np.vstack([s1, s2]) if endpoints else np.zeros((0, s1.shape[1]))
    t = np.arange(1, segs, dtype=np.float64) / float(segs)
    mid = space.interpolate(np.repeat(s1, len(t), 0), np.repeat(s2, len(t), 0), t)
    return np.vstack([s1, mid, s2]) if endpoints else mid


def check_path_batch(checker, mv, path_rows):
    """PathGeometric::check, geometric/src/PathGeometric.cpp:159-180."""
    path_rows = np.ascontiguousarray(path_rows, dtype=np.float64)
    if len(path_rows) == 0:
        return True
    if not checker.isValidBatch(path_rows[:1])[0]:
        return False
    if len(path_rows) < 2:
        return True
    return bool(mv.checkMotionBatch(path_rows[:-1], path_rows[1:]).all())


# ----------------------------------------------------------------------
# BIT* (SURVEY 8(f) row 3): sample batches, neighbourhoods of a set of vertices, bulk pruning
# ----------------------------------------------------------------------
def update_samples_batch(checker, nn, draw, num_have, num_required, allowed_failed_factor=100, chunk=16384):
    """The sampling loop of BITstar::ImplicitGraph::updateSamples,
    geometric/planners/informedtrees/bitstar/src/ImplicitGraph.cpp:965-993, followed by addToSamples (:996):

        for (tries = 0; tries < factor * numRequired && numSamples_ < numRequired; ++tries)
            draw a state; ++numStateCollisionChecks_; if isValid: keep it, ++numSamples_

    with the per-state isValid replaced by isValidBatch over chunks of the same random stream; states are
    accepted in drawing order, so the kept set, `tries` and the collision-check count equal the
    sequential loop's.  The kept states go into the free-sample structure with ONE bulk add.
    Returns (rows, ids, tries)."""
    limit = allowed_failed_factor * num_required
    kept, tries, have = [], 0, num_have
    while tries < limit and have < num_required:
        n = min(chunk, limit - tries)
        cand = np.ascontiguousarray(draw(n), dtype=np.float64)
        idx = np.flatnonzero(checker.isValidBatch(cand))
        if have + len(idx) >= num_required:
            idx = idx[: num_required - have]
            tries += int(idx[-1]) + 1
            have = num_required
            kept.append(cand[idx])
            break
        tries += len(cand)
        have += len(idx)
        kept.append(cand[idx])
    rows = np.vstack(kept) if kept else np.zeros((0, nn.space.stride))
    ids = np.zeros(0, dtype=np.uint64)
    if len(rows):
        first = nn.add_rows(rows)
        ids = first + np.arange(len(rows), dtype=np.uint64)
    return rows, ids, tries


def nearest_samples_batch(nn, vertex_rows, k=None, r=None):
    """BITstar::ImplicitGraph::nearestSamples (ImplicitGraph.cpp:303-321) for every vertex about to be
    expanded: samples_->nearestK(vertex, k_) or nearestR(vertex, r_), one batched call.  Returns a list of
    id arrays sorted by (distance, id)."""
    if k:
        ids, _ = nn.nearestKBatch(vertex_rows, k)
        return [row for row in ids]
    ids, _, offs = nn.nearestRBatch(vertex_rows, r)
    return [ids[offs[i]:offs[i + 1]] for i in range(len(vertex_rows))]


def prune_samples_batch(nn, ids):
    """The removals of BITstar::ImplicitGraph::pruneSamples (samples_->remove(sample) per pruned sample,
    ImplicitGraph.cpp, pruneSamples) as one bulk removal: tombstones on the device, compaction when more
    than half of the slots are dead.  Returns the number removed."""
    return nn.remove_ids(ids)
