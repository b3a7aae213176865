"""Sharded roadmaps: the host-side plumbing of DESIGN.md section 7.

Rank r of N owns the states whose insertion indices fall in shard_range(n, r, N)
(NearestNeighborsCUDA.set_id_base(lo) makes its ids global), answers every query
against its shard, and the per-rank (distance, id) lists are exchanged with ONE
all-gather.  The merge itself is a device kernel (omplcu_merge_topk_device); the
functions here are backend-agnostic torch.distributed calls, so the same code runs
over NCCL on GPUs and over gloo in the CPU tests (tests/test_sharding_gloo.py).
Motion / state validity batches are split with the same shard_range and need no
exchange.
"""
import torch
import torch.distributed as dist


def shard_range(n, rank, world):
    """Contiguous block [lo, hi) of insertion indices owned by `rank`; blocks tile [0, n)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return n * rank // world, n * (rank + 1) // world


def grid_shape(world, n_states, min_shard=500_000):
    """(S, R) with S * R = world: the roadmap is cut into S state shards and the query batch into R blocks;
    rank = s * R + r holds shard s and answers block r.  S is the largest divisor of `world` that keeps
    n_states / S >= min_shard, but at least 2 once there are 2 ranks (the roadmap IS sharded): below
    ~500K states per shard the scan's fixed cost per (query, shard) list outweighs the shorter stream,
    so further ranks are better spent on disjoint query blocks.  One all-gather still
    moves every rank's lists; block r is merged from the S ranks s * R + r."""
    best = 1
    for s in range(1, world + 1):
        if world % s == 0 and (n_states // s >= min_shard or s <= 2):
            best = s
    best = min(best, world)
    return best, world // best


def grid_coords(rank, world, n_states, min_shard=500_000):
    """(s, r, S, R) of `rank` in the grid of grid_shape: state shard s of S, query block r of R."""
    S, R = grid_shape(world, n_states, min_shard)
    s, r = divmod(rank, R)
    return s, r, S, R


def block_part_ranks(r, S, R):
    """The ranks whose lists make up query block r (one per state shard), in shard order.  In the
    all-gathered buffer they are R rank-buffers apart, which is the part_stride the merge kernel gets."""
    return [s * R + r for s in range(S)]


def alloc_gather_buffers(world, nq, k, device):
    """[world, nq, k] int64 ids and float64 distances, the layout merge_topk expects (parts-major)."""
    return (torch.empty((world, nq, k), dtype=torch.int64, device=device),
            torch.empty((world, nq, k), dtype=torch.float64, device=device))


def all_gather_lists(ids, dists, g_ids, g_dists, group=None):
    """Every rank contributes its [nq, k] lists; afterwards g_*[r] holds rank r's lists on every rank.
    Short lists (shard smaller than k) must be padded by the caller with id = -1 / distance = +inf,
    which is what omplcu_nn_knn_device writes."""
    if ids.shape != g_ids.shape[1:] or dists.shape != g_dists.shape[1:]:
        raise ValueError("gather buffers do not match the local lists")
    # concatenation along dim 0 is the one layout every backend accepts: [world*nq, k] is the same memory
    dist.all_gather_into_tensor(g_ids.view(-1, g_ids.shape[-1]), ids.contiguous(), group=group)
    dist.all_gather_into_tensor(g_dists.view(-1, g_dists.shape[-1]), dists.contiguous(), group=group)


class PackedLists:
    """ids and distances of one rank in ONE [2, nq, k] int64 buffer (row 0: ids, row 1: the float64
    distances' bits), so that a single all-gather moves both; `gathered` is [world, 2, nq, k].
    omplcu_merge_topk_strided_device reads it in place with part_stride = 2 * nq * k."""

    def __init__(self, world, nq, k, device):
        self.world, self.nq, self.k = world, nq, k
        self.local = torch.empty((2, nq, k), dtype=torch.int64, device=device)
        self.gathered = torch.empty((world, 2, nq, k), dtype=torch.int64, device=device)

    @property
    def ids(self):
        return self.local[0]

    @property
    def dists(self):
        return self.local[1].view(torch.float64)

    @property
    def part_stride(self):
        return 2 * self.nq * self.k

    def gathered_ids(self):
        return self.gathered[:, 0]

    def gathered_dists(self):
        return self.gathered[:, 1].view(torch.float64)

    def all_gather(self, group=None):
        dist.all_gather_into_tensor(self.gathered.view(-1, self.k), self.local.view(-1, self.k), group=group)


# ----------------------------------------------------------------------
# radius queries over a sharded roadmap (all-gather of counts, then of padded lists)
# ----------------------------------------------------------------------
def all_gather_radius(ids, dists, offsets, device="cpu", group=None):
    """Merge per-shard nearestR results into the answer of the whole roadmap, on every rank.

    ids / dists / offsets: this rank's NearestNeighborsCUDA.nearestRBatch output for ALL queries against
    its shard (numpy; ids are global insertion indices via set_id_base, each query's slice sorted by
    (distance, id), offsets = CSR boundaries, length nq + 1).  Two exchanges, as the lists are ragged:
    the per-query counts of every rank, then every rank's lists padded to the longest rank total, ids
    and the distances' bits packed in ONE int64 buffer.  `device` is where the exchanged tensors live
    ("cpu" for gloo, a CUDA device for NCCL).  Returns (ids, dists, offsets) of the merged result; the
    order (distance, id) is total and ids are insertion indices, so it equals the single-index answer
    (NearestNeighborsLinear::nearestR, NearestNeighborsLinear.h:135-142, with ties by insertion index).
    """
    import numpy as np

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    ids = np.ascontiguousarray(ids, dtype=np.uint64)
    dists = np.ascontiguousarray(dists, dtype=np.float64)
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    nq = len(offsets) - 1
    if nq < 0 or len(ids) != offsets[-1] or len(dists) != offsets[-1]:
        raise ValueError("ids / dists / offsets do not describe one CSR result")
    if world == 1:
        return ids, dists, offsets
    counts = torch.from_numpy(np.diff(offsets)).to(device)
    g_counts = torch.empty((world, nq), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(g_counts.view(-1), counts, group=group)
    g_counts = g_counts.cpu().numpy()
    totals = g_counts.sum(axis=1)
    pad = int(totals.max())
    merged_off = np.concatenate([[0], np.cumsum(g_counts.sum(axis=0))]).astype(np.int64)
    if pad == 0:
        return np.empty(0, np.uint64), np.empty(0, np.float64), merged_off
    local = np.zeros((2, pad), dtype=np.int64)
    local[0, : len(ids)] = ids.view(np.int64)
    local[1, : len(dists)] = dists.view(np.int64)
    g = torch.empty((world, 2, pad), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(g.view(world * 2, pad), torch.from_numpy(local).to(device), group=group)
    g = g.cpu().numpy()
    parts = [(g[r, 0, : totals[r]].view(np.uint64), g[r, 1, : totals[r]].view(np.float64), g_counts[r])
             for r in range(world)]
    return merge_radius_parts(parts)


def merge_radius_parts(parts):
    """parts: per shard (ids, dists, counts per query).  Returns the merged (ids, dists, offsets): per query
    the union of the shards' neighbours ordered by (distance, id)."""
    import numpy as np

    nq = len(parts[0][2])
    qidx = np.concatenate([np.repeat(np.arange(nq), np.asarray(c, dtype=np.int64)) for _, _, c in parts])
    all_ids = np.concatenate([np.asarray(i, dtype=np.uint64) for i, _, _ in parts])
    all_d = np.concatenate([np.asarray(d, dtype=np.float64) for _, d, _ in parts])
    order = np.lexsort((all_ids, all_d, qidx))  # last key is the primary one
    total = np.sum([np.asarray(c, dtype=np.int64) for _, _, c in parts], axis=0)
    return all_ids[order], all_d[order], np.concatenate([[0], np.cumsum(total)]).astype(np.int64)
