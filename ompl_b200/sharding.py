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
