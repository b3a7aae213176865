#!/usr/bin/env python
"""Per-phase device times of one sharded kNN step (local scan / all-gather / merge), under torchrun."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ompl_b200 as ob  # noqa: E402
from ompl_b200 import sharding  # noqa: E402
from ompl_b200._lib import check, lib  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
N, Q, K, D = 1_000_000, 65536, 10, 7
r = np.random.default_rng(1)
pts, qs = r.random((N, D)), r.random((Q, D))
lo, hi = sharding.shard_range(N, rank, world)
nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(D, device=local))
nn.set_id_base(lo)
nn.add_rows(pts[lo:hi])
pk = sharding.PackedLists(world, Q, K, dev)
dq = torch.from_numpy(qs).to(dev)
m_i = torch.empty((Q, K), dtype=torch.int64, device=dev)
m_d = torch.empty((Q, K), dtype=torch.float64, device=dev)
L = lib()
s = torch.cuda.current_stream()
sp = C.c_void_p(s.cuda_stream)
acc = np.zeros(4)
for it in range(13):
    dist.barrier()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record(s)
    check(L.omplcu_nn_knn_device(nn.handle, dq.data_ptr(), Q, K, pk.ids.data_ptr(), pk.dists.data_ptr(), sp))
    ev[1].record(s)
    pk.all_gather()
    ev[2].record(s)
    check(L.omplcu_merge_topk_strided_device(local, pk.gathered_dists().data_ptr(), pk.gathered_ids().data_ptr(), world,
                                             pk.part_stride, Q, K, m_d.data_ptr(), m_i.data_ptr(), sp))
    ev[3].record(s)
    ev[3].synchronize()
    if it >= 3:
        acc += [ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3]), ev[0].elapsed_time(ev[3])]
t = torch.tensor(acc / 10, device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"world={world}: local knn {t[0]:.3f} ms  all-gather {t[1]:.3f} ms  merge {t[2]:.3f} ms  step {t[3]:.3f} ms (max over ranks)")
dist.destroy_process_group()
