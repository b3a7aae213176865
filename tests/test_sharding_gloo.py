"""world_size-2 gloo test of the sharded-roadmap plumbing (ompl_b200/sharding.py).

No GPU here, so each rank's local top-k comes from the oracle (tests may use it); what is
under test is the product's host-side logic: shard ranges, global id bases, the all-gather
layout handed to the merge, and that merging per-shard lists by (distance, id) reproduces
the single-index answer bit for bit -- including ties that straddle shards."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from ompl_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_tile_the_index_space():
    for n in (0, 1, 7, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            r = [sharding.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(hi - lo for lo, hi in r) - min(hi - lo for lo, hi in r) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def test_grid_shape_covers_every_rank_once():
    for world in (1, 2, 4, 8):
        for n in (100_000, 1_000_000, 4_000_000, 16_000_000):
            S, R = sharding.grid_shape(world, n)
            assert S * R == world and S >= min(world, 2)
            assert S == world or n // (2 * S) < 500_000 or world % (2 * S)   # no finer cut would keep 500K per shard
            seen = set()
            for rank in range(world):
                s, r = divmod(rank, R)
                seen.add((sharding.shard_range(n, s, S), sharding.shard_range(65536, r, R)))
            assert len(seen) == world
    assert sharding.grid_shape(8, 1_000_000) == (2, 4) and sharding.grid_shape(8, 16_000_000) == (8, 1)


WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np
    import torch
    import torch.distributed as dist
    sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
    import helpers
    from ompl_b200 import sharding
    from oracle import oracle_py as orc

    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n, nq, k, d = 5000, 64, 10, 7
    pts = helpers.with_duplicates(helpers.uniform_points(1, n, d), 5, frac=0.05)   # ties across shards
    qs = np.vstack([helpers.uniform_points(2, nq - 8, d), pts[:8]])                # and exact hits
    lo, hi = sharding.shard_range(n, rank, world)
    sp = orc.space_rv(d)
    ids, dd = orc.knn(sp, pts[lo:hi], qs, k)
    ids = ids.astype(np.int64) + lo                                                # set_id_base(lo)
    g_ids, g_d = sharding.alloc_gather_buffers(world, nq, k, "cpu")
    sharding.all_gather_lists(torch.from_numpy(ids), torch.from_numpy(dd), g_ids, g_d)
    # merge by (distance, id): what merge_topk_kernel does on the device
    cat_i = g_ids.numpy().transpose(1, 0, 2).reshape(nq, world * k)
    cat_d = g_d.numpy().transpose(1, 0, 2).reshape(nq, world * k)
    order = np.lexsort((cat_i, cat_d), axis=1)[:, :k]
    m_i, m_d = np.take_along_axis(cat_i, order, 1), np.take_along_axis(cat_d, order, 1)
    w_i, w_d = orc.knn(sp, pts, qs, k)
    ok = bool((m_i == w_i.astype(np.int64)).all() and (m_d == w_d).all())
    # the packed variant bench.py uses: one buffer, one all-gather
    pk = sharding.PackedLists(world, nq, k, "cpu")
    pk.ids.copy_(torch.from_numpy(ids))
    pk.dists.copy_(torch.from_numpy(dd))
    pk.all_gather()
    ok = ok and bool((pk.gathered_ids() == g_ids).all() and (pk.gathered_dists() == g_d).all())
    ok = ok and pk.part_stride == 2 * nq * k
    # every rank holds the same gathered buffers
    chk = torch.tensor([float(g_d.sum()), float(g_ids.sum())], dtype=torch.float64)
    all_chk = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(all_chk, chk)
    same = all(bool((c == all_chk[0]).all()) for c in all_chk)
    # motion batches: split by segment, results concatenated in rank order equal the unsplit run
    a, b = helpers.se2_states(3, 400), helpers.se2_states(4, 400)
    circ = np.array([[0.5, 0.5, 0.2], [0.2, 0.8, 0.1]])
    osp, ochk = orc.space_se2(), orc.checker_circles(circ)
    slo, shi = sharding.shard_range(len(a), rank, world)
    v, t = orc.check_motions(osp, ochk, a[slo:shi], b[slo:shi], last_valid=True)
    parts = [None] * world
    dist.all_gather_object(parts, (slo, v, t))
    parts.sort(key=lambda p: p[0])
    fv, ft = orc.check_motions(osp, ochk, a, b, last_valid=True)
    mot = bool((np.concatenate([p[1] for p in parts]) == fv).all() and (np.concatenate([p[2] for p in parts]) == ft).all())
    # radius queries over the sharded roadmap: counts, then padded lists, merged by (distance, id)
    def csr(points, base, radii):
        ii, dd, off = [], [], [0]
        for qv, r in zip(qs, radii):
            i, x = orc.radius(sp, points, qv, r)
            ii.append(i.astype(np.uint64) + np.uint64(base)); dd.append(x); off.append(off[-1] + len(i))
        return np.concatenate(ii), np.concatenate(dd), np.array(off)
    radii = np.where(np.arange(nq) % 5 == 0, 0.0, 0.45)     # some queries find nothing (or only exact hits)
    r_i, r_d, r_o = sharding.all_gather_radius(*csr(pts[lo:hi], lo, radii))
    w_i, w_d, w_o = csr(pts, 0, radii)
    rad = bool((r_o == w_o).all() and (r_i == w_i).all() and (r_d == w_d).all() and r_o[-1] > 1000)
    e_i, e_d, e_o = sharding.all_gather_radius(np.empty(0, np.uint64), np.empty(0), np.zeros(nq + 1, np.int64))
    rad = rad and len(e_i) == 0 and (e_o == 0).all()
    print(f"RANK{{rank}} merge_ok={{ok}} same={{same}} motion_ok={{mot}} radius_ok={{rad}}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok and same and mot and rad else 1)
""")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_world_size_2_merge_equals_single_index(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            p.kill()
            out, _ = p.communicate()
        outs.append(out)
    for rank, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {rank} failed:\n{out[-2000:]}"
        assert f"RANK{rank} merge_ok=True same=True motion_ok=True radius_ok=True" in out


GRID_WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np
    import torch
    import torch.distributed as dist
    sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
    import helpers
    from ompl_b200 import sharding
    from oracle import oracle_py as orc

    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n, nq, k, d = 4000, 64, 10, 7
    pts = helpers.with_duplicates(helpers.uniform_points(21, n, d), 22, frac=0.05)
    qs = np.vstack([helpers.uniform_points(23, nq - 8, d), pts[:8]])
    # the layout bench.py uses for N > 1: S state shards x R query blocks, rank = s * R + r
    s, r, S, R = sharding.grid_coords(rank, world, n, min_shard=n // 2)
    assert (S, R) == (2, 2) and S * R == world
    lo, hi = sharding.shard_range(n, s, S)
    nqb = nq // R
    sp = orc.space_rv(d)
    ids, dd = orc.knn(sp, pts[lo:hi], qs[r * nqb:(r + 1) * nqb], k)
    pk = sharding.PackedLists(world, nqb, k, "cpu")
    pk.ids.copy_(torch.from_numpy(ids.astype(np.int64) + lo))
    pk.dists.copy_(torch.from_numpy(dd))
    pk.all_gather()
    g_i, g_d = pk.gathered_ids().numpy(), pk.gathered_dists().numpy()
    m_i = np.empty((nq, k), dtype=np.int64)
    m_d = np.empty((nq, k), dtype=np.float64)
    for rb in range(R):   # what omplcu_merge_topk_strided_device does per block, part_stride = R rank buffers
        parts = sharding.block_part_ranks(rb, S, R)
        assert parts == [rb + j * R for j in range(S)]
        cat_i = np.concatenate([g_i[p] for p in parts], axis=1)
        cat_d = np.concatenate([g_d[p] for p in parts], axis=1)
        order = np.lexsort((cat_i, cat_d), axis=1)[:, :k]
        m_i[rb * nqb:(rb + 1) * nqb] = np.take_along_axis(cat_i, order, 1)
        m_d[rb * nqb:(rb + 1) * nqb] = np.take_along_axis(cat_d, order, 1)
    w_i, w_d = orc.knn(sp, pts, qs, k)
    ok = bool((m_i == w_i.astype(np.int64)).all() and (m_d == w_d).all())
    print(f"RANK{{rank}} grid_ok={{ok}}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)
""")


def test_world_size_4_grid_of_state_shards_and_query_blocks(tmp_path):
    script = tmp_path / "grid_worker.py"
    script.write_text(GRID_WORKER.format(root=ROOT))
    port = _free_port()
    procs = []
    for rank in range(4):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="4", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    for rank, p in enumerate(procs):
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            p.kill()
            out, _ = p.communicate()
        assert p.returncode == 0, f"rank {rank} failed:\n{out[-2000:]}"
        assert f"RANK{rank} grid_ok=True" in out
