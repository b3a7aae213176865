"""CPU models of two pieces of device arithmetic whose correctness is combinatorial rather than numerical, kept next to
the GPU parity tests that exercise the kernels themselves:

* the spread order of motion_group_kernel (ompl_b200/csrc/validity.cu, `spread`): every state of a segment is visited
  exactly once, whatever the lane-group width;
* the rank rule of merge_sorted_kernel (ompl_b200/csrc/nn.cu): binary searches by distance in sorted per-shard lists plus
  the run of equal distances by id give the ranks of a full sort by (distance, id), with absent entries, ties across
  and inside lists, and the early stop at k - rank entries.
"""
import math

import numpy as np
import pytest

NONE = 2**64 - 1


def _spread_visits(cnt, G):
    blk = (cnt + G - 1) // G
    stepc = max((blk * 5 + 4) // 8, 1)
    while not (math.gcd(stepc, blk) == 1 or blk == 1):
        stepc += 1
    assert stepc <= max(blk, 1)  # one conditional subtraction keeps the offset inside the block
    seen, off, passes = [], 0, 0
    while passes < blk:
        for gl in range(G):
            m = gl * blk + off
            if m < cnt:
                seen.append(m)
        passes += 1
        off += stepc
        if off >= blk:
            off -= blk
    return seen


@pytest.mark.parametrize("G", [4, 8, 16, 32])
def test_spread_order_visits_every_state_once(G):
    for cnt in list(range(1, 400)) + [1000, 4097, 65536]:
        seen = _spread_visits(cnt, G)
        assert sorted(seen) == list(range(cnt)), (cnt, G)


def test_spread_order_first_pass_spans_the_segment():
    seen = _spread_visits(72, 4)
    assert seen[:4] == [0, 18, 36, 54]
    # successive passes do not visit neighbours of the previous pass
    assert abs(seen[4] - seen[0]) > 4


def _merge_sorted(d, i, k):
    P = len(d)
    out_i, out_d = [NONE] * k, [math.inf] * k
    for p in range(P):
        for j in range(k):
            ident = i[p][j]
            if ident == NONE:
                continue
            key, rank, keep = d[p][j], 0, True
            for t in range(P):
                if not keep:
                    break
                pp = p if t == 0 else (t - 1 if t <= p else t)
                lo, hi = 0, k - rank
                while lo < hi:
                    mid = (lo + hi) // 2
                    if i[pp][mid] != NONE and d[pp][mid] < key:
                        lo = mid + 1
                    else:
                        hi = mid
                rank += lo
                e = lo
                while e < k and rank < k:
                    if i[pp][e] == NONE or d[pp][e] != key:
                        break
                    rank += 1 if i[pp][e] < ident else 0
                    e += 1
                keep = rank < k
            if keep:
                assert out_i[rank] == NONE, "two candidates with one rank"
                out_i[rank], out_d[rank] = ident, key
    return out_i, out_d


def test_merge_sorted_rank_rule_equals_a_full_sort():
    rng = np.random.default_rng(0)
    for _ in range(400):
        P, k = int(rng.integers(1, 7)), int(rng.integers(1, 24))
        ids = rng.permutation(4000)[: P * k]
        d, i, allc = [], [], []
        for p in range(P):
            m = int(rng.integers(0, k + 1))
            dd = np.sort(rng.integers(0, 6, m).astype(float))  # few distinct distances: long runs of ties
            ii = [int(x) for x in ids[p * k: p * k + m]]  # ids inside a run in any order
            d.append(list(dd) + [math.inf] * (k - m))
            i.append(ii + [NONE] * (k - m))
            allc += list(zip([float(x) for x in dd], ii))
        allc.sort()
        oi, od = _merge_sorted(d, i, k)
        got = [(a, b) for a, b in zip(od, oi) if b != NONE]
        assert got == allc[:k]
