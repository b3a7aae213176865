"""Diagnostics for the tensor-core scan (knn_scan_tc.cu): dump the raw scan values of a small
problem through the `tc_scan_values_out` option and compare them with an FP64 evaluation of
    s = |p - c|^2 - 2 (q - c).(p - c)
then check the neighbour sets against the oracle.
Run on the GPU box:  python scripts/debug_tc_scan.py [variant ...]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ompl_b200 as ob  # noqa: E402
from oracle import oracle_py as orc  # noqa: E402


def run(variant, d=100, n=1000, nq=256, k=10):
    r = np.random.default_rng(5)
    pts = r.uniform(-np.pi, np.pi, size=(n, d))
    qs = r.uniform(-np.pi, np.pi, size=(nq, d))
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    nn.set_option("tc_scan", 1)
    nn.set_option("tc_variant", variant)
    units = (n + 127) // 128
    qpad = (nq + 127) // 128 * 128
    buf = torch.full((qpad, units * 128), float("nan"), dtype=torch.float32, device="cuda")
    nn.set_option("tc_scan_values_out", buf.data_ptr())
    ids, dist = nn.nearestKBatch(qs, k)
    nn.set_option("tc_scan_values_out", 0)
    torch.cuda.synchronize()
    got = buf.cpu().numpy().astype(np.float64)[:nq, :n]
    c = pts.mean(axis=0)
    pt = (pts - c).astype(np.float32).astype(np.float64)
    qt = (qs - c).astype(np.float32).astype(np.float64)
    want = (pt * pt).sum(1)[None, :] - 2.0 * qt @ pt.T
    err = np.abs(got - want)
    scale = (pt * pt).sum(1).max() + 2 * np.sqrt((qt * qt).sum(1).max() * (pt * pt).sum(1).max())
    oi, od = orc.knn(orc.space_rv(d), pts, qs, k, nthreads=4)
    print(
        f"variant {variant}: last_tc={nn.stat('last_tc')} nan={int(np.isnan(got).sum())} "
        f"max|err|={np.nanmax(err):.3e} ({np.nanmax(err) / scale / 2**-24:.1f} u of {scale:.1f}) "
        f"ids_equal={bool((ids == oi).all())} dists_equal={bool((dist == od).all())} "
        f"fallbacks={nn.stat('last_fallbacks')}",
        flush=True,
    )
    if np.nanmax(err) > 1.0:
        bad = np.argwhere(err > 1.0)
        print("   first bad entries (q, p):", bad[:6].tolist(), "got", got[tuple(bad[0])], "want", want[tuple(bad[0])])
        print("   bad rows:", np.unique(bad[:, 0])[:16].tolist(), "bad cols:", np.unique(bad[:, 1])[:16].tolist())


if __name__ == "__main__":
    # usage: debug_tc_scan.py [d ...]   (variant 1 swaps the descriptor fields and must fail)
    for d in [int(a) for a in sys.argv[1:]] or [100]:
        run(0, d=d)
