#!/usr/bin/env python
"""Secondary workloads of SURVEY.md 8(d) (M1 at 4M / 16M states, M1 with ties, M2 hypercube, M4 chain),
each timed through the host-pointer API (what a caller sees) with a parity sample against the oracle.
bench.py carries the headline (M1 at 1M states, M3); this script produces the table in DESIGN.md 9.

    python scripts/bench_workloads.py [m1 m1ties m2 m4] [--log=profiles/r2_workloads.log] > profiles/r2_workloads.jsonl
(--log: the same lines as a tools::Benchmark log for the reference's scripts/ompl_benchmark_statistics.py)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
import ompl_b200 as ob  # noqa: E402
from oracle import oracle_py as orc  # noqa: E402

CORES = os.cpu_count() or 1
PI = np.pi


def timed(fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        best = min(best, time.perf_counter() - t0)
    return best, out


RECORDS = []


def emit(**kw):
    RECORDS.append(kw)
    print(json.dumps(kw), flush=True)


def write_benchmark_log(path):
    """The emitted lines as a tools::Benchmark log (one "planner" per workload line, one run each), which the
    reference's scripts/ompl_benchmark_statistics.py turns into its sqlite database."""
    from ompl_b200 import benchmark_log as bl

    log = bl.BenchmarkLog("ompl_b200_workloads", run_count=1)
    log.setup_info = "SURVEY 8(d) workloads M1-M4 through the host-pointer API of libomplcu.so\n"
    log.cpu_info = f"{CORES} host threads\n"
    for i, r in enumerate(RECORDS):
        name = "_".join(str(r.get(k)) for k in ("workload", "n_states", "k", "dim", "links") if r.get(k) is not None)
        pl = log.add_planner(f"NearestNeighborsCUDA_{name}_{i}", {"nearest_neighbors STRING": "NearestNeighborsCUDA"})
        run = {}
        for k, v in r.items():
            if isinstance(v, bool):
                run[f"{k.replace('_', ' ')} BOOLEAN"] = "1" if v else "0"
            elif isinstance(v, int):
                run[f"{k.replace('_', ' ')} INTEGER"] = str(v)
            elif isinstance(v, float):
                run[f"{k.replace('_', ' ')} REAL"] = repr(v)
        if "ms" in r:
            run["time REAL"] = repr(r["ms"] * 1e-3)
            log.total_duration += r["ms"] * 1e-3
        pl.runs.append(run)
    with open(path, "w") as f:
        log.save(f)


def m1(n_list=(1_000_000, 4_000_000, 16_000_000), nq=65536, k=10, d=7):
    qs = helpers.uniform_points(1001, nq, d)
    for n in n_list:
        pts = helpers.uniform_points(1, n, d)
        nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
        t0 = time.perf_counter()
        nn.add_rows(pts)  # ids = insertion indices 0..n-1 (a clear() would not restart them)
        t_add = time.perf_counter() - t0
        t, (ids, dd) = timed(lambda: nn.nearestKBatch(qs, k))
        m = 64
        oi, od = orc.knn(orc.space_rv(d), pts, qs[:m], k, nthreads=CORES)
        emit(workload="M1", n_states=n, n_queries=nq, k=k, dim=d, queries_per_s=nq / t, ms=t * 1e3,
             add_ms=t_add * 1e3, fallbacks=nn.stat("last_fallbacks"),
             parity_sample=m, ids_equal=bool((ids[:m] == oi).all()), dists_equal=bool((dd[:m] == od).all()))
        del nn, pts


def m1ties(n=1_000_000, nq=65536, k=10, d=7):
    pts = helpers.with_duplicates(helpers.uniform_points(2, n, d), 5, frac=0.01)
    qs = np.vstack([helpers.uniform_points(1002, nq - 1024, d), pts[:1024]])  # 1024 queries sit on duplicated points
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    t, (ids, dd) = timed(lambda: nn.nearestKBatch(qs, k))
    sel = np.r_[0:32, nq - 32:nq]
    oi, od = orc.knn(orc.space_rv(d), pts, qs[sel], k, nthreads=CORES)
    emit(workload="M1 + 1% exact duplicates", n_states=n, n_queries=nq, k=k, queries_per_s=nq / t, ms=t * 1e3,
         fallbacks=nn.stat("last_fallbacks"), parity_sample=len(sel), ids_equal=bool((ids[sel] == oi).all()),
         dists_equal=bool((dd[sel] == od).all()))


def m2(n=100_000, d=8):
    """HypercubeBenchmark: 100K valid samples by rejection, PRM* (k=36) and FMT* (k=2100) neighbourhoods,
    then every (sample, neighbour) edge through the motion validator at resolution 0.001."""
    sp = ob.RealVectorStateSpace(d)
    sp.setLongestValidSegmentFraction(0.001)
    chk = ob.HypercubeValidityChecker(d)
    r = helpers.rng(8)
    got = []
    tried = 0
    t0 = time.perf_counter()
    while sum(len(g) for g in got) < n:
        # the free space is a thin corridor: mix corridor samples with uniform ones like a bridge sampler would
        cand = np.vstack([helpers.hypercube_valid_states(int(r.integers(1 << 30)), 200_000, d) +
                          r.normal(0, 0.02, (200_000, d)), r.uniform(0, 1, (200_000, d))]).clip(0, 1)
        tried += len(cand)
        got.append(cand[chk.isValidBatch(cand)])
    t_samp = time.perf_counter() - t0
    s = np.vstack(got)[:n]
    emit(workload="M2 sampling", tried=tried, valid=n, states_per_s=tried / t_samp)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.add_rows(s)
    osp = orc.space_rv(d, 0.0, 1.0, 0.001)
    for k in (36, 2100):
        t, (ids, dd) = timed(lambda: nn.nearestKBatch(s, k), reps=1 if k > 100 else 2)
        m = 16
        oi, od = orc.knn(osp, s, s[:m], k, nthreads=CORES)
        emit(workload=f"M2 kNN k={k}", n_states=n, n_queries=n, queries_per_s=n / t, ms=t * 1e3,
             path="scan" if k <= 40 else "exact (segmented radix sort)", parity_sample=m,
             ids_equal=bool((ids[:m] == oi).all()), dists_equal=bool((dd[:m] == od).all()))
        if k == 36:
            nb = ids[:, 1:].astype(np.int64)  # drop self
            a = np.repeat(s, nb.shape[1], axis=0)
            b = s[nb.reshape(-1)]
            mv = ob.BatchedDiscreteMotionValidator(sp, chk)
            t, (v, lt, _) = timed(lambda: mv.checkMotionBatch(a, b, last_valid=True), reps=2)
            mm = 20000
            ov, ot = orc.check_motions(osp, orc.checker_hypercube(d), a[:mm], b[:mm], last_valid=True, nthreads=CORES)
            emit(workload="M2 motion: all (sample, neighbour) edges, resolution 0.001", n_segments=len(a),
                 segments_per_s=len(a) / t, ms=t * 1e3, valid_fraction=float(v.mean()),
                 states_in_segments=chk.stat("last_states"), states_evaluated=chk.stat("last_states_evaluated"),
                 lanes_per_segment=chk.stat("last_group"), parity_sample=mm,
                 valid_equal=bool((v[:mm] == ov).all()), last_t_equal=bool((lt[:mm][~ov] == ot[~ov]).all()))


def m4(n=100_000):
    for links in (10, 100):
        L = 1.0 / links
        env = orc.horn_environment(links, np.log(links) / links)
        chk, ochk = ob.KinematicChainValidityChecker(links, L, env), orc.checker_chain(links, L, env)
        r = helpers.rng(links)
        s = np.vstack([r.uniform(-PI, PI, (n // 2, links)), r.normal(0, 0.15, (n // 2, links))])
        t, v = timed(lambda: chk.isValidBatch(s), reps=2)
        m = 2000 if links == 10 else 200
        sel = np.r_[0:m // 2, n - m // 2:n]
        t0 = time.perf_counter()
        ov = orc.check_states(ochk, s[sel], nthreads=CORES)
        cpu = (time.perf_counter() - t0) / len(sel)
        emit(workload=f"M4 chain isValid, {links} links", n_states=n, states_per_s=n / t, ms=t * 1e3,
             valid_fraction=float(v.mean()), cpu_states_per_s=1.0 / cpu, cpu_threads=CORES, parity_sample=len(sel),
             valid_equal=bool((v[sel] == ov).all()))
        # nearestR with the chain metric (exact path): 1024 queries against 100K states
        sp, osp = ob.KinematicChainSpace(links, L), orc.space_chain(links, L)
        nn = ob.NearestNeighborsCUDA(sp)
        nn.add_rows(s)
        q = s[::n // 1024][:1024]
        rad = 0.5 if links == 10 else 0.35
        t, (rids, rd, offs) = timed(lambda: nn.nearestRBatch(q, rad), reps=1)
        want = orc.radius(osp, s, q[3], rad)
        got = rids[offs[3]:offs[4]]
        emit(workload=f"M4 chain nearestR(r={rad}), {links} links", n_states=n, n_queries=len(q),
             queries_per_s=len(q) / t, ms=t * 1e3, mean_neighbours=float(offs[-1]) / len(q),
             parity_query_ids_equal=bool(list(got) == list(want[0])), neighbours_of_query_3=len(want[0]))
    # plain L2 at d = 100 (SURVEY: "separately, plain L2 at d = 100"): register-tiled FP32 scan (knn_scan_hd.cuh)
    d = 100
    pts, qs = helpers.uniform_points(4, n, d), helpers.uniform_points(5, 65536, d)
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    t, (ids, dd) = timed(lambda: nn.nearestKBatch(qs, 10), reps=1)
    oi, od = orc.knn(orc.space_rv(d), pts, qs[:16], 10, nthreads=CORES)
    emit(workload="M4 plain L2, d=100, k=10 (high-dimensional scan)", n_states=n, n_queries=len(qs), queries_per_s=len(qs) / t,
         ms=t * 1e3, fallbacks=nn.stat("last_fallbacks"), scan_path=nn.stat("last_path") == 1,
         ids_equal=bool((ids[:16] == oi).all()), dists_equal=bool((dd[:16] == od).all()))


if __name__ == "__main__":
    which = [a for a in sys.argv[1:] if not a.startswith("--log=")] or ["m1", "m1ties", "m2", "m4"]
    for w in which:
        {"m1": m1, "m1ties": m1ties, "m2": m2, "m4": m4}[w]()
    for a in sys.argv[1:]:
        if a.startswith("--log="):
            write_benchmark_log(a[len("--log="):])
