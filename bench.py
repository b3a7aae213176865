#!/usr/bin/env python
"""bench.py -- the reference's headline metric for the hot path on B200.

metric   kNN queries/s, k = 10, 1M states (BASELINE.json `metric`)
workload BASELINE configs[2]: RealVectorStateSpace(7), 1M states uniform in
         [0,1]^7, one batch of 65536 queries, k = 10 (synthetic, seeded)
step     one batched nearestK over the whole query batch

  value      device-resident: queries / ids / distances live in HBM
  e2e        through the host-pointer C-ABI call omplcu_nn_knn (what the C++
             adapter NearestNeighborsCUDA::nearestKBatch calls): H2D of the
             queries and D2H of ids + distances inside the timed region
  roofline   the dominant kernel: the tensor-core filter scan (knn_scan_tc_kernel:
             tcgen05.mma on BF16 split operands, selection straight from TMEM); its
             duration is measured live with CUDA events on the launching stream and
             put against the measured dense BF16 peak (MEASURED_PEAKS.json).
             roofline_fp32_scan: the same step with the FP32-pipe scan kernel
             (knn_scan_kernel, option tc_scan = 0) against the FP32 FMA peak
  cpu_baseline / --impl reference
             the reference's own NearestNeighborsGNAT (oracle/_ref, compiled
             from the reference headers) on all host threads; falls back to the
             oracle port of NearestNeighborsLinear when _ref was not built

N > 1 (torchrun, one rank per GPU): the 1M states are sharded by insertion
index across ranks (strong scaling); every rank answers all queries on its
shard, the per-rank top-k lists are exchanged with one NCCL all-gather and
merged by (distance, id).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_STATES, N_QUERIES, K, DIM = 1_000_000, 65536, 10, 7
N_SEGMENTS = 1_000_000
MOTION_METRIC = "motion checks/sec (1M SE3 segments, 1000 obstacles, resolution 0.01)"
METRIC = "kNN queries/sec (k=10, 1M states)"
TC_CAPTURE = "r2_scan_tc_d7_ncu_summary.csv"    # ncu --set full of the tensor-core scan (roofline.traffic is read from it)
FP32_CAPTURE = "r1_scan_v3_ncu_summary.csv"     # ... of the FP32-pipe scan
MOTION_CAPTURE = "r2_motion_v6_ncu_summary.csv"  # ... of motion_group_kernel
UNIT = "queries/s"


def bench_config():
    """The workload, with the SAME keys in both arms (the driver compares the two dicts)."""
    return {"workload": "synthetic batched kNN: RealVectorStateSpace(7), "
                        f"{N_STATES // 1_000_000}M states, 64K queries, k=10",
            "n_states": N_STATES, "n_queries": N_QUERIES, "k": K, "dim": DIM}


def ncu_traffic(summary_csv):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch, read from a committed ncu summary under
    profiles/ (scripts/ncu_summary.py format); None if the file or the metrics are absent."""
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot, seen = 0.0, 0
    try:
        for ln in open(os.path.join(ROOT, "profiles", summary_csv)):
            f = [x.strip().strip('"') for x in ln.split(",")]
            if len(f) >= 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and f[1] in unit:
                tot += float(f[2]) * unit[f[1]]
                seen += 1
    except (OSError, ValueError):
        return None
    return tot if seen == 2 else None


def make_inputs(n_states=None, n_queries=N_QUERIES):
    n_states = N_STATES if n_states is None else n_states
    pts = np.random.default_rng(1).random((n_states, DIM))
    qs = np.random.default_rng(1001).random((n_queries, DIM))
    return pts, qs


# ---------------------------------------------------------------------------
# clocks: sample nvidia-smi during the timed region
# ---------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([x.strip() for x in ln.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------
# reference arm / cpu baseline (the ONLY place bench.py executes oracle/)
# ---------------------------------------------------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def build_cpu_reference(pts):
    from oracle import oracle_py as orc

    sp = orc.space_rv(DIM)
    if orc.have_ref():
        t0 = time.perf_counter()
        nn = orc.RefNN(orc.RefNN.GNAT, sp)
        nn.add(pts, reserve_query_slots=4096)
        build_s = time.perf_counter() - t0

        def run(q, nthreads):
            t0 = time.perf_counter()
            nn.knn(q, K, nthreads)
            return time.perf_counter() - t0

        return "reference", "NearestNeighborsGNAT (reference headers, oracle/_ref)", build_s, run

    def run(q, nthreads):
        t0 = time.perf_counter()
        orc.knn(sp, pts, q, K, nthreads)
        return time.perf_counter() - t0

    return "port", "NearestNeighborsLinear restatement (oracle port)", 0.0, run


def cpu_sample_size(kind, cores):
    # ~10-20 s of CPU work: GNAT ~800 q/s/thread, Linear port ~150 q/s/thread at N = 1M
    per = 8192 if kind == "reference" else 1024
    return per * cores


def cpu_motion_baseline(a, b, robot, sph, box, cores, m):
    """DiscreteMotionValidator::checkMotion(s1, s2, lastValid) on the host cores (oracle port, pthreads)."""
    from oracle import oracle_py as orc

    osp, ochk = orc.space_se3(-1.0, 1.0), orc.checker_se3(robot, sph, box)
    t0 = time.perf_counter()
    ov, ot = orc.check_motions(osp, ochk, a[:m], b[:m], last_valid=True, nthreads=cores)
    return time.perf_counter() - t0, ov, ot


MOTION_CPU_WHAT = ("oracle restatement of DiscreteMotionValidator::checkMotion(s1,s2,lastValid) + SE3 obstacle "
                   "predicate (bisection order, early exit)")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pts, qs_all = make_inputs()
    cores = host_threads()
    kind, what, build_s, run = build_cpu_reference(pts)
    # a bounded sample of the SAME 64K queries per step: all of them for GNAT on >= 16 cores (about 3.5 s
    # per step), a slice for fewer cores or for the Linear port
    per_step = min(max(cores * (4096 if kind == "reference" else 128), 64), N_QUERIES)
    nblocks = N_QUERIES // per_step

    def block(i):
        j = i % nblocks
        return qs_all[j * per_step:(j + 1) * per_step]

    for w in range(args.warmup):
        run(block(w), cores)
    t = 0.0
    for s in range(args.warmup, args.warmup + args.steps):
        t += run(block(s), cores)
    value = per_step * args.steps / t
    sample = (f"{per_step} of the 65536 queries per step x {args.steps} steps vs the same 1M states, {cores} threads, "
              f"{what}")
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": bench_config(),
        "queries_per_step": per_step,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
                         "build_s": build_s},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.no_motion:
        from ompl_b200 import workloads

        m = min(N_SEGMENTS, 12000 * cores)
        s1, s2 = workloads.m3_segments(N_SEGMENTS)
        sph, box = workloads.se3_obstacles()
        secs, _, _ = cpu_motion_baseline(s1, s2, workloads.se3_robot(), sph, box, cores, m)
        out["motion"] = {"metric": MOTION_METRIC, "value": m / secs, "unit": "segments/s",
                         "cpu_baseline": {"value": m / secs, "unit": "segments/s", "cores": cores, "kind": "port",
                                          "sample": f"first {m} of the 1M segments, {cores} threads, "
                                                    + MOTION_CPU_WHAT}}
    print(json.dumps(out))


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch

    import ompl_b200 as ob
    from ompl_b200._lib import check, lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    nccl = None  # process group of the device-timed legs (created after the one-process e2e leg, see below)
    if world > 1:
        import torch.distributed as dist

        # default group: gloo (host-side barriers and the rendezvous); the NCCL group comes later
        dist.init_process_group("gloo")
    dev = torch.device("cuda", local)
    L = lib()

    from ompl_b200 import sharding

    pts, qs = make_inputs()
    # N > 1: S state shards x R query blocks (sharding.grid_shape); rank = s * R + r
    S_SH, R_QB = (1, 1) if args.shard == "single" else sharding.grid_shape(world, N_STATES)
    if args.shard == "states":
        S_SH, R_QB = world, 1
    sh_s, sh_r = divmod(rank, R_QB)  # sharding.grid_coords; block r is merged from sharding.block_part_ranks(r, S, R)
    lo, hi = sharding.shard_range(N_STATES, sh_s, S_SH)
    assert N_QUERIES % R_QB == 0
    NQB = N_QUERIES // R_QB           # queries answered by this rank
    qlo = sh_r * NQB
    e2e_note = "omplcu_nn_knn on one handle: host buffers in, host buffers out"
    e2e_wall_ms, motion_e2e, e2e_ids = None, None, None
    if world > 1:
        # End to end at N GPUs = what ONE caller process gets from a roadmap sharded over the N GPUs:
        # omplcu_nn_create_sharded(devices[0..N-1]) and the same host-pointer call.  Rank 0 drives all GPUs (the
        # library has one host thread per shard) while the other ranks wait at a host-side barrier.  This leg
        # runs BEFORE the NCCL communicator exists: with torch's NCCL peer mappings in place, rank 0's own
        # contexts on the other GPUs made cudaMalloc fail with "resource already mapped".
        dist.barrier()
        if rank == 0:
            e2e_wall_ms, motion_e2e, e2e_ids = sharded_e2e(args, torch, world, pts, qs, not args.no_motion)
        dist.barrier()
        e2e_note = (f"omplcu_nn_create_sharded over {world} GPUs in ONE process (rank 0), omplcu_nn_knn with host "
                    "buffers: one H2D, NVLink peer copies to the shards, per-shard scan, peer gather, device merge, "
                    "one D2H; results identical to the device-timed leg")
        # NCCL prints its version banner on stdout when the first communicator comes up; stdout is
        # reserved for the one JSON line, so park it on stderr until the communicator exists
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            nccl = dist.new_group(backend="nccl", device_id=torch.device("cuda", local))
            warm = torch.zeros(1, device=torch.device("cuda", local))
            dist.all_reduce(warm, group=nccl)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    sp = ob.RealVectorStateSpace(DIM, device=local)
    nn = ob.NearestNeighborsCUDA(sp)
    nn.set_id_base(lo)
    nn.add_rows(pts[lo:hi])
    nn.set_option("profile", 1)
    if args.scan_ctas:
        nn.set_option("scan_ctas_per_sm", args.scan_ctas)
    if args.no_sample_thr:
        nn.set_option("sample_thresholds", 0)
    if args.fp32_scan:
        nn.set_option("tc_scan", 0)

    qs_all = qs
    qs = np.ascontiguousarray(qs_all[qlo:qlo + NQB])
    d_q = torch.from_numpy(qs).to(dev)
    packed = sharding.PackedLists(world, NQB, K, dev)  # ids + distances in one buffer: one all-gather
    d_ids, d_d = packed.ids, packed.dists
    if world > 1:
        m_ids = torch.empty((N_QUERIES, K), dtype=torch.int64, device=dev)
        m_d = torch.empty((N_QUERIES, K), dtype=torch.float64, device=dev)

    def merge_blocks():
        # block r: the S lists of ranks s * R + r, part_stride = R rank buffers apart
        g_i, g_d = packed.gathered_ids(), packed.gathered_dists()
        for r in range(R_QB):
            check(L.omplcu_merge_topk_strided_device(local, g_d[r].data_ptr(), g_i[r].data_ptr(), S_SH,
                                                     R_QB * packed.part_stride, NQB, K,
                                                     m_d[r * NQB:].data_ptr(), m_ids[r * NQB:].data_ptr(), sptr))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream()
    sptr = C.c_void_p(stream.cuda_stream)

    def step_device():
        check(L.omplcu_nn_knn_device(nn.handle, d_q.data_ptr(), NQB, K, d_ids.data_ptr(), d_d.data_ptr(), sptr))
        if world > 1:
            packed.all_gather(group=nccl)
            merge_blocks()

    h_ids = np.empty((NQB, K), dtype=np.uint64)
    h_d = np.empty((NQB, K), dtype=np.float64)
    h_cnt = np.empty(NQB, dtype=np.uint32)
    h_all_ids = np.empty((N_QUERIES, K), dtype=np.uint64) if world > 1 else h_ids
    h_all_d = np.empty((N_QUERIES, K), dtype=np.float64) if world > 1 else h_d
    pin = lambda a: torch.cuda.cudart().cudaHostRegister(a.ctypes.data, a.nbytes, 0)  # noqa: E731
    for a in (qs, h_ids, h_d):
        pin(a)

    def step_e2e():
        # host buffers in, host buffers out: the reference-facing call
        check(L.omplcu_nn_knn(nn.handle, qs.ctypes.data, NQB, K, h_ids.ctypes.data, h_d.ctypes.data,
                              h_cnt.ctypes.data))

    def barrier():
        if world > 1:
            dist.barrier(group=nccl)
        torch.cuda.synchronize()

    scan_mark = [0, 0]

    def timed(step, steps, warmup, sampler=None):
        for _ in range(warmup):
            flush.fill_(1)
            step()
        barrier()
        # the kernel time reported with the roofline is the mean over the TIMED steps only
        scan_mark[0], scan_mark[1] = nn.stat("scan_ns_total"), nn.stat("scan_count")
        if sampler:
            sampler.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        wall = 0.0
        for a, b in ev:
            flush.fill_(1)  # L2 flush between timed iterations (not timed)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            a.record(stream)
            step()
            b.record(stream)
            b.synchronize()
            wall += time.perf_counter() - t0
        barrier()
        clocks = sampler.stop() if sampler else None
        dev_ms = sum(a.elapsed_time(b) for a, b in ev)
        # the host-pointer path synchronises inside the call; wall and device time agree there
        ms = max(dev_ms, 0.0)
        t = torch.tensor([ms, wall * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=nccl)
        return float(t[0]), float(t[1]), clocks

    l0 = nn.stat("launches")
    nn.set_option("profile", 1)
    sampler = ClockSampler(local) if rank == 0 else None
    dev_ms, _, clocks = timed(step_device, args.steps, args.warmup, sampler)
    launches = nn.stat("launches") - l0 + (args.steps + args.warmup) * (R_QB if world > 1 else 0)
    launches_timed = launches * args.steps // (args.steps + args.warmup)
    scan_ns = (nn.stat("scan_ns_total") - scan_mark[0]) / max(nn.stat("scan_count") - scan_mark[1], 1)
    fallbacks = nn.stat("last_fallbacks")
    used_tc = nn.stat("last_tc") == 1
    res_ids = (m_ids if world > 1 else d_ids).cpu().numpy().astype(np.uint64)
    if world == 1:
        _, e2e_wall_ms, _ = timed(step_e2e, args.steps, max(args.warmup, 1))
        # sanity: device-resident and host-pointer results agree (same code path underneath)
        assert (res_ids == h_all_ids).all(), "device and host entry points disagree"

    # the same step on the FP32-pipe scan kernel (the round-1 default), for the second roofline
    fp32_leg = None
    if used_tc:
        nn.set_option("tc_scan", 0)
        nn.set_option("profile", 1)
        f_ms, _, _ = timed(step_device, max(args.steps // 2, 2), 3)
        fp32_leg = (f_ms / max(args.steps // 2, 2),
                    (nn.stat("scan_ns_total") - scan_mark[0]) / max(nn.stat("scan_count") - scan_mark[1], 1))
        nn.set_option("tc_scan", -1)

    motion = None if args.no_motion else motion_block(args, torch, dist, nccl, dev, local, rank, world, timed)

    if world > 1 and rank == 0:
        assert (e2e_ids == res_ids).all(), "sharded handle (one process) and the torchrun ranks disagree"
        if motion is not None and motion_e2e is not None:
            motion["e2e"] = motion_e2e

    if rank == 0:
        t_step = dev_ms / args.steps
        value = N_QUERIES / (t_step * 1e-3)
        e2e_val = N_QUERIES / (e2e_wall_ms / args.steps * 1e-3)
        # rooflines of the dominant kernel (the filter scan)
        ffma, ffma2 = C.c_double(), C.c_double()
        check(L.omplcu_measure_fp32_peak(local, C.byref(ffma), C.byref(ffma2)))
        fp32_peak = max(ffma.value, ffma2.value)
        pairs = float(NQB) * (hi - lo)
        alg_flops = 3.0 * DIM * pairs           # sub, mul, add per dimension
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        fp32_src = ("FP32 FMA pipe measured live on this GPU (omplcu_measure_fp32_peak: FFMA "
                    f"{ffma.value:.1f}, FFMA2 {ffma2.value:.1f} TFLOP/s); MEASURED_PEAKS.json has no FP32 figure")

        def fp32_roofline(kernel_ns, step_ms):
            exe = 2.0 * DIM * pairs             # what the kernel issues: DIM FMAs per pair
            a = alg_flops / (kernel_ns * 1e-9) / 1e12
            scan_bytes = 4.0 * 8 * (hi - lo) * (NQB / 512.0)  # every query tile streams the filter copy
            return {
                "bound": "fp32", "kernel": "knn_scan_kernel<7,4,12>", "achieved": a, "peak": fp32_peak,
                "unit": "TFLOP/s", "frac": a / fp32_peak if fp32_peak else None,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full
                # (profiles/r1_scan_v3_ncu_summary.csv); algorithmic: 32 MB filter copy + 1.8 MB queries + lists
                "traffic": ncu_traffic(FP32_CAPTURE) if world == 1 else None,
                "traffic_unit": f"bytes/launch (profiles/{FP32_CAPTURE})",
                "peak_source": fp32_src, "flops_per_pair_algorithmic": 3 * DIM, "flops_per_pair_executed": 2 * DIM,
                "frac_executed": exe / (kernel_ns * 1e-9) / 1e12 / fp32_peak if fp32_peak else None,
                "kernel_ms": kernel_ns * 1e-6, "kernel_share_of_step": kernel_ns * 1e-6 / step_ms,
                "hbm_view": {"l2_stream_gbs": scan_bytes / (kernel_ns * 1e-9) / 1e9, "hbm_peak_gbs": hbm_peak,
                             "note": "filter copy (32 MB) is L2-resident; HBM is not the bound for batched queries"},
            }

        roofline_fp32 = None
        if used_tc:
            kpad = (3 * DIM + 3 + 15) // 16 * 16  # K issued: ceil((3 d + 3) / 16) tcgen05.mma of K = 16
            tpeak = peaks.get("bf16_tflops", 2250.0)
            ach = alg_flops / (scan_ns * 1e-9) / 1e12
            exe = 2.0 * kpad * pairs            # MMA flops issued: K = 3d + 3 padded to 16
            roofline = {
                "bound": "tensor", "kernel": "knn_scan_tc_kernel<16,4> (+ query image kernel)", "achieved": ach,
                "peak": tpeak, "unit": "TFLOP/s", "frac": ach / tpeak,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full
                # (profiles/r1_scan_tc_ncu_summary.csv); algorithmic: 64 MB operand image + queries + lists
                "traffic": ncu_traffic(TC_CAPTURE) if world == 1 else None,
                "traffic_unit": f"bytes/launch (profiles/{TC_CAPTURE})",
                "peak_source": "MEASURED_PEAKS.json bf16_tflops (dense BF16 cuBLAS, burst: the kernel is timed alone)"
                if "bf16_tflops" in peaks else "fallback 2250 TFLOP/s (B200_PROFILING.md): MEASURED_PEAKS.json absent",
                "flops_per_pair_algorithmic": 3 * DIM, "flops_per_pair_executed": 2 * kpad,
                "frac_executed": exe / (scan_ns * 1e-9) / 1e12 / tpeak,
                "kernel_ms": scan_ns * 1e-6, "kernel_share_of_step": scan_ns * 1e-6 / t_step,
                "vs_fp32_pipe_peak": ach / fp32_peak if fp32_peak else None,
                "note": "K = 32 per accumulator, so the tensor pipe is not the bound (ncu round 2: 38 % busy) and neither is HBM: "
                        "the kernel is bound by instruction issue and latency in the selection warps (issue slots 56 % busy; of the "
                        "executed instructions 30 % are the FMNMX3 min-trees over the accumulators, 23 % branch / sync / nop overhead "
                        "of the per-half-tile TMEM hand-over, 19 % integer bookkeeping and list insertions, 9 % compares, 15 % "
                        "mbarrier polls; profiles/r2_scan_tc_d7_instruction_mix.txt). Every FP32 accumulator value costs at least 1 "
                        "issue clock per SM sub-partition whatever instruction looks at it (scripts/ubench_reduce.cu), i.e. 1.76 ms "
                        "for this launch; achieved exceeds the FP32 FMA peak (vs_fp32_pipe_peak > 1), which bounded the FP32 scan "
                        "(roofline_fp32_scan)",
            }
            if fp32_leg:
                roofline_fp32 = fp32_roofline(fp32_leg[1], fp32_leg[0])
                roofline_fp32["ms_per_step"] = fp32_leg[0]
                roofline_fp32["queries_per_s"] = N_QUERIES / (fp32_leg[0] * 1e-3)
        else:
            roofline = fp32_roofline(scan_ns, t_step)
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None,
            "dtype": "bf16x2 split filter on tensor cores (f32 accumulate) + f64 exact re-rank" if used_tc
            else "f32 filter + f64 exact re-rank", "data": "synthetic",
            "config": bench_config(),
            "sharding": f"{S_SH} state shard(s) by insertion index x {R_QB} query block(s) over {world} ranks; "
                        "one NCCL all-gather of the packed (distance, id) lists + merge by (distance, id)"
            if world > 1 else "single GPU",
            "l2": "flushed between timed iterations (256 MB write)",
            "clocks": clocks,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(qs_all.nbytes),
                    "d2h_bytes_per_step": int(N_QUERIES * K * 16), "path": e2e_note},
            "gpu_launches": int(launches_timed),
            "roofline": roofline,
            "exact_fallback_queries": int(fallbacks),
        }
        if roofline_fp32 is not None:
            out["roofline_fp32_scan"] = roofline_fp32
        if motion is not None:
            out["motion"] = motion
        if world == 1 and not args.no_motion:
            out["other_calls"] = other_calls(nn, pts, qs, local)
        if world == 1 and not args.no_cpu:
            cores = host_threads()
            kind, what, build_s, run = build_cpu_reference(pts)
            nsamp = min(cpu_sample_size(kind, cores), N_QUERIES)
            secs = run(qs[:nsamp], cores)
            out["cpu_baseline"] = {"value": nsamp / secs, "unit": UNIT, "cores": cores, "kind": kind,
                                   "sample": f"first {nsamp} of the 65536 queries vs the same 1M states, "
                                             f"{cores} threads, {what}; build {build_s:.2f} s not counted"}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def other_calls(nn, pts, qs, local):
    """Secondary figures, N = 1: the literal one-query call a planner makes (PRM::addMilestone, prm/src/PRM.cpp:575;
    RRT's nn_->nearest, rrt/src/RRT.cpp:170) and kNN on an SE(3) roadmap, all through the host-pointer C-ABI call
    (wall clock, pageable host buffers as a caller has them).  Every figure comes with a parity check against the
    oracle on a sample."""
    import ompl_b200 as ob
    from ompl_b200._lib import check, lib
    from oracle import oracle_py as orc

    L = lib()
    out = {}

    def per_call(h, q, k, reps):
        nq = len(q)
        ib, db, cb = np.empty((nq, k), np.uint64), np.empty((nq, k)), np.empty(nq, np.uint32)
        check(L.omplcu_nn_knn(h, q.ctypes.data, nq, k, ib.ctypes.data, db.ctypes.data, cb.ctypes.data))
        t0 = time.perf_counter()
        for _ in range(reps):
            check(L.omplcu_nn_knn(h, q.ctypes.data, nq, k, ib.ctypes.data, db.ctypes.data, cb.ctypes.data))
        return (time.perf_counter() - t0) / reps, ib, db

    # one nearestK on the metric's own roadmap (1M x 7-D)
    q1 = qs[:1].copy()  # pageable, as a caller has it
    t, ib, db = per_call(nn.handle, q1, K, 200)
    oi, od = orc.knn(orc.space_rv(DIM), pts, q1, K, nthreads=host_threads())
    out["single_nearestK_rv7_1M"] = {"ms_per_call": t * 1e3, "matches_oracle": bool((ib == oi).all() and (db == od).all()),
                                     "kernel": "knn_few_kernel<1> + knn_rerank_select_kernel"}
    # SE(3): 1M states, uniform positions in [-1,1]^3 and uniform rotations
    r = np.random.default_rng(77)

    def se3(m):
        q = r.normal(size=(m, 4))
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        return np.ascontiguousarray(np.hstack([r.uniform(-1, 1, (m, 3)), q]))

    n3, nq3 = 1_000_000, 4096
    p3, q3 = se3(n3), se3(nq3)
    sp = ob.SE3StateSpace(-1.0, 1.0, device=local)
    nn3 = ob.NearestNeighborsCUDA(sp)
    nn3.add_rows(p3)
    t, ib, db = per_call(nn3.handle, q3, K, 5)
    osp = orc.space_se3(-1.0, 1.0)
    oi, od = orc.knn(osp, p3, q3[:32], K, nthreads=host_threads())
    out["knn_se3_1M"] = {"queries_per_s": nq3 / t, "ms_per_call": t * 1e3, "n_queries": nq3, "k": K,
                         "matches_oracle": bool((ib[:32] == oi).all() and np.allclose(db[:32], od, rtol=4e-16, atol=1e-15)),
                         "filtered": int(nn3.stat("last_filter")),
                         "path": "tensor-core list scan + threshold scan on the Euclidean lower-bound embedding, exact SE(3) "
                                 "distance on the kept slots (DESIGN.md 3.8)"}
    t, ib, db = per_call(nn3.handle, np.ascontiguousarray(q3[:1]), K, 100)
    out["single_nearestK_se3_1M"] = {"ms_per_call": t * 1e3, "matches_oracle": bool((ib == oi[:1]).all()),
                                     "kernel": "exact_few_kernel<1> + exact_few_select_kernel"}
    return out


def sharded_e2e(args, torch, world, pts, qs_all, with_motion):
    """Rank 0 only.  kNN (and motion checks) end to end through the multi-GPU entry points of the library
    with pinned host buffers; wall clock around the synchronous calls, L2 of every GPU flushed between
    timed iterations.  Returns (kNN ms over args.steps, motion e2e dict or None, the neighbour ids)."""
    import ompl_b200 as ob
    from ompl_b200 import workloads
    from ompl_b200._lib import check, lib

    L = lib()
    devs = list(range(world))

    pinned = []

    def pin(a):
        rc = torch.cuda.cudart().cudaHostRegister(a.ctypes.data, a.nbytes, 0)
        assert int(rc) == 0, f"cudaHostRegister failed: {rc}"
        pinned.append(a)

    def flush_all():
        for d in devs:
            check(L.omplcu_device_flush_l2(d))

    def wall(step, steps, warmup):
        for _ in range(warmup):
            flush_all()
            step()
        t = 0.0
        for _ in range(steps):
            flush_all()
            t0 = time.perf_counter()
            step()
            t += time.perf_counter() - t0
        return t * 1e3

    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(DIM, device=0), devices=devs)
    nn.add_rows(pts)
    q = qs_all.copy()  # fresh buffers: parts of the originals are already registered by the per-rank legs
    h_ids = np.empty((N_QUERIES, K), dtype=np.uint64)
    h_d = np.empty((N_QUERIES, K), dtype=np.float64)
    h_cnt = np.empty(N_QUERIES, dtype=np.uint32)
    for a in (q, h_ids, h_d):
        pin(a)

    def step():
        check(L.omplcu_nn_knn(nn.handle, q.ctypes.data, N_QUERIES, K, h_ids.ctypes.data, h_d.ctypes.data,
                              h_cnt.ctypes.data))

    ms = wall(step, args.steps, max(args.warmup, 2))
    del nn
    motion_e2e = None
    if with_motion:
        s1, s2 = workloads.m3_segments(N_SEGMENTS)
        sph, box = workloads.se3_obstacles()
        robot = workloads.se3_robot()
        sp = ob.SE3StateSpace(-1.0, 1.0, device=0)
        chks = [ob.SE3ObstacleValidityChecker(robot, sph, box, device=d) for d in devs]
        hs = (C.c_void_p * world)(*[c.handle.value for c in chks])
        a, b = s1.copy(), s2.copy()
        hv = np.empty(N_SEGMENTS, dtype=np.uint8)
        ht = np.empty(N_SEGMENTS, dtype=np.float64)
        for arr in (a, b, hv, ht):
            pin(arr)

        def mstep():
            check(L.omplcu_check_motions_multi(hs, world, sp.handle, a.ctypes.data, b.ctypes.data, N_SEGMENTS,
                                               hv.ctypes.data, ht.ctypes.data, None))

        mms = wall(mstep, args.steps, max(args.warmup, 2))
        motion_e2e = {"value": N_SEGMENTS / (mms / args.steps * 1e-3), "unit": "segments/s",
                      "h2d_bytes_per_step": int(a.nbytes + b.nbytes), "d2h_bytes_per_step": int(hv.nbytes + ht.nbytes),
                      "path": f"omplcu_check_motions_multi over {world} per-GPU checkers in one process (rank 0), "
                              "segments split by index, host buffers", "valid_segments": int(hv.sum())}
    # the arrays die with this frame: unregister them, or a later allocation that reuses their pages fails
    # to register ("part or all of the requested memory range is already mapped")
    h_ids = h_ids.copy()
    for a in pinned:
        torch.cuda.cudart().cudaHostUnregister(a.ctypes.data)
    return ms, motion_e2e, h_ids


# ---------------------------------------------------------------------------
# second half of BASELINE.json's metric: motion checks/sec (workload M3)
# ---------------------------------------------------------------------------
def motion_block(args, torch, dist, nccl, dev, local, rank, world, timed):
    """1M SE3 segments against 500 spheres + 500 boxes at resolution 0.01, split by
    segment over the ranks (no collective on the data path).  Returns the "motion"
    object of the JSON line (rank 0) or None."""
    import ompl_b200 as ob
    from ompl_b200 import workloads
    from ompl_b200._lib import check, lib

    L = lib()
    s1, s2 = workloads.m3_segments(N_SEGMENTS)
    sph, box = workloads.se3_obstacles()
    robot = workloads.se3_robot()
    from ompl_b200 import sharding

    lo, hi = sharding.shard_range(N_SEGMENTS, rank, world)
    n = hi - lo
    a, b = np.ascontiguousarray(s1[lo:hi]), np.ascontiguousarray(s2[lo:hi])
    sp = ob.SE3StateSpace(-1.0, 1.0, device=local)
    chk = ob.SE3ObstacleValidityChecker(robot, sph, box, device=local)
    da, db = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
    dv = torch.empty(n, dtype=torch.uint8, device=dev)
    dt = torch.empty(n, dtype=torch.float64, device=dev)
    sptr = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    hv = np.empty(n, dtype=np.uint8)
    ht = np.empty(n, dtype=np.float64)
    for arr in (a, b, hv, ht):
        torch.cuda.cudart().cudaHostRegister(arr.ctypes.data, arr.nbytes, 0)

    def step_device():
        check(L.omplcu_check_motions_device(chk.handle, sp.handle, da.data_ptr(), db.data_ptr(), n, dv.data_ptr(),
                                            dt.data_ptr(), None, sptr))

    def step_e2e():
        check(L.omplcu_check_motions(chk.handle, sp.handle, a.ctypes.data, b.ctypes.data, n, hv.ctypes.data,
                                     ht.ctypes.data, None))

    def step_device_bool():  # checkMotion(s1, s2): no lastValid, the states of a segment are visited in spread order
        check(L.omplcu_check_motions_device(chk.handle, sp.handle, da.data_ptr(), db.data_ptr(), n, dvb.data_ptr(),
                                            None, None, sptr))

    dvb = torch.empty(n, dtype=torch.uint8, device=dev)
    bool_ms, _, _ = timed(step_device_bool, args.steps, args.warmup)
    bool_eval = chk.stat("last_states_evaluated")
    l0 = chk.stat("launches")
    dev_ms, _, _ = timed(step_device, args.steps, args.warmup)
    launches = (chk.stat("launches") - l0) * args.steps // (args.steps + args.warmup)
    assert bool((dv == dvb).all()), "checkMotion(s1, s2) and checkMotion(s1, s2, lastValid) disagree"
    e2e_ms = float("nan")
    if world == 1:
        _, e2e_ms, _ = timed(step_e2e, args.steps, max(args.warmup, 1))
        assert (dv.cpu().numpy() == hv).all() and (dt.cpu().numpy() == ht).all(), "device and host entry points disagree"
    else:
        hv[:] = dv.cpu().numpy()  # e2e at N > 1 is measured by rank 0 through omplcu_check_motions_multi
        ht[:] = dt.cpu().numpy()
    tot = torch.tensor([chk.stat("last_states"), chk.stat("last_states_evaluated"), int(hv.sum())],
                       dtype=torch.float64, device=dev)
    for arr in (a, b, hv, ht):
        # registrations outlive the arrays otherwise, and a later numpy allocation that lands on such a range makes
        # cudaMemcpyAsync fail with "invalid argument"
        torch.cuda.cudart().cudaHostUnregister(arr.ctypes.data)
    if world > 1:
        dist.all_reduce(tot, group=nccl)
    if rank != 0:
        return None
    states, evaluated, nvalid = (int(x) for x in tot.tolist())
    t_step = dev_ms / args.steps
    # 8 flops per (robot sphere, obstacle sphere) test, 14 per box test, every state of every segment
    alg = float(states) * len(robot) * (8.0 * len(sph) + 14.0 * len(box))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    # algorithmic bytes: both endpoints in (2 x 56 B), valid flag + lastValid time out (9 B) per segment
    alg_bytes = float(n) * (2 * 7 * 8 + 9)
    gbs = alg_bytes / (t_step * 1e-3) / 1e9
    out = {
        "metric": MOTION_METRIC,
        "value": N_SEGMENTS / (t_step * 1e-3), "unit": "segments/s", "ms_per_step": t_step, "scaling": "strong",
        "dtype": "f32 broad phase + f32 narrow phase, f64 reference expressions inside the ambiguity band",
        "config": {"workload": "M3: SE3 [-1,1]^3, uniform endpoints, 500 spheres + 500 boxes, 4-sphere robot",
                   "n_segments": N_SEGMENTS, "states_in_segments": states, "states_evaluated": evaluated,
                   "valid_segments": nvalid, "lanes_per_segment": chk.stat("last_group"),
                   "grid_cells": chk.stat("grid_cells"), "grid_items": chk.stat("grid_items"),
                   "sharding": f"segments split over {world} rank(s), no collective" if world > 1 else "single GPU"},
        "e2e": {"value": N_SEGMENTS / (e2e_ms / args.steps * 1e-3), "unit": "segments/s",
                "h2d_bytes_per_step": int(a.nbytes + b.nbytes), "d2h_bytes_per_step": int(hv.nbytes + ht.nbytes)},
        "gpu_launches": int(launches),
        "bool_overload": {"what": "checkMotion(s1, s2) without lastValid on the same segments (per rank, device-timed): "
                                  "any failing state decides, so the states are visited in spread order",
                          "value": n / (bool_ms / args.steps * 1e-3), "unit": "segments/s",
                          "ms_per_step": bool_ms / args.steps, "states_evaluated": int(bool_eval),
                          "equal_to_lastValid_overload": True},
        "roofline": {
            "bound": "hbm", "kernel": "motion_group_kernel<SE3, 8, lanes_per_segment, SE(3)-specialised>", "peak": hbm_peak, "unit": "GB/s",
            "achieved": gbs, "frac": gbs / hbm_peak, "traffic": ncu_traffic(MOTION_CAPTURE) if world == 1 else None,
            "traffic_unit": f"bytes/launch (ncu, profiles/{MOTION_CAPTURE})",
            "brute_force_equivalent_tflops": alg / (t_step * 1e-3) / 1e12,
            "note": "per rank. Inputs are read once (ncu: 132 MB in, 8 MB out per 1M segments), so HBM is the only "
                    "hardware roof the algorithmic figure can be put against, and the kernel is nowhere near it: "
                    "it is instruction-issue / latency bound (ncu, capture at 8 lanes per segment: issue slots 65 % "
                    "busy, 700 M warp instructions, 16.5 of 32 lanes active; by source line "
                    "profiles/r2_motion_v6_source_lines.txt: narrow phase 48 %, cell walk 17 %, interpolation 14 %). "
                    "brute_force_equivalent_tflops is SURVEY 8(d)'s flop count (every state "
                    "x every robot/obstacle pair) over the time; the grid, the sphere filter and the early exit avoid "
                    "almost all of it, so it exceeds the FP32 pipe peak by construction.",
        },
    }
    if world == 1 and not args.no_cpu:
        cores = host_threads()
        m = min(N_SEGMENTS, 12000 * cores)
        secs, ov, ot = cpu_motion_baseline(a, b, robot, sph, box, cores, m)
        ok = bool((ov == hv[:m].astype(bool)).all() and (ot[~ov] == ht[:m][~ov]).all())
        out["cpu_baseline"] = {"value": m / secs, "unit": "segments/s", "cores": cores, "kind": "port",
                               "sample": f"first {m} of the 1M segments, {cores} threads, " + MOTION_CPU_WHAT,
                               "matches_gpu": ok}
    return out


def run_m4(args):
    """Second driver-visible line: kNN on RealVectorStateSpace(100), 100K states, 65536 queries, k = 10 (SURVEY 8(d),
    M4 'plain L2 at d = 100'): the shape north_star reserves for tensor cores.  One GPU; same timing rules."""
    import torch

    import ompl_b200 as ob
    from ompl_b200._lib import check, lib

    n, nq, k, d = 100_000, 65536, 10, 100
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device")
    dev = torch.device("cuda", 0)
    L = lib()
    pts, qs = np.random.default_rng(4).random((n, d)), np.random.default_rng(5).random((nq, d))
    nn = ob.NearestNeighborsCUDA(ob.RealVectorStateSpace(d))
    nn.add_rows(pts)
    nn.set_option("profile", 1)
    d_q = torch.from_numpy(qs).to(dev)
    d_ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
    d_d = torch.empty((nq, k), dtype=torch.float64, device=dev)
    h_ids, h_d, h_cnt = np.empty((nq, k), np.uint64), np.empty((nq, k)), np.empty(nq, np.uint32)
    for a in (qs, h_ids, h_d):
        torch.cuda.cudart().cudaHostRegister(a.ctypes.data, a.nbytes, 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    sptr = C.c_void_p(stream.cuda_stream)

    def step_device():
        check(L.omplcu_nn_knn_device(nn.handle, d_q.data_ptr(), nq, k, d_ids.data_ptr(), d_d.data_ptr(), sptr))

    def step_e2e():
        check(L.omplcu_nn_knn(nn.handle, qs.ctypes.data, nq, k, h_ids.ctypes.data, h_d.ctypes.data, h_cnt.ctypes.data))

    def timed(step, steps, warmup):
        for _ in range(warmup):
            flush.fill_(1)
            step()
        torch.cuda.synchronize()
        ms = 0.0
        for _ in range(steps):
            flush.fill_(1)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            step()
            b.record(stream)
            b.synchronize()
            ms += a.elapsed_time(b)
        return ms

    l0 = nn.stat("launches")
    sampler = ClockSampler(0)
    for _ in range(args.warmup):
        step_device()
    sampler.start()
    mark = (nn.stat("scan_ns_total"), nn.stat("scan_count"))  # kernel time: mean over the timed steps only
    dev_ms = timed(step_device, args.steps, 0)
    clocks = sampler.stop()
    launches = (nn.stat("launches") - l0) * args.steps // (args.steps + args.warmup)
    scan_ns = (nn.stat("scan_ns_total") - mark[0]) / max(nn.stat("scan_count") - mark[1], 1)
    e2e_ms = timed(step_e2e, args.steps, max(args.warmup, 1))
    assert (d_ids.cpu().numpy().astype(np.uint64) == h_ids).all()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    tpeak = peaks.get("bf16_tflops", 2250.0)
    kpad = (3 * d + 3 + 15) // 16 * 16  # K issued: ceil((3 d + 3) / 16) tcgen05.mma of K = 16
    pairs = float(nq) * n
    t_step = dev_ms / args.steps
    ach = 3.0 * d * pairs / (scan_ns * 1e-9) / 1e12
    out = {
        "metric": "kNN queries/sec (k=10, 100K states, d=100)", "value": nq / (t_step * 1e-3), "unit": UNIT, "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16x2 split filter on tensor cores (f32 accumulate) + f64 exact re-rank", "data": "synthetic",
        "config": {"workload": "M4 plain L2: RealVectorStateSpace(100), 100K states, 64K queries, k=10",
                   "n_states": n, "n_queries": nq, "k": k, "dim": d},
        "l2": "flushed between timed iterations (256 MB write)", "clocks": clocks,
        "e2e": {"value": nq / (e2e_ms / args.steps * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(qs.nbytes),
                "d2h_bytes_per_step": int(h_ids.nbytes + h_d.nbytes)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": "knn_scan_tc_kernel<16,1,TS,PAIR> (query tile in TMEM, CTA pairs share "
                                                  "the point chunks by TMA multicast)", "achieved": ach,
                     "peak": tpeak, "unit": "TFLOP/s", "frac": ach / tpeak,
                     "traffic": ncu_traffic("r2_scan_tc_d100_ncu_summary.csv"),
                     "traffic_unit": "bytes/launch (profiles/r2_scan_tc_d100_ncu_summary.csv)",
                     "flops_per_pair_algorithmic": 3 * d, "flops_per_pair_executed": 2 * kpad,
                     "frac_executed": 2.0 * kpad * pairs / (scan_ns * 1e-9) / 1e12 / tpeak,
                     "kernel_ms": scan_ns * 1e-6, "kernel_share_of_step": scan_ns * 1e-6 / t_step,
                     "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst)" if "bf16_tflops" in peaks
                     else "fallback 2250 TFLOP/s",
                     "frac_executed_of_sustained_peak": (2.0 * kpad * pairs / (scan_ns * 1e-9) / 1e12 /
                                                         peaks["bf16_tflops_sustained"])
                     if "bf16_tflops_sustained" in peaks else None,
                     "note": "tensor-bound: without the selection the same MMAs take 2.6 of the 3.1 ms (profiles/"
                             "r2_experiments.md), the rate of the sustained cuBLAS figure of MEASURED_PEAKS.json (the chip "
                             "runs at its power cap under dense BF16 MMA); ncu: tensor pipe 63 % of cycles at the capture's "
                             "clocks, DRAM 0.3 GB per launch"},
        "exact_fallback_queries": int(nn.stat("last_fallbacks")),
    }
    if not args.no_cpu:
        from oracle import oracle_py as orc

        cores = host_threads()
        m = 64 * cores
        t0 = time.perf_counter()
        oi, od = orc.knn(orc.space_rv(d), pts, qs[:m], k, nthreads=cores)
        secs = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": m / secs, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"first {m} of the 65536 queries, {cores} threads, NearestNeighborsLinear "
                                         "restatement (oracle port)",
                               "matches_gpu": bool((h_ids[:m] == oi).all() and (h_d[:m] == od).all())}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scan-ctas", type=int, default=0, help="tuning: cap resident scan CTAs per SM")
    ap.add_argument("--no-sample-thr", action="store_true", help="tuning: disable the threshold pre-pass")
    ap.add_argument("--shard", default="grid", choices=["grid", "states", "single"],
                    help="N > 1: 'grid' = state shards x query blocks (sharding.grid_shape), 'states' = N state shards")
    ap.add_argument("--states", type=int, default=0, help="roadmap size for scaling studies (default: the metric's 1M)")
    ap.add_argument("--fp32-scan", action="store_true", help="use the FP32-pipe scan kernel instead of the tensor-core scan")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs (profiling runs)")
    ap.add_argument("--no-motion", action="store_true", help="skip the motion-validity block (profiling runs)")
    ap.add_argument("--workload", default="m1", choices=["m1", "m4"],
                    help="m1: the metric's workload (default). m4: SURVEY 8(d) 'plain L2 at d = 100' (100K states, 65536 "
                         "queries, k = 10), the dense-contraction shape of the tensor-core path; same JSON contract, N = 1")
    args = ap.parse_args()
    if args.states:
        global N_STATES
        N_STATES = args.states  # scaling studies on larger roadmaps; the default run is BASELINE's 1M states
    if args.workload == "m4":
        run_m4(args)
    elif args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
