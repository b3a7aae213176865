"""ctypes binding of the parity oracle (oracle/libompl_oracle.so) and, when it
was built, of the compiled reference headers (oracle/_ref/libompl_ref.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs, never by ompl_b200.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libompl_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libompl_ref.so")

RV, SO2, SO3, CHAIN = 0, 1, 2, 3
CHK_ALLVALID, CHK_HYPERCUBE, CHK_CHAIN, CHK_SE3, CHK_CIRCLES = range(5)
MAX_COMP = 8


class OrcComponent(C.Structure):
    _fields_ = [("type", C.c_int), ("dim", C.c_int), ("weight", C.c_double), ("lvs", C.c_double),
                ("factor", C.c_uint), ("param", C.c_double)]


class OrcSpace(C.Structure):
    _fields_ = [("ncomp", C.c_int), ("compound", C.c_int), ("stride", C.c_int), ("comp", OrcComponent * MAX_COMP)]


class OrcChecker(C.Structure):
    _fields_ = [
        ("kind", C.c_int), ("dim", C.c_int), ("edge_width", C.c_double),
        ("nlinks", C.c_int), ("link_length", C.c_double), ("nenv", C.c_int), ("env", C.c_void_p),
        ("nrobot", C.c_int), ("robot", C.c_void_p), ("nsph", C.c_int), ("sph", C.c_void_p),
        ("nbox", C.c_int), ("box", C.c_void_p), ("ncirc", C.c_int), ("circ", C.c_void_p),
    ]


_o = None
_r = None


def oracle():
    global _o
    if _o is None:
        if not os.path.exists(ORACLE_SO):
            raise ImportError(f"{ORACLE_SO} missing: run `make -C oracle`")
        o = C.CDLL(ORACLE_SO)
        o.orc_distance.restype = C.c_double
        o.orc_distance.argtypes = [C.POINTER(OrcSpace), C.c_void_p, C.c_void_p]
        o.orc_interpolate.restype = None
        o.orc_interpolate.argtypes = [C.POINTER(OrcSpace), C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
        o.orc_valid_segment_count.restype = C.c_uint
        o.orc_valid_segment_count.argtypes = [C.POINTER(OrcSpace), C.c_void_p, C.c_void_p]
        o.orc_is_valid.restype = C.c_int
        o.orc_is_valid.argtypes = [C.POINTER(OrcChecker), C.c_void_p]
        o.orc_check_motion.restype = C.c_int
        o.orc_check_motion.argtypes = [C.POINTER(OrcSpace), C.POINTER(OrcChecker), C.c_void_p, C.c_void_p]
        o.orc_check_motion_last.restype = C.c_int
        o.orc_check_motion_last.argtypes = [C.POINTER(OrcSpace), C.POINTER(OrcChecker), C.c_void_p, C.c_void_p,
                                            C.POINTER(C.c_double), C.c_void_p]
        o.orc_check_states_batch.restype = None
        o.orc_check_states_batch.argtypes = [C.POINTER(OrcChecker), C.c_int, C.c_void_p, C.c_size_t, C.c_void_p,
                                             C.c_int]
        o.orc_check_motions_batch.restype = None
        o.orc_check_motions_batch.argtypes = [C.POINTER(OrcSpace), C.POINTER(OrcChecker), C.c_void_p, C.c_void_p,
                                              C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]
        o.orc_knn.restype = C.c_size_t
        o.orc_knn.argtypes = [C.POINTER(OrcSpace), C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p,
                              C.c_void_p]
        o.orc_radius.restype = C.c_size_t
        o.orc_radius.argtypes = [C.POINTER(OrcSpace), C.c_void_p, C.c_size_t, C.c_void_p, C.c_double, C.c_void_p,
                                 C.c_void_p, C.c_size_t]
        o.orc_knn_batch.restype = None
        o.orc_knn_batch.argtypes = [C.POINTER(OrcSpace), C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t,
                                    C.c_void_p, C.c_void_p, C.c_int]
        o.orc_horn_environment.restype = C.c_int
        o.orc_horn_environment.argtypes = [C.c_uint, C.c_double, C.c_void_p]
        o.orc_rv_max_extent.restype = C.c_double
        o.orc_rv_max_extent.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        for name, args in {
            "orc_space_rv": [C.POINTER(OrcSpace), C.c_int, C.c_void_p, C.c_void_p, C.c_double],
            "orc_space_so2": [C.POINTER(OrcSpace), C.c_double],
            "orc_space_so3": [C.POINTER(OrcSpace), C.c_double],
            "orc_space_se2": [C.POINTER(OrcSpace), C.c_void_p, C.c_void_p, C.c_double],
            "orc_space_se3": [C.POINTER(OrcSpace), C.c_void_p, C.c_void_p, C.c_double],
            "orc_space_chain": [C.POINTER(OrcSpace), C.c_int, C.c_double, C.c_double],
        }.items():
            getattr(o, name).restype = None
            getattr(o, name).argtypes = args
        _o = o
    return _o


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _r
    if _r is None:
        r = C.CDLL(REF_SO)
        r.ref_nn_create.restype = C.c_void_p
        r.ref_nn_create.argtypes = [C.c_int, C.POINTER(OrcSpace)]
        r.ref_nn_destroy.restype = None
        r.ref_nn_destroy.argtypes = [C.c_void_p]
        r.ref_nn_add.restype = None
        r.ref_nn_add.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]
        r.ref_nn_remove.restype = C.c_int
        r.ref_nn_remove.argtypes = [C.c_void_p, C.c_ulong]
        r.ref_nn_size.restype = C.c_size_t
        r.ref_nn_size.argtypes = [C.c_void_p]
        r.ref_nn_list.restype = C.c_size_t
        r.ref_nn_list.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        r.ref_nn_knn.restype = C.c_double
        r.ref_nn_knn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_int]
        r.ref_nn_radius.restype = C.c_size_t
        r.ref_nn_radius.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_size_t]
        r.ref_nn_nearest.restype = C.c_long
        r.ref_nn_nearest.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
        _r = r
    return _r


def _arr(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a):
    return a.ctypes.data if a is not None else None


# ---- spaces ----
def space_rv(d, low=0.0, high=1.0, fraction=0.01):
    sp = OrcSpace()
    lo = _arr(np.broadcast_to(low, (d,)))
    hi = _arr(np.broadcast_to(high, (d,)))
    oracle().orc_space_rv(C.byref(sp), d, _p(lo), _p(hi), fraction)
    return sp


def space_so2(fraction=0.01):
    sp = OrcSpace()
    oracle().orc_space_so2(C.byref(sp), fraction)
    return sp


def space_so3(fraction=0.01):
    sp = OrcSpace()
    oracle().orc_space_so3(C.byref(sp), fraction)
    return sp


def space_se2(low=0.0, high=1.0, fraction=0.01):
    sp = OrcSpace()
    lo, hi = _arr(np.broadcast_to(low, (2,))), _arr(np.broadcast_to(high, (2,)))
    oracle().orc_space_se2(C.byref(sp), _p(lo), _p(hi), fraction)
    return sp


def space_se3(low=0.0, high=1.0, fraction=0.01):
    sp = OrcSpace()
    lo, hi = _arr(np.broadcast_to(low, (3,))), _arr(np.broadcast_to(high, (3,)))
    oracle().orc_space_se3(C.byref(sp), _p(lo), _p(hi), fraction)
    return sp


def space_chain(links, link_length, fraction=0.01):
    sp = OrcSpace()
    oracle().orc_space_chain(C.byref(sp), links, link_length, fraction)
    return sp


# ---- math ----
def distance(sp, a, b):
    a, b = _arr(a).reshape(-1, sp.stride), _arr(b).reshape(-1, sp.stride)
    o = oracle()
    return np.array([o.orc_distance(C.byref(sp), _p(a[i]), _p(b[i])) for i in range(len(a))])


def interpolate(sp, a, b, t):
    a, b = _arr(a).reshape(-1, sp.stride), _arr(b).reshape(-1, sp.stride)
    t = np.broadcast_to(np.asarray(t, dtype=np.float64), (len(a),))
    out = np.empty_like(a)
    o = oracle()
    for i in range(len(a)):
        o.orc_interpolate(C.byref(sp), _p(a[i]), _p(b[i]), float(t[i]), _p(out[i]))
    return out


def segment_count(sp, a, b):
    a, b = _arr(a).reshape(-1, sp.stride), _arr(b).reshape(-1, sp.stride)
    o = oracle()
    return np.array([o.orc_valid_segment_count(C.byref(sp), _p(a[i]), _p(b[i])) for i in range(len(a))],
                    dtype=np.uint32)


# ---- checkers (keep the numpy arrays alive on the returned object) ----
class Checker:
    def __init__(self, c, keep):
        self.c = c
        self._keep = keep


def checker_allvalid():
    c = OrcChecker()
    c.kind = CHK_ALLVALID
    return Checker(c, [])


def checker_hypercube(dim, edge_width=0.1):
    c = OrcChecker()
    c.kind, c.dim, c.edge_width = CHK_HYPERCUBE, dim, edge_width
    return Checker(c, [])


def horn_environment(d, eps):
    segs = np.zeros((2 * (d - 1), 4))
    n = oracle().orc_horn_environment(d, eps, _p(segs))
    return segs[:n]


def checker_chain(links, link_length, env):
    env = _arr(env).reshape(-1, 4)
    c = OrcChecker()
    c.kind, c.nlinks, c.link_length, c.nenv, c.env = CHK_CHAIN, links, link_length, len(env), _p(env)
    return Checker(c, [env])


def checker_se3(robot, sph, box):
    robot, sph, box = _arr(robot).reshape(-1, 4), _arr(sph).reshape(-1, 4), _arr(box).reshape(-1, 6)
    c = OrcChecker()
    c.kind = CHK_SE3
    c.nrobot, c.robot, c.nsph, c.sph, c.nbox, c.box = len(robot), _p(robot), len(sph), _p(sph), len(box), _p(box)
    return Checker(c, [robot, sph, box])


def checker_circles(circ):
    circ = _arr(circ).reshape(-1, 3)
    c = OrcChecker()
    c.kind, c.ncirc, c.circ = CHK_CIRCLES, len(circ), _p(circ)
    return Checker(c, [circ])


def check_states(chk, states, nthreads=1):
    s = _arr(states)
    out = np.empty(len(s), dtype=np.uint8)
    oracle().orc_check_states_batch(C.byref(chk.c), s.shape[1], _p(s), len(s), _p(out), nthreads)
    return out.astype(bool)


def check_motions(sp, chk, s1, s2, last_valid=False, nthreads=1):
    a, b = _arr(s1).reshape(-1, sp.stride), _arr(s2).reshape(-1, sp.stride)
    v = np.empty(len(a), dtype=np.uint8)
    t = np.empty(len(a), dtype=np.float64) if last_valid else None
    oracle().orc_check_motions_batch(C.byref(sp), C.byref(chk.c), _p(a), _p(b), len(a), _p(v), _p(t), nthreads)
    return (v.astype(bool), t) if last_valid else v.astype(bool)


def check_motion_last(sp, chk, s1, s2):
    """(valid, t, state) of the lastValid overload for one segment."""
    a, b = _arr(s1), _arr(s2)
    t = C.c_double(1.0)
    st = np.zeros(sp.stride)
    ok = oracle().orc_check_motion_last(C.byref(sp), C.byref(chk.c), _p(a), _p(b), C.byref(t), _p(st))
    return bool(ok), t.value, st


# ---- nearest neighbours ----
def knn(sp, pts, queries, k, nthreads=1):
    pts, q = _arr(pts).reshape(-1, sp.stride), _arr(queries).reshape(-1, sp.stride)
    kk = min(k, len(pts))
    idx = np.empty((len(q), kk), dtype=np.uint64)
    d = np.empty((len(q), kk), dtype=np.float64)
    if kk:
        oracle().orc_knn_batch(C.byref(sp), _p(pts), len(pts), _p(q), len(q), k, _p(idx), _p(d), nthreads)
    return idx, d


def radius(sp, pts, query, r):
    pts, q = _arr(pts).reshape(-1, sp.stride), _arr(query)
    idx = np.empty(len(pts), dtype=np.uint64)
    d = np.empty(len(pts), dtype=np.float64)
    n = oracle().orc_radius(C.byref(sp), _p(pts), len(pts), _p(q), float(r), _p(idx), _p(d), len(pts))
    return idx[:n], d[:n]


class RefNN:
    """The reference's own NearestNeighborsLinear / NearestNeighborsGNAT (compiled
    unmodified from /root/reference into oracle/_ref)."""

    LINEAR, GNAT, GNAT_SMALL = 0, 1, 2

    def __init__(self, kind, sp):
        self.sp = sp
        self.h = ref().ref_nn_create(kind, C.byref(sp))

    def __del__(self):
        try:
            ref().ref_nn_destroy(self.h)
        except Exception:
            pass

    def add(self, rows, reserve_query_slots=256):
        rows = _arr(rows).reshape(-1, self.sp.stride)
        ref().ref_nn_add(self.h, _p(rows), len(rows), reserve_query_slots)

    def remove(self, i):
        return bool(ref().ref_nn_remove(self.h, int(i)))

    def size(self):
        return ref().ref_nn_size(self.h)

    def list(self):
        n = self.size()
        out = np.empty(max(n, 1), dtype=np.uint64)
        got = ref().ref_nn_list(self.h, _p(out), n)
        return out[:got]

    def knn(self, queries, k, nthreads=1):
        q = _arr(queries).reshape(-1, self.sp.stride)
        kk = min(k, self.size())
        out = np.empty((len(q), kk), dtype=np.uint64)
        secs = ref().ref_nn_knn(self.h, _p(q), len(q), k, _p(out), nthreads)
        return out, secs

    def radius(self, query, r):
        q = _arr(query)
        out = np.empty(max(self.size(), 1), dtype=np.uint64)
        n = ref().ref_nn_radius(self.h, _p(q), float(r), _p(out), self.size())
        return out[:n]

    def nearest(self, query):
        q = _arr(query)
        err = C.create_string_buffer(256)
        r = ref().ref_nn_nearest(self.h, _p(q), err, 256)
        if r < 0:
            raise RuntimeError(err.value.decode())
        return r


# ---- the reference's own state spaces + DiscreteMotionValidator (oracle/_ref/libompl_refbase.so) ----
REFBASE_SO = os.path.join(HERE, "_ref", "libompl_refbase.so")
_rb = None


def have_refbase():
    return os.path.exists(REFBASE_SO)


def refbase():
    global _rb
    if _rb is None:
        r = C.CDLL(REFBASE_SO)
        r.refb_create.restype = C.c_void_p
        r.refb_create.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_uint]
        r.refb_destroy.restype = None
        r.refb_destroy.argtypes = [C.c_void_p]
        r.refb_stride.restype = C.c_uint
        r.refb_stride.argtypes = [C.c_void_p]
        r.refb_max_extent.restype = C.c_double
        r.refb_max_extent.argtypes = [C.c_void_p]
        r.refb_longest_valid_segment.restype = C.c_double
        r.refb_longest_valid_segment.argtypes = [C.c_void_p]
        r.refb_distance.restype = None
        r.refb_distance.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        r.refb_interpolate.restype = None
        r.refb_interpolate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        r.refb_segment_count.restype = None
        r.refb_segment_count.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        r.refb_check_motions.restype = None
        r.refb_check_motions.argtypes = [C.c_void_p, C.POINTER(OrcChecker), C.c_void_p, C.c_void_p, C.c_size_t,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _rb = r
    return _rb


class RefBase:
    """A reference state space (RealVector / SO2 / SO3 / SE2 / SE3) after setup(), with the
    reference's DiscreteMotionValidator on top; rows are copyToReals order."""

    KINDS = {"rv": 0, "so2": 1, "so3": 2, "se2": 3, "se3": 4}

    def __init__(self, kind, dim=0, low=0.0, high=1.0, fraction=0.01, factor=1):
        nb = {"rv": dim, "se2": 2, "se3": 3}.get(kind, 0)
        lo = _arr(np.broadcast_to(low, (max(nb, 1),)))
        hi = _arr(np.broadcast_to(high, (max(nb, 1),)))
        self.h = refbase().refb_create(self.KINDS[kind], dim, _p(lo), _p(hi), fraction, factor)
        self.stride = refbase().refb_stride(self.h)

    def __del__(self):
        try:
            refbase().refb_destroy(self.h)
        except Exception:
            pass

    def max_extent(self):
        return refbase().refb_max_extent(self.h)

    def longest_valid_segment(self):
        return refbase().refb_longest_valid_segment(self.h)

    def distance(self, a, b):
        a, b = _arr(a).reshape(-1, self.stride), _arr(b).reshape(-1, self.stride)
        out = np.empty(len(a))
        refbase().refb_distance(self.h, _p(a), _p(b), len(a), _p(out))
        return out

    def interpolate(self, a, b, t):
        a, b = _arr(a).reshape(-1, self.stride), _arr(b).reshape(-1, self.stride)
        t = _arr(np.broadcast_to(np.asarray(t, dtype=np.float64), (len(a),)))
        out = np.empty_like(a)
        refbase().refb_interpolate(self.h, _p(a), _p(b), _p(t), len(a), _p(out))
        return out

    def segment_count(self, a, b):
        a, b = _arr(a).reshape(-1, self.stride), _arr(b).reshape(-1, self.stride)
        out = np.empty(len(a), dtype=np.uint32)
        refbase().refb_segment_count(self.h, _p(a), _p(b), len(a), _p(out))
        return out

    def check_motions(self, chk, s1, s2, last_valid=False, last_state=False):
        a, b = _arr(s1).reshape(-1, self.stride), _arr(s2).reshape(-1, self.stride)
        v = np.empty(len(a), dtype=np.uint8)
        t = np.empty(len(a)) if last_valid else None
        st = np.zeros_like(a) if (last_valid and last_state) else None
        cnt = np.zeros(2, dtype=np.uint32)
        refbase().refb_check_motions(self.h, C.byref(chk.c), _p(a), _p(b), len(a), _p(v), _p(t), _p(st), _p(cnt))
        self.counts = cnt
        if last_valid:
            return v.astype(bool), t, st
        return v.astype(bool)
