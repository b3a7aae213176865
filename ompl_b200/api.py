"""Host-side mirror of the reference interfaces for the hot path, over the C ABI.

The reference is C++; its toolchain (Boost, Eigen) is absent here, so the C++
adapters in include/ompl_b200/*.h are what an OMPL build links, and this module
is the same surface in Python for the tests and the benchmark: same names,
argument meaning and error behaviour as

  ompl::NearestNeighbors<_T>          datastructures/NearestNeighbors.h:46-115
  base::StateSpace (distance / interpolate / validSegmentCount / setup)
                                      base/src/StateSpace.cpp:237-248,828-854,996-1003,1068-1116
  base::StateValidityChecker          base/StateValidityChecker.h:89-182
  base::MotionValidator / DiscreteMotionValidator
                                      base/MotionValidator.h:64-143, base/src/DiscreteMotionValidator.cpp:48-145

States are rows of float64 in StateSpace::copyToReals order
(base/src/StateSpace.cpp:329-335); SO3 is (x, y, z, w).
"""
import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import Component, OmplCuError, check, lib

PI = math.pi  # == boost::math::constants::pi<double>() as a double
DBL_EPS = np.finfo(np.float64).eps


class Exception_(RuntimeError):
    """ompl::Exception (util/Exception.h)."""


def _rows(a, stride, name="states"):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(1, -1) if stride != 1 or a.size == 1 else a.reshape(-1, 1)
    if a.ndim != 2 or a.shape[1] != stride:
        raise ValueError(f"{name}: expected rows of {stride} doubles, got shape {a.shape}")
    return a


def _ptr(a):
    return a.ctypes.data if a is not None else None


# ----------------------------------------------------------------------
# state spaces
# ----------------------------------------------------------------------
class StateSpace:
    """A (possibly compound) state space described to the device by its
    components; longestValidSegment_ follows StateSpace::setup()."""

    def __init__(self, comps, compound, device=0):
        # comps: list of dict(type, dim, weight, extent, param)
        self._comps = comps
        self._compound = bool(compound)
        self._fraction = 0.01  # StateSpace.cpp:92
        self._factor = 1  # StateSpace.cpp:94
        self._handle = None
        self.device = device
        self.stride = sum(c["dim"] for c in comps)

    # -- reference setters --
    def setLongestValidSegmentFraction(self, f):
        # StateSpace.cpp:828-834
        if f < DBL_EPS or f > 1.0 - DBL_EPS:
            raise Exception_("The fraction of the extent must be larger than 0 and less than 1")
        self._fraction = float(f)
        self._drop()

    def setValidSegmentCountFactor(self, factor):
        # StateSpace.cpp:820-826 (a compound space does NOT forward this to its
        # components; CompoundStateSpace::validSegmentCount uses the components' own)
        if factor < 1:
            raise Exception_(
                "The multiplicative factor for the valid segment count between two states must be strictly positive"
            )
        self._factor = int(factor)
        self._drop()

    def getLongestValidSegmentFraction(self):
        return self._fraction

    def getValidSegmentCountFactor(self):
        return self._factor

    def getDimensionValues(self):
        return self.stride

    def getMaximumExtent(self):
        if not self._compound:
            return self._comps[0]["extent"]
        # CompoundStateSpace::getMaximumExtent, StateSpace.cpp:996-1003
        e = 0.0
        for c in self._comps:
            if c["weight"] >= DBL_EPS:
                e += c["weight"] * c["extent"]
        return e

    def getLongestValidSegmentLength(self):
        return self.getMaximumExtent() * self._fraction

    def components(self):
        """omplcu_component array as setup() would leave the reference objects."""
        arr = (Component * len(self._comps))()
        for i, c in enumerate(self._comps):
            arr[i].type = c["type"]
            arr[i].dim = c["dim"]
            arr[i].weight = c["weight"]
            arr[i].lvs = c["extent"] * self._fraction  # StateSpace.cpp:239-240 per (sub)space
            # a plain space uses its own factor; components of a compound keep the default 1
            arr[i].factor = self._factor if not self._compound else 1
            arr[i].param = c.get("param", 0.0)
        return arr

    def _drop(self):
        if self._handle is not None:
            lib().omplcu_space_destroy(self._handle)
            self._handle = None

    @property
    def handle(self):
        if self._handle is None:
            arr = self.components()
            for c in arr:
                if c.lvs < DBL_EPS:  # StateSpace.cpp:242-248
                    raise Exception_("The longest valid segment for state space must be positive.")
            h = C.c_void_p()
            check(lib().omplcu_space_create(arr, len(self._comps), int(self._compound), C.byref(h)))
            self._handle = h
        return self._handle

    def __del__(self):
        try:
            self._drop()
        except Exception:
            pass

    # -- batched math on the device --
    def distance(self, a, b):
        a, b = _rows(a, self.stride), _rows(b, self.stride)
        out = np.empty(len(a), dtype=np.float64)
        check(lib().omplcu_space_distance(self.handle, self.device, _ptr(a), _ptr(b), len(a), _ptr(out)))
        return out

    def interpolate(self, a, b, t):
        a, b = _rows(a, self.stride), _rows(b, self.stride)
        t = np.ascontiguousarray(np.broadcast_to(np.asarray(t, dtype=np.float64), (len(a),)))
        out = np.empty_like(a)
        check(lib().omplcu_space_interpolate(self.handle, self.device, _ptr(a), _ptr(b), _ptr(t), len(a), _ptr(out)))
        return out

    def validSegmentCount(self, a, b):
        a, b = _rows(a, self.stride), _rows(b, self.stride)
        out = np.empty(len(a), dtype=np.uint32)
        check(lib().omplcu_space_segment_count(self.handle, self.device, _ptr(a), _ptr(b), len(a), _ptr(out)))
        return out


def _rv_extent(low, high):
    # RealVectorStateSpace::getMaximumExtent, RealVectorStateSpace.cpp:167-176
    e = 0.0
    for lo, hi in zip(low, high):
        d = float(hi) - float(lo)
        e += d * d
    return math.sqrt(e)


def _bounds(dim, low, high):
    low = [float(low)] * dim if np.isscalar(low) else [float(x) for x in low]
    high = [float(high)] * dim if np.isscalar(high) else [float(x) for x in high]
    if len(low) != dim or len(high) != dim:
        raise Exception_("Bounds do not match dimension of state space")
    return low, high


def RealVectorStateSpace(dim, low=0.0, high=1.0, device=0):
    low, high = _bounds(dim, low, high)
    sp = StateSpace([dict(type=_lib.RV, dim=dim, weight=1.0, extent=_rv_extent(low, high))], False, device)
    sp.low, sp.high = low, high
    return sp


def SO2StateSpace(device=0):
    return StateSpace([dict(type=_lib.SO2, dim=1, weight=1.0, extent=PI)], False, device)


def SO3StateSpace(device=0):
    return StateSpace([dict(type=_lib.SO3, dim=4, weight=1.0, extent=0.5 * PI)], False, device)


def SE2StateSpace(low=0.0, high=1.0, device=0):
    # R^2 weight 1.0 + SO2 weight 0.5, base/spaces/SE2StateSpace.h:108-112
    low, high = _bounds(2, low, high)
    return StateSpace(
        [
            dict(type=_lib.RV, dim=2, weight=1.0, extent=_rv_extent(low, high)),
            dict(type=_lib.SO2, dim=1, weight=0.5, extent=PI),
        ],
        True,
        device,
    )


def SE3StateSpace(low=0.0, high=1.0, device=0):
    # R^3 weight 1.0 + SO3 weight 1.0, base/spaces/SE3StateSpace.h constructor
    low, high = _bounds(3, low, high)
    return StateSpace(
        [
            dict(type=_lib.RV, dim=3, weight=1.0, extent=_rv_extent(low, high)),
            dict(type=_lib.SO3, dim=4, weight=1.0, extent=0.5 * PI),
        ],
        True,
        device,
    )


def KinematicChainSpace(links, link_length, device=0):
    # demos/KinematicChain.h:88-99: RealVectorStateSpace(links), bounds [-pi, pi]
    ext = _rv_extent([-PI] * links, [PI] * links)
    return StateSpace([dict(type=_lib.CHAIN, dim=links, weight=1.0, extent=ext, param=float(link_length))], False,
                      device)


def CompoundStateSpace(parts, weights, device=0):
    comps = []
    for p, w in zip(parts, weights):
        if p._compound or len(p._comps) != 1:
            raise Exception_("nested compound spaces are flattened by the caller")
        c = dict(p._comps[0])
        c["weight"] = float(w)
        comps.append(c)
    return StateSpace(comps, True, device)


# ----------------------------------------------------------------------
# nearest neighbours
# ----------------------------------------------------------------------
class NearestNeighborsCUDA:
    """ompl::NearestNeighbors<_T> with the element set mirrored on the GPU.

    Elements are arbitrary hashable Python objects; `state_accessor(elem)`
    returns the element's state row (the reference hands the structure only an
    opaque distance function; a GPU structure needs coordinates).
    Results are NearestNeighborsLinear's with ties broken by insertion order.
    """

    def __init__(self, space, state_accessor=None, device=None, devices=None):
        """`devices` (a list of GPU indices) shards the states over several GPUs of this process
        (omplcu_nn_create_sharded); results are identical to the single-device structure."""
        self.space = space
        self.device = space.device if device is None else device
        self.devices = None if devices is None else [int(d) for d in devices]
        self._accessor = state_accessor or (lambda e: e)
        self._dist_fun = None
        self._elem = {}  # id -> element
        self._ids_of = {}  # element key -> [ids] (insertion order)
        h = C.c_void_p()
        if self.devices is None:
            check(lib().omplcu_nn_create(space.handle, self.device, C.byref(h)))
        else:
            arr = (C.c_int * len(self.devices))(*self.devices)
            check(lib().omplcu_nn_create_sharded(space.handle, arr, len(self.devices), C.byref(h)))
            self.device = self.devices[0]
        self._h = h

    def __del__(self):
        try:
            if getattr(self, "_h", None) is not None:
                lib().omplcu_nn_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # -- reference interface --
    def setDistanceFunction(self, f):
        self._dist_fun = f  # kept for getDistanceFunction(); queries use the space's metric

    def getDistanceFunction(self):
        return self._dist_fun

    def reportsSortedResults(self):
        return True

    def clear(self):
        check(lib().omplcu_nn_clear(self._h))
        self._elem.clear()
        self._ids_of.clear()

    @staticmethod
    def _key(e):
        try:
            hash(e)
            return e
        except TypeError:
            return id(e)

    def add(self, data):
        """add(const _T&) or add(const std::vector<_T>&)."""
        elems = data if isinstance(data, (list, tuple)) and not self._looks_like_state(data) else [data]
        if not elems:
            return
        rows = _rows(np.array([np.asarray(self._accessor(e), dtype=np.float64) for e in elems]), self.space.stride)
        first = C.c_uint64()
        check(lib().omplcu_nn_add(self._h, _ptr(rows), len(rows), C.byref(first)))
        for i, e in enumerate(elems):
            self._elem[first.value + i] = e
            self._ids_of.setdefault(self._key(e), []).append(first.value + i)

    def _looks_like_state(self, data):
        return len(data) > 0 and np.isscalar(data[0])

    def add_rows(self, rows, elems=None):
        """Bulk add of an (n x stride) array; elements default to their ids."""
        rows = _rows(rows, self.space.stride)
        first = C.c_uint64()
        check(lib().omplcu_nn_add(self._h, _ptr(rows), len(rows), C.byref(first)))
        if elems is not None:
            for i, e in enumerate(elems):
                self._elem[first.value + i] = e
                self._ids_of.setdefault(self._key(e), []).append(first.value + i)
        return first.value

    def remove(self, data):
        # NearestNeighborsLinear::remove erases the LAST equal element (:86-96)
        ids = self._ids_of.get(self._key(data))
        if not ids:
            return False
        i = ids.pop()
        if not ids:
            del self._ids_of[self._key(data)]
        del self._elem[i]
        rc = lib().omplcu_nn_remove(self._h, i)
        if rc == _lib.ERR_NOT_FOUND:
            return False
        check(rc)
        return True

    def remove_id(self, i):
        rc = lib().omplcu_nn_remove(self._h, int(i))
        if rc == _lib.ERR_NOT_FOUND:
            return False
        check(rc)
        e = self._elem.pop(int(i), None)
        if e is not None:
            ids = self._ids_of.get(self._key(e), [])
            if int(i) in ids:
                ids.remove(int(i))
        return True

    def remove_ids(self, ids):
        """Bulk removal by id (one kernel, one synchronisation); returns how many were present."""
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        n = C.c_size_t()
        check(lib().omplcu_nn_remove_batch(self._h, _ptr(ids), len(ids), C.byref(n)))
        for i in ids.tolist():
            e = self._elem.pop(i, None)
            if e is not None:
                lst = self._ids_of.get(self._key(e), [])
                if i in lst:
                    lst.remove(i)
        return n.value

    def size(self):
        n = C.c_size_t()
        check(lib().omplcu_nn_size(self._h, C.byref(n)))
        return n.value

    def list_ids(self):
        n = self.size()
        out = np.empty(n, dtype=np.uint64)
        got = C.c_size_t()
        check(lib().omplcu_nn_list(self._h, _ptr(out), n, C.byref(got)))
        return out[: got.value]

    def get_row(self, i):
        """Stored coordinates of the element with id `i` (omplcu_nn_get)."""
        row = np.empty(self.space.stride, dtype=np.float64)
        check(lib().omplcu_nn_get(self._h, int(i), _ptr(row)))
        return row

    def list(self):
        return [self._elem.get(int(i), int(i)) for i in self.list_ids()]

    def _elems(self, ids):
        return [self._elem.get(int(i), int(i)) for i in ids]

    def nearest(self, data):
        if self.size() == 0:
            # NearestNeighborsLinear.h:115
            raise Exception_("No elements found in nearest neighbors data structure")
        ids, _ = self.nearestKBatch(np.asarray(self._accessor(data), dtype=np.float64).reshape(1, -1), 1)
        return self._elems(ids[0])[0]

    def nearestK(self, data, k):
        q = np.asarray(self._accessor(data), dtype=np.float64).reshape(1, -1)
        ids, _ = self.nearestKBatch(q, k)
        return self._elems(ids[0])

    def nearestR(self, data, radius):
        q = np.asarray(self._accessor(data), dtype=np.float64).reshape(1, -1)
        ids, _, offs = self.nearestRBatch(q, [radius])
        return self._elems(ids[offs[0]:offs[1]])

    # -- batch-submit entry points (what PRM/FMT*/BIT* call sites use) --
    def nearestKBatch(self, queries, k, return_dists=True):
        """ids (nq x min(k,size)) uint64 and exact distances, rows sorted by (distance, id)."""
        q = _rows(queries, self.space.stride, "queries")
        nq, k = len(q), int(k)
        ids = np.empty((nq, k), dtype=np.uint64)
        d = np.empty((nq, k), dtype=np.float64) if return_dists else None
        cnt = np.empty(nq, dtype=np.uint32)
        check(lib().omplcu_nn_knn(self._h, _ptr(q), nq, k, _ptr(ids), _ptr(d), _ptr(cnt)))
        kk = int(cnt[0]) if nq else 0
        return ids[:, :kk], (d[:, :kk] if d is not None else None)

    def nearestRBatch(self, queries, radii):
        q = _rows(queries, self.space.stride, "queries")
        r = np.ascontiguousarray(np.broadcast_to(np.asarray(radii, dtype=np.float64), (len(q),)))
        offs = np.zeros(len(q) + 1, dtype=np.uint64)
        pi, pd = C.c_void_p(), C.c_void_p()
        check(lib().omplcu_nn_radius(self._h, _ptr(q), len(q), _ptr(r), C.byref(pi), C.byref(pd), _ptr(offs)))
        tot = int(offs[-1])
        try:
            if tot == 0 or not pi.value:
                ids, d = np.empty(0, dtype=np.uint64), np.empty(0, dtype=np.float64)
            else:
                ids = np.ctypeslib.as_array(C.cast(pi, C.POINTER(C.c_uint64)), shape=(tot,)).copy()
                d = np.ctypeslib.as_array(C.cast(pd, C.POINTER(C.c_double)), shape=(tot,)).copy()
        finally:
            lib().omplcu_free(pi)
            lib().omplcu_free(pd)
        return ids, d, offs.astype(np.int64)

    def set_option(self, key, value):
        check(lib().omplcu_nn_set_option(self._h, key.encode(), int(value)))

    def stat(self, key):
        v = C.c_long()
        check(lib().omplcu_nn_get_stat(self._h, key.encode(), C.byref(v)))
        return v.value

    def set_id_base(self, base):
        check(lib().omplcu_nn_set_id_base(self._h, int(base)))

    @property
    def handle(self):
        return self._h


# ----------------------------------------------------------------------
# validity
# ----------------------------------------------------------------------
class StateValidityChecker:
    """base::StateValidityChecker with a batched isValid on the device."""

    def __init__(self, handle, stride_hint=None, device=0):
        self._h = handle
        self.device = device
        self._stride = stride_hint

    def __del__(self):
        try:
            if getattr(self, "_h", None) is not None:
                lib().omplcu_checker_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def isValid(self, state):
        s = np.asarray(state, dtype=np.float64).reshape(1, -1)
        return bool(self.isValidBatch(s)[0])

    def isValidBatch(self, states):
        s = np.ascontiguousarray(states, dtype=np.float64)
        if s.ndim != 2:
            raise ValueError("states must be (n x stride)")
        out = np.empty(len(s), dtype=np.uint8)
        check(lib().omplcu_check_states(self._h, s.shape[1], _ptr(s), len(s), _ptr(out)))
        return out.astype(bool)

    def stat(self, key):
        v = C.c_long()
        check(lib().omplcu_checker_get_stat(self._h, key.encode(), C.byref(v)))
        return v.value


def AllValidStateValidityChecker(device=0):
    h = C.c_void_p()
    check(lib().omplcu_checker_create_allvalid(device, C.byref(h)))
    return StateValidityChecker(h, device=device)


def HypercubeValidityChecker(dim, edge_width=0.1, device=0):
    """demos/HypercubeBenchmark.cpp:51-72."""
    h = C.c_void_p()
    check(lib().omplcu_checker_create_hypercube(device, int(dim), float(edge_width), C.byref(h)))
    return StateValidityChecker(h, dim, device)


def KinematicChainValidityChecker(links, link_length, env_segments, device=0):
    """demos/KinematicChain.h:193-276; env_segments is (n x 4): x0,y0,x1,y1."""
    env = np.ascontiguousarray(env_segments, dtype=np.float64).reshape(-1, 4)
    h = C.c_void_p()
    check(lib().omplcu_checker_create_chain(device, int(links), float(link_length), _ptr(env), len(env), C.byref(h)))
    return StateValidityChecker(h, links, device)


def SE3ObstacleValidityChecker(robot_spheres, sphere_obstacles, box_obstacles, device=0):
    """Rigid body of spheres vs sphere / axis-aligned-box obstacles."""
    rb = np.ascontiguousarray(robot_spheres, dtype=np.float64).reshape(-1, 4)
    sp = np.ascontiguousarray(sphere_obstacles, dtype=np.float64).reshape(-1, 4)
    bx = np.ascontiguousarray(box_obstacles, dtype=np.float64).reshape(-1, 6)
    h = C.c_void_p()
    check(lib().omplcu_checker_create_se3(device, _ptr(rb), len(rb), _ptr(sp) if len(sp) else None, len(sp),
                                          _ptr(bx) if len(bx) else None, len(bx), C.byref(h)))
    return StateValidityChecker(h, 7, device)


def Circles2DValidityChecker(circles, device=0):
    """tests/resources/circles2D.h:139-149; circles is (n x 3): x, y, r."""
    c = np.ascontiguousarray(circles, dtype=np.float64).reshape(-1, 3)
    h = C.c_void_p()
    check(lib().omplcu_checker_create_circles(device, _ptr(c) if len(c) else None, len(c), C.byref(h)))
    return StateValidityChecker(h, 2, device)


class BatchedDiscreteMotionValidator:
    """base::DiscreteMotionValidator with a batched checkMotion on the device.

    Keeps the valid_/invalid_ counters of base::MotionValidator
    (base/MotionValidator.h:102-139): one count per segment, as
    DiscreteMotionValidator.cpp:139-142 does."""

    def __init__(self, space, checker, checkers=None):
        """`checkers`: one checker per GPU with the same obstacle set; batches are then split by
        segment over them (omplcu_check_motions_multi), results unchanged."""
        self.space = space
        self.checker = checker if checker is not None else checkers[0]
        self.checkers = list(checkers) if checkers else None
        self.valid_ = 0
        self.invalid_ = 0

    def checkMotion(self, s1, s2, lastValid=None):
        """checkMotion(s1, s2) or checkMotion(s1, s2, lastValid).  `lastValid`
        is a 2-item list [state_or_None, t] updated in place on failure."""
        a = np.asarray(s1, dtype=np.float64).reshape(1, -1)
        b = np.asarray(s2, dtype=np.float64).reshape(1, -1)
        if lastValid is None:
            return bool(self.checkMotionBatch(a, b)[0])
        want_state = lastValid[0] is not None
        v, t, st = self.checkMotionBatch(a, b, last_valid=True, last_state=want_state)
        if not v[0]:
            lastValid[1] = float(t[0])
            if want_state:
                lastValid[0] = st[0].copy()
        return bool(v[0])

    def checkMotionBatch(self, s1, s2, last_valid=False, last_state=False):
        a, b = _rows(s1, self.space.stride), _rows(s2, self.space.stride)
        n = len(a)
        v = np.empty(n, dtype=np.uint8)
        t = np.empty(n, dtype=np.float64) if last_valid else None
        st = np.zeros_like(a) if (last_valid and last_state) else None
        if self.checkers:
            hs = (C.c_void_p * len(self.checkers))(*[c.handle.value for c in self.checkers])
            check(lib().omplcu_check_motions_multi(hs, len(self.checkers), self.space.handle, _ptr(a), _ptr(b), n,
                                                   _ptr(v), _ptr(t), _ptr(st)))
        else:
            check(lib().omplcu_check_motions(self.checker.handle, self.space.handle, _ptr(a), _ptr(b), n, _ptr(v),
                                             _ptr(t), _ptr(st)))
        nv = int(v.sum())
        self.valid_ += nv
        self.invalid_ += n - nv
        vb = v.astype(bool)
        if last_valid:
            return vb, t, st
        return vb

    def getValidMotionCount(self):
        return self.valid_

    def getInvalidMotionCount(self):
        return self.invalid_

    def getCheckedMotionCount(self):
        return self.valid_ + self.invalid_

    def resetMotionCounter(self):
        self.valid_ = self.invalid_ = 0
