"""omplcu_check_motions_device_noalloc: the allocation-free variant of omplcu_check_motions_device
(scratch in a file-scope static __device__ array of libomplcu.so, see include/omplcu.h).  It has no
counterpart in the reference.

The file name sorts last on purpose: a failure here must not hide
any other test.  The test passes alone in a fresh process and after the rest of the suite.

Bar, as in tests/test_validity_gpu.py: validity booleans, lastValid.second and lastValid.first `==`
(bit-exact) between the two entry points, on two calls in a row (the static array keeps the first
call's contents; the second call has different segments and a different count).
"""
import ctypes as C

import numpy as np
import pytest

import ompl_b200 as ob
from ompl_b200._lib import OmplCuError, check, lib

pytestmark = pytest.mark.gpu


def _se3_world(seed):
    r = np.random.default_rng(seed)
    sph = np.hstack([r.uniform(-1, 1, size=(60, 3)), r.uniform(0.02, 0.08, size=(60, 1))])
    box = np.hstack([r.uniform(-1, 1, size=(60, 3)), r.uniform(0.02, 0.08, size=(60, 3))])
    robot = np.array([[0.0, 0.0, 0.0, 0.03], [0.05, 0.0, 0.0, 0.02], [0.0, -0.04, 0.01, 0.02]])
    return robot, sph, box


def _se3_states(r, n, short):
    q = r.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    a = np.hstack([r.uniform(-1, 1, size=(n, 3)), q])
    if not short:
        q2 = r.normal(size=(n, 4))
        q2 /= np.linalg.norm(q2, axis=1, keepdims=True)
        return a, np.hstack([r.uniform(-1, 1, size=(n, 3)), q2])
    # short edges: many valid segments, so CODE_VALID and lastValid = 1.0 are exercised too
    q2 = q + 0.05 * r.normal(size=(n, 4))
    q2 /= np.linalg.norm(q2, axis=1, keepdims=True)
    return a, np.hstack([a[:, :3] + 0.03 * r.normal(size=(n, 3)), q2])


def _both(L, chk, sp, a, b, stride, with_state=True):
    import torch

    n = len(a)
    dev = torch.device("cuda:0")
    da, db = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
    sptr = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out = []
    for fn in (L.omplcu_check_motions_device, L.omplcu_check_motions_device_noalloc):
        # poisoned outputs: every element must be written by the call
        dv = torch.full((n,), 7, dtype=torch.uint8, device=dev)
        dt = torch.full((n,), -5.0, dtype=torch.float64, device=dev)
        ds = torch.full((n, stride), -9.0, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        check(fn(chk.handle, sp.handle, da.data_ptr(), db.data_ptr(), n, dv.data_ptr(), dt.data_ptr(),
                 ds.data_ptr() if with_state else None, sptr))
        out.append((dv.cpu().numpy(), dt.cpu().numpy(), ds.cpu().numpy()))
    (v0, t0, s0), (v1, t1, s1) = out
    assert set(np.unique(v1)) <= {0, 1}
    assert (v0 == v1).all(), f"validity differs in {(v0 != v1).sum()} of {n} segments"
    assert (t0 == t1).all(), "lastValid.second differs"
    if with_state:
        bad = v1 == 0
        assert (s0[bad] == s1[bad]).all(), "lastValid.first differs"
    return v1


def test_noalloc_motion_check_equals_the_allocating_entry_point_twice_in_a_row(gpu):
    L = lib()
    robot, sph, box = _se3_world(11)
    se3 = ob.SE3StateSpace(-1.0, 1.0)
    chk = ob.SE3ObstacleValidityChecker(robot, sph, box)
    r = np.random.default_rng(5)
    # call 1: long random segments (nearly all invalid), a count that is not a multiple of anything
    a, b = _se3_states(r, 50001, short=False)
    v = _both(L, chk, se3, a, b, 7)
    assert 0 < int((v == 0).sum())
    # call 2, straight after, fewer and different segments: the array still holds call 1's contents
    a, b = _se3_states(r, 20011, short=True)
    v = _both(L, chk, se3, a, b, 7)
    assert 0 < int(v.sum()) < len(v), "the second batch should mix valid and invalid segments"
    # a second checker of another kind on the same device shares the same array
    hc = ob.RealVectorStateSpace(8)
    hc.setLongestValidSegmentFraction(0.001)
    cube = ob.HypercubeValidityChecker(8, 0.1)
    pa, pb = r.random((30000, 8)), r.random((30000, 8))
    pb[:15000] = pa[:15000] + 0.002 * r.normal(size=(15000, 8))
    _both(L, cube, hc, pa, pb, 8)
    # and the bool-only overload (no lastValid outputs)
    import torch

    dev = torch.device("cuda:0")
    da, db = torch.from_numpy(pa).to(dev), torch.from_numpy(pb).to(dev)
    dv0 = torch.empty(len(pa), dtype=torch.uint8, device=dev)
    dv1 = torch.empty(len(pa), dtype=torch.uint8, device=dev)
    sptr = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    check(L.omplcu_check_motions_device(cube.handle, hc.handle, da.data_ptr(), db.data_ptr(), len(pa),
                                        dv0.data_ptr(), None, None, sptr))
    check(L.omplcu_check_motions_device_noalloc(cube.handle, hc.handle, da.data_ptr(), db.data_ptr(), len(pa),
                                                dv1.data_ptr(), None, None, sptr))
    assert (dv0.cpu().numpy() == dv1.cpu().numpy()).all()


def test_noalloc_motion_check_rejects_batches_beyond_the_limit(gpu):
    import torch

    L = lib()
    n = 262144 + 1  # OMPLCU_NOALLOC_MAX_SEGMENTS + 1
    sp = ob.RealVectorStateSpace(2)
    cube = ob.HypercubeValidityChecker(2, 0.1)
    dev = torch.device("cuda:0")
    da = torch.zeros((n, 2), dtype=torch.float64, device=dev)
    dv = torch.zeros(n, dtype=torch.uint8, device=dev)
    sptr = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    with pytest.raises(OmplCuError, match="OMPLCU_NOALLOC_MAX_SEGMENTS"):
        check(L.omplcu_check_motions_device_noalloc(cube.handle, sp.handle, da.data_ptr(), da.data_ptr(), n,
                                                    dv.data_ptr(), None, None, sptr))
    # exactly the limit is accepted, and n == 0 is a no-op
    check(L.omplcu_check_motions_device_noalloc(cube.handle, sp.handle, da.data_ptr(), da.data_ptr(), n - 1,
                                                dv.data_ptr(), None, None, sptr))
    check(L.omplcu_check_motions_device_noalloc(cube.handle, sp.handle, da.data_ptr(), da.data_ptr(), 0,
                                                dv.data_ptr(), None, None, sptr))
