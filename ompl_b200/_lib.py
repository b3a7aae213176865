"""ctypes binding of libomplcu.so (the C ABI declared in include/omplcu.h).

The library is the product; there is no CPU fallback.  Importing this module
without a built library, or calling a compute entry point without a GPU, raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libomplcu.so")

OK, ERR_INVALID, ERR_CUDA, ERR_NOT_FOUND, ERR_EMPTY, ERR_INTERNAL = range(6)
ID_NONE = 0xFFFFFFFFFFFFFFFF
RV, SO2, SO3, CHAIN = 0, 1, 2, 3


class Component(C.Structure):
    _fields_ = [
        ("type", C.c_int32),
        ("dim", C.c_int32),
        ("weight", C.c_double),
        ("lvs", C.c_double),
        ("factor", C.c_uint32),
        ("_pad", C.c_uint32),
        ("param", C.c_double),
    ]


class OmplCuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


_vp, _sz, _i, _u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64
_pd = C.c_void_p  # double* / uint64_t* etc. passed as raw addresses

# name -> argtypes (all return int unless noted); kept in sync with include/omplcu.h
SIGNATURES = {
    "omplcu_version": [],
    "omplcu_device_count": [C.POINTER(C.c_int)],
    "omplcu_space_create": [C.POINTER(Component), _i, _i, C.POINTER(_vp)],
    "omplcu_space_destroy": [_vp],
    "omplcu_space_stride": [_vp, C.POINTER(C.c_int)],
    "omplcu_space_distance": [_vp, _i, _pd, _pd, _sz, _pd],
    "omplcu_space_interpolate": [_vp, _i, _pd, _pd, _pd, _sz, _pd],
    "omplcu_space_segment_count": [_vp, _i, _pd, _pd, _sz, _pd],
    "omplcu_space_satisfies_bounds": [_vp, _i, _pd, _pd, _pd, _sz, _pd],
    "omplcu_space_maximum_extent": [_vp, _pd, _pd, C.POINTER(C.c_double)],
    "omplcu_nn_create": [_vp, _i, C.POINTER(_vp)],
    "omplcu_nn_create_sharded": [_vp, C.POINTER(C.c_int), _i, C.POINTER(_vp)],
    "omplcu_nn_destroy": [_vp],
    "omplcu_nn_set_id_base": [_vp, _u64],
    "omplcu_nn_add": [_vp, _pd, _sz, C.POINTER(_u64)],
    "omplcu_nn_remove": [_vp, _u64],
    "omplcu_nn_remove_batch": [_vp, _pd, _sz, C.POINTER(_sz)],
    "omplcu_nn_clear": [_vp],
    "omplcu_nn_size": [_vp, C.POINTER(_sz)],
    "omplcu_nn_list": [_vp, _pd, _sz, C.POINTER(_sz)],
    "omplcu_nn_get": [_vp, _u64, _pd],
    "omplcu_nn_knn": [_vp, _pd, _sz, _sz, _pd, _pd, _pd],
    "omplcu_nn_radius": [_vp, _pd, _sz, _pd, C.POINTER(_vp), C.POINTER(_vp), _pd],
    "omplcu_nn_knn_device": [_vp, _pd, _sz, _sz, _pd, _pd, _vp],
    "omplcu_merge_topk_device": [_i, _pd, _pd, _i, _sz, _sz, _pd, _pd, _vp],
    "omplcu_merge_topk_strided_device": [_i, _pd, _pd, _i, _sz, _sz, _sz, _pd, _pd, _vp],
    "omplcu_nn_set_option": [_vp, C.c_char_p, C.c_long],
    "omplcu_nn_get_stat": [_vp, C.c_char_p, C.POINTER(C.c_long)],
    "omplcu_checker_create_allvalid": [_i, C.POINTER(_vp)],
    "omplcu_checker_create_hypercube": [_i, _i, C.c_double, C.POINTER(_vp)],
    "omplcu_checker_create_chain": [_i, _i, C.c_double, _pd, _i, C.POINTER(_vp)],
    "omplcu_checker_create_se3": [_i, _pd, _i, _pd, _i, _pd, _i, C.POINTER(_vp)],
    "omplcu_checker_create_circles": [_i, _pd, _i, C.POINTER(_vp)],
    "omplcu_checker_destroy": [_vp],
    "omplcu_checker_get_stat": [_vp, C.c_char_p, C.POINTER(C.c_long)],
    "omplcu_check_states": [_vp, _i, _pd, _sz, _pd],
    "omplcu_check_motions": [_vp, _vp, _pd, _pd, _sz, _pd, _pd, _pd],
    "omplcu_check_states_multi": [C.POINTER(_vp), _i, _i, _pd, _sz, _pd],
    "omplcu_check_motions_multi": [C.POINTER(_vp), _i, _vp, _pd, _pd, _sz, _pd, _pd, _pd],
    "omplcu_check_states_device": [_vp, _i, _pd, _sz, _pd, _vp],
    "omplcu_check_motions_device": [_vp, _vp, _pd, _pd, _sz, _pd, _pd, _pd, _vp],
    "omplcu_check_motions_device_noalloc": [_vp, _vp, _pd, _pd, _sz, _pd, _pd, _pd, _vp],
    "omplcu_device_flush_l2": [_i],
    "omplcu_measure_fp32_peak": [_i, C.POINTER(C.c_double), C.POINTER(C.c_double)],
}

_lib = None


def lib():
    """Load libomplcu.so (once).  Raises if the CUDA extension was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)"
            )
        l = C.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = C.c_int
        l.omplcu_last_error.restype = C.c_char_p
        l.omplcu_last_error.argtypes = []
        l.omplcu_free.restype = None
        l.omplcu_free.argtypes = [_vp]
        _lib = l
    return _lib


def check(rc):
    if rc != OK:
        raise OmplCuError(rc, lib().omplcu_last_error().decode("utf-8", "replace"))
    return rc
