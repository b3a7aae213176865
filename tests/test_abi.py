"""The C-ABI library loads and exports every symbol include/omplcu.h declares.
No compute calls: this runs without a GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "omplcu.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(omplcu_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    names = declared_symbols()
    for need in ("omplcu_nn_add", "omplcu_nn_remove", "omplcu_nn_knn", "omplcu_nn_radius", "omplcu_nn_list",
                 "omplcu_check_states", "omplcu_check_motions", "omplcu_space_create", "omplcu_last_error"):
        assert need in names


def test_library_exports_every_declared_symbol():
    from ompl_b200 import _lib

    assert os.path.exists(_lib.LIB_PATH), "libomplcu.so not built (run __graft_entry__.build())"
    raw = C.CDLL(_lib.LIB_PATH)
    missing = [n for n in declared_symbols() if not hasattr(raw, n)]
    assert not missing, f"declared in omplcu.h but not exported: {missing}"


def test_python_binding_covers_the_header():
    from ompl_b200 import _lib

    bound = set(_lib.SIGNATURES) | {"omplcu_last_error", "omplcu_free"}
    assert set(declared_symbols()) <= bound


def test_no_cpu_fallback_when_library_missing(monkeypatch):
    from ompl_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libomplcu.so")
    with pytest.raises(ImportError):
        _lib.lib()


def test_product_never_touches_the_oracle():
    """ompl_b200/ must not import, link or name the oracle."""
    pkg = os.path.join(ROOT, "ompl_b200")
    for dp, _, fs in os.walk(pkg):
        if os.path.basename(dp) in ("build", "lib", "__pycache__"):
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                for ln in open(os.path.join(dp, f), errors="replace"):
                    code = ln.split("//")[0].split("#", 1)[0] if not ln.lstrip().startswith("#include") else ln
                    uses = re.search(r"(#include|import|CDLL|dlopen|-l|\.so).*(oracle|ompl_ref)", code)
                    assert not uses, (dp, f, ln)


def test_status_codes_without_gpu_are_errors_not_fallbacks():
    """Without a device, create calls must fail with a CUDA status (never compute on the CPU)."""
    from ompl_b200 import _lib

    l = _lib.lib()
    n = C.c_int(-1)
    rc = l.omplcu_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a GPU is present")
    comp = (_lib.Component * 1)()
    comp[0].type, comp[0].dim, comp[0].weight, comp[0].lvs, comp[0].factor = _lib.RV, 3, 1.0, 0.01, 1
    sp = C.c_void_p()
    assert l.omplcu_space_create(comp, 1, 0, C.byref(sp)) == 0  # host-only object
    h = C.c_void_p()
    rc = l.omplcu_nn_create(sp, 0, C.byref(h))
    assert rc != 0 and l.omplcu_last_error()
    l.omplcu_space_destroy(sp)
