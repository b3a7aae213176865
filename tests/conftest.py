import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import ctypes as C

        from ompl_b200 import _lib

        n = C.c_int(0)
        return _lib.lib().omplcu_device_count(C.byref(n)) == 0 and n.value > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu():
    """GPU tests must fail loudly (not skip) when the extension is missing on a
    GPU box; without a device they cannot run at all."""
    from ompl_b200 import _lib

    _lib.lib()  # ImportError if libomplcu.so was not built
    if not _have_gpu():
        pytest.fail("no CUDA device visible: -m gpu tests need a GPU")
    return 0
