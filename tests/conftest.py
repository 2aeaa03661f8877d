import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_device_count():
    """Devices the CUDA runtime sees (0 without a driver). Asked of libcudart directly: a missing or stale libb200vr.so must
    stay a test failure, never turn into a skip."""
    import ctypes
    for name in ("libcudart.so", "libcudart.so.12", "/usr/local/cuda/lib64/libcudart.so"):
        try:
            rt = ctypes.CDLL(name)
        except OSError:
            continue
        n = ctypes.c_int(0)
        return n.value if rt.cudaGetDeviceCount(ctypes.byref(n)) == 0 else 0
    return 0


def pytest_collection_modifyitems(config, items):
    if any("gpu" in it.keywords for it in items) and _cuda_device_count() == 0:
        skip = pytest.mark.skip(reason="no CUDA device on this machine (the B200 volume device has no CPU fallback)")
        for it in items:
            if "gpu" in it.keywords:
                it.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the oracle (test infrastructure) and, when nvcc is here and the .so is stale, the CUDA library."""
    from oracle import pyoracle
    pyoracle.build()
    import shutil
    if shutil.which("nvcc"):
        import ovr_b200
        ovr_b200.capi.build()
    yield
