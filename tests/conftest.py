import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


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
