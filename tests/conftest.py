import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference compiled into oracle/_ref (SSE2 build) — the parity checker."""
    from oracle import refpy
    if not refpy.available("sse2"):
        pytest.skip("oracle/_ref/libjplref_sse2.so not built (needs /root/reference at build time)")
    return refpy.Ref("sse2")


@pytest.fixture(scope="session")
def gpu_ctx():
    import jplspatial_b200 as J
    if J.device_count() == 0:
        pytest.fail("GPU test selected but no CUDA device is visible: the product has no CPU fallback")
    return J.Context(0)


@pytest.fixture(scope="session")
def host_ctx():
    import jplspatial_b200 as J
    return J.Context(-1)
