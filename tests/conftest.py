import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
HOSTSHIM = os.path.join(ROOT, "tests", "hostshim")  # phh_* facade + ctypes binding that let pytest drive the C++ mirror
if HOSTSHIM not in sys.path:
    sys.path.insert(0, HOSTSHIM)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle

    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def emu_lib():
    """The product .cu sources compiled with g++ against tests/emu's CUDA-runtime shim (test harness only)."""
    import __graft_entry__ as g
    from phylo_rs_b200 import capi

    alt = os.environ.get("PHB_EMU_LIB")  # e.g. an AddressSanitizer build of the same sources (LD_PRELOAD libasan)
    return capi.Library(alt if alt else g.build_emu())


@pytest.fixture(scope="session")
def gpu_lib():
    """The real sm_100a library; the gpu tests call through its C ABI."""
    from phylo_rs_b200 import capi

    return capi.default_library()
