import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.join(ROOT, "tests") not in sys.path:
    sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Build the product library and the oracle once per session."""
    import fcntl
    import __graft_entry__ as ge
    with open(os.path.join(ROOT, "tests", ".build.lock"), "w") as lock:   # one pytest-xdist worker builds, the others wait
        fcntl.flock(lock, fcntl.LOCK_EX)
        ge.build()
    return True


@pytest.fixture(scope="session")
def emu_lib(built):
    """TEST-ONLY host harness of the CUDA kernels (tests/emu): same .cu sources
    compiled with g++ and executed thread by thread."""
    from regcm_b200 import capi
    d = os.path.join(ROOT, "tests", "emu")
    import fcntl
    with open(os.path.join(d, ".build.lock"), "w") as lock:   # pytest-xdist workers would otherwise rebuild it at once
        fcntl.flock(lock, fcntl.LOCK_EX)
        subprocess.check_call(["make", "-s", "-C", d])
    return capi.CoreLib(os.path.join(d, "libregcm_dyn_emu.so"), "rcm_dyn_")


@pytest.fixture(scope="session")
def cuda_lib(built):
    from regcm_b200 import capi
    return capi.cuda_lib()
