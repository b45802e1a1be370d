import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_api():
    from oracle import pyoracle
    if not os.path.exists(pyoracle.PORT_LIB) and not os.path.exists(pyoracle.REF_LIB):
        pyoracle.build()
    return pyoracle.load()


def _build_hostsim():
    """g++ build of the kernels' per-thread bodies (tests/hostsim) — test harness only."""
    import ctypes
    d = os.path.join(ROOT, "tests", "hostsim")
    so = os.path.join(d, "libhostsim.so")
    csrc = os.path.join(ROOT, "pupiloptixlab_b200", "csrc")
    srcs = [os.path.join(d, "hostsim.cpp")] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cuh", ".h"))]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        fma = ["-mfma"] if "fma" in open("/proc/cpuinfo").read() else []
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", *fma, "-shared",
                               "-Wno-unknown-pragmas", *os.environ.get("HS_EXTRA", "").split(), "-o", so, os.path.join(d, "hostsim.cpp")])
    return ctypes.CDLL(so)


@pytest.fixture(scope="session")
def hostsim_lib():
    return _build_hostsim()
