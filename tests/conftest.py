import ctypes
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle.loader import load_oracle
    return load_oracle()


@pytest.fixture(scope="session")
def hostcheck_lib():
    """The product's per-ray stage code compiled for the host (tests/hostcheck)."""
    src = os.path.join(ROOT, "tests", "hostcheck", "hostcheck.cpp")
    out = os.path.join(ROOT, "tests", "hostcheck", "libhostcheck.so")
    csrc = os.path.join(ROOT, "photon_b200", "csrc")
    newest = max(os.path.getmtime(os.path.join(csrc, f)) for f in os.listdir(csrc))
    newest = max(newest, os.path.getmtime(src))
    if not os.path.exists(out) or os.path.getmtime(out) < newest:
        gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.run([gxx, "-O2", "-std=c++17", "-DPH_NO_FMA", "-ffp-contract=off", "-fPIC", "-shared", "-x", "c++", "-o", out, src, "-lz"], check=True)
    return ctypes.CDLL(out)


@pytest.fixture(scope="session")
def cuda_lib():
    from photon_b200.lib import load_library
    return load_library(require_gpu=True)
