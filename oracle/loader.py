"""Loads the CPU oracle / host-compiled reference.  TEST INFRASTRUCTURE ONLY: importable
from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def build_oracle():
    subprocess.run(["make", "-C", HERE, "-s"], check=True)


def load_oracle(build=True):
    so = os.path.join(HERE, "libphoton_oracle.so")
    if build and (not os.path.exists(so)
                  or os.path.getmtime(so) < os.path.getmtime(os.path.join(HERE, "photon_oracle.cpp"))):
        build_oracle()
    lib = ctypes.CDLL(so)
    lib.photon_oracle_last_ray_steps.restype = ctypes.c_longlong
    lib.photon_oracle_last_rays.restype = ctypes.c_longlong
    lib.photon_oracle_last_trace_seconds.restype = ctypes.c_double
    lib.photon_oracle_last_threads.restype = ctypes.c_int
    return lib


def ref_cpu_path():
    return os.path.join(HERE, "_ref", "libphoton_ref_cpu.so")


def load_ref_cpu():
    """The reference's own sources compiled for the host (oracle/build_ref.sh); None if absent."""
    p = ref_cpu_path()
    return ctypes.CDLL(p) if os.path.exists(p) else None
