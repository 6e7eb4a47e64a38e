"""Builds photon_b200/lib/libparallel_ray_tracing.so with nvcc for sm_100a (in-tree)."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libparallel_ray_tracing.so")
SOURCES = ["driver.cu"]
HEADERS = ["vecmath.cuh", "ray_stages.cuh", "ray_optics.cuh", "trace_core.cuh", "trace_kernel.cuh",
           "scene_args.h", "nrrd_reader.h", "../../include/photon_abi.h", "../../include/photon_b200.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "-fmad=false",     # FMA only where the code asks for it (vecmath.cuh)
]


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found")
    return p


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False, defines=(), out=None):
    """Compile the CUDA library if it is missing or older than its sources.  ``defines``/``out`` build
    an experimental variant beside it (tools/gpu_sweep.sh selects one with PHOTON_B200_LIB)."""
    if not force and not is_stale() and out is None:
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    flags = list(NVCC_FLAGS) + ["-D" + d for d in defines]
    out = out or LIB
    cmd = [nvcc_path()] + flags + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", out] + [os.path.join(CSRC, s) for s in SOURCES] + ["-lz"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return out


if __name__ == "__main__":
    print(build(force=True, verbose=True))
