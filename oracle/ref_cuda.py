"""Oracle B -- the reference's OWN CUDA path (oracle/_ref/libphoton_ref_cuda.so, built by oracle/build_ref_cuda.sh
from the sources under /root/reference with texture references patched to texture objects) run on the GPU box.
TEST INFRASTRUCTURE ONLY: importable from tests/ and from bench.py's cpu_baseline / --impl reference legs.

Every call runs in its OWN subprocess: the reference writes image[(ii-1)*nx + jj-1] with ii or jj == 0 (an
out-of-bounds store) and calls exit() on an NRRD failure -- neither may take the test process'
CUDA context with it.  The call travels as an .npz (photon_b200.recording), the outputs come back as files.
"""
import glob
import os
import subprocess
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def ref_cuda_path():
    return os.path.join(HERE, "_ref", "libphoton_ref_cuda.so")


def available():
    return os.path.exists(ref_cuda_path())


def run(call, outdir, dumps=True, timeout=900, env=None):
    """-> dict(image, pos, dir, wall_s, call_s).  pos/dir are None when dumps is False."""
    from photon_b200 import recording
    os.makedirs(outdir, exist_ok=True)
    for f in glob.glob(os.path.join(outdir, "*.bin")) + glob.glob(os.path.join(outdir, "*.npy")):
        os.remove(f)
    import copy
    c = copy.copy(call)
    c.save_lightrays = bool(dumps)
    c.lightray_position_save_path = outdir
    c.lightray_direction_save_path = outdir
    spn = min(c.source_point_number, c.num_particles)
    c.num_lightrays_save = spn * c.lightray_number_per_particle if dumps else 0
    npz = os.path.join(outdir, "call.npz")
    recording.save_npz(npz, c)
    e = dict(os.environ)
    e.update(env or {})
    e["PYTHONPATH"] = ROOT + os.pathsep + e.get("PYTHONPATH", "")
    t0 = time.perf_counter()
    r = subprocess.run([sys.executable, "-m", "oracle.ref_cuda", npz, outdir], cwd=ROOT, env=e, capture_output=True,
                       text=True, timeout=timeout)
    wall = time.perf_counter() - t0
    if r.returncode != 0 or not os.path.exists(os.path.join(outdir, "image.npy")):
        raise RuntimeError("reference CUDA path failed (rc %d):\n%s\n%s" % (r.returncode, r.stdout[-3000:], r.stderr[-3000:]))
    out = {"image": np.load(os.path.join(outdir, "image.npy")), "pos": None, "dir": None, "wall_s": wall,
           "call_s": float(np.load(os.path.join(outdir, "call_seconds.npy")))}
    if dumps:
        out["pos"] = np.concatenate([np.fromfile(f, np.float32).reshape(-1, 3)
                                     for f in sorted(glob.glob(os.path.join(outdir, "pos_*.bin")))])
        out["dir"] = np.concatenate([np.fromfile(f, np.float32).reshape(-1, 3)
                                     for f in sorted(glob.glob(os.path.join(outdir, "dir_*.bin")))])
    return out


def _worker(npz, outdir):
    import ctypes
    from photon_b200 import recording
    call, _ = recording.load_npz(npz)
    lib = ctypes.CDLL(ref_cuda_path())
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    sys.stdout.flush()
    os.dup2(devnull, 1)                       # the reference prints progress lines on stdout
    try:
        t0 = time.perf_counter()
        img = call.run(lib)
        dt = time.perf_counter() - t0
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    np.save(os.path.join(outdir, "call_seconds.npy"), np.float64(dt))
    np.save(os.path.join(outdir, "image.npy"), img)


if __name__ == "__main__":
    _worker(sys.argv[1], sys.argv[2])
