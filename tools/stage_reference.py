#!/usr/bin/env python
"""Run the reference's UNCHANGED Python scripts (sample_run_script.sh piv|bos ->
batch_run_simulation.py -> run_simulation_02.py -> perform_ray_tracing_03.py) against a
libparallel_ray_tracing.so of our choosing, from a writable stage under /tmp.

The scripts are copied, never edited.  What the stage adds:
  * cuda_codes/Debug/libparallel_ray_tracing.so = tools/record_stub.c, which records the 29
    arguments of every start_ray_tracing call and forwards to --forward <lib.so> (the CUDA
    library on a GPU box, the CPU oracle here) or, with no --forward, sets one pixel;
  * a sitecustomize.py that restores the numpy aliases the 2017 scripts use (np.NAN ...)
    and pins np.random.seed(None) so that recordings are reproducible;
  * LD_LIBRARY_PATH (pr03.py:1885 reads it).

usage: python tools/stage_reference.py {piv|bos|bos_rotated} --out DIR [--forward LIB] [--reference /root/reference]

kind bos_rotated = sample-data/bos/sample-parameters.mat (camera turned 150 degrees about y, Euler,
thin lens, 512 x 512, one point per dot on a regular grid) -- the shipped parameter file whose rays DO cross
sample-density.nrrd.  Its two absolute /scratch paths are re-pointed into the stage (density file, image
directory) and the empty camera_design.pixel_gain is filled from bos_pattern.pixel_gain; every other field is
written back as loaded.
Needs /root/reference, so it runs in the build container; the recordings it writes are what
travels (tests/golden/recorded_*.npz, see photon_b200/recording.py).
"""
import argparse
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SHIM = r'''
import numpy as np
for _n, _v in (("NAN", np.nan), ("NaN", np.nan), ("infty", np.inf), ("Inf", np.inf), ("float", float),
               ("int", int), ("object", object), ("bool", bool)):
    if not hasattr(np, _n):
        try:
            setattr(np, _n, _v)
        except Exception:
            pass
try:
    np.__all__ = [n for n in np.__all__ if n not in ("max", "min", "round")]   # bhmie.py: from numpy import *
except Exception:
    pass
_seed = np.random.seed
_count = [0]
def _pinned_seed(s=None, *a, **k):
    if s is None:
        _count[0] += 1
        s = 20170518 + _count[0]
    return _seed(s, *a, **k)
np.random.seed = _pinned_seed
'''


def rotated_parameter_file(bos_dir):
    """sample-data/bos/sample-parameters.mat with its two /scratch paths re-pointed into the stage."""
    import numpy as np
    import scipy.io as sio

    def plain(v):
        if hasattr(v, "_fieldnames"):
            return {k: plain(getattr(v, k)) for k in v._fieldnames}
        return v
    m = sio.loadmat(os.path.join(bos_dir, "sample-parameters.mat"), squeeze_me=True, struct_as_record=False)
    m = {k: plain(v) for k, v in m.items() if not k.startswith("__")}
    m["density_gradients"]["density_gradient_filename"] = "../sample-data/bos/sample-density.nrrd"
    m["output_data"]["image_directory"] = "../sample-data/bos/images_rotated/"
    # camera_design.pixel_gain is an EMPTY array in the shipped file (pr03.py:1859 cannot convert it); the file keeps the
    # gain under bos_pattern (25 dB, the value of the other sample file).  It only scales the 16-bit export.
    if np.size(m["camera_design"]["pixel_gain"]) == 0:
        m["camera_design"]["pixel_gain"] = float(m["bos_pattern"]["pixel_gain"])
    os.makedirs(os.path.join(bos_dir, "parameters_rotated"), exist_ok=True)
    os.makedirs(os.path.join(bos_dir, "images_rotated"), exist_ok=True)
    sio.savemat(os.path.join(bos_dir, "parameters_rotated", "sample-parameters.mat"), m, format="5", long_field_names=True)


def stage(kind, out, forward, reference):
    st = os.path.join(out, "stage")
    if os.path.exists(st):
        shutil.rmtree(st)
    os.makedirs(os.path.join(st, "cuda_codes", "Debug"))
    os.makedirs(os.path.join(st, "cuda_codes", "data"))
    shutil.copytree(os.path.join(reference, "python_codes"), os.path.join(st, "python_codes"),
                    ignore=shutil.ignore_patterns("*.pyc", "__pycache__"))
    shutil.copytree(os.path.join(reference, "sample-data"), os.path.join(st, "sample-data"))
    with open(os.path.join(st, "sitecustomize.py"), "w") as f:
        f.write(SHIM)
    stub = os.path.join(st, "cuda_codes", "Debug", "libparallel_ray_tracing.so")
    subprocess.run(["gcc", "-O1", "-fPIC", "-shared", "-o", stub, os.path.join(ROOT, "tools", "record_stub.c"), "-ldl"],
                   check=True)
    env = dict(os.environ)
    env["PYTHONPATH"] = st + os.pathsep + env.get("PYTHONPATH", "")
    env["LD_LIBRARY_PATH"] = env.get("LD_LIBRARY_PATH", "") or "/usr/lib"
    env["PHOTON_RECORD_DIR"] = os.path.join(out, "recording_" + kind)
    env["PHOTON_RECORD_FORWARD"] = os.path.abspath(forward) if forward else ""
    env["MPLBACKEND"] = "Agg"
    if os.path.exists(env["PHOTON_RECORD_DIR"]):
        shutil.rmtree(env["PHOTON_RECORD_DIR"])
    # sample_run_script.sh <kind>  ==  python3 batch_run_simulation.py ../sample-data/<kind>/parameters 1 1
    pdir = "../sample-data/%s/parameters" % kind
    if kind == "bos_rotated":
        pdir = "../sample-data/bos/parameters_rotated"
        rotated_parameter_file(os.path.join(st, "sample-data", "bos"))
    cmd = [sys.executable, "batch_run_simulation.py", pdir, "1", "1"]
    r = subprocess.run(cmd, cwd=os.path.join(st, "python_codes"), env=env, capture_output=True, text=True)
    with open(os.path.join(out, "run_%s.log" % kind), "w") as f:
        f.write(r.stdout + "\n--- stderr ---\n" + r.stderr)
    return r.returncode, env["PHOTON_RECORD_DIR"], st


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["piv", "bos", "bos_rotated"])
    ap.add_argument("--out", required=True)
    ap.add_argument("--forward", default="")
    ap.add_argument("--reference", default="/root/reference")
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    rc, rec, st = stage(a.kind, a.out, a.forward, a.reference)
    print("exit code", rc, "| recording:", rec, "| stage:", st, "| log:", os.path.join(a.out, "run_%s.log" % a.kind))
    sys.exit(rc)


if __name__ == "__main__":
    main()
