"""Parity of the trace-kernel variants AT THE BENCHMARK'S SIZE (C3, 512^3 field) against the reference: the host build
(oracle/_ref/libphoton_ref_cpu.so) and, when present, the reference's own CUDA path (oracle/_ref/libphoton_ref_cuda.so).
The reference runs once per process; every variant named on the command line is compared with it.
usage: python tools/parity_probe512.py [--grid 512] [--points 96] [variant ...]      variant = name[:ENV=VAL[,ENV=VAL]]
"""
import argparse, json, os, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from photon_b200 import scenes
from photon_b200.lib import load_library
from tests import util

ap = argparse.ArgumentParser()
ap.add_argument("--grid", type=int, default=512)
ap.add_argument("--points", type=int, default=96)
ap.add_argument("--rays", type=int, default=1000)
ap.add_argument("variants", nargs="*", default=["cells", "coherent:PHOTON_CELLS=0", "fanout:PHOTON_HOT=4", "texture:PHOTON_SAMPLER=texture"])
args = ap.parse_args()
d = os.path.join(tempfile.gettempdir(), "photon_bench"); os.makedirs(d, exist_ok=True)
nrrd = os.path.join(d, f"c3_blob_{args.grid}.nrrd")
call = scenes.c3_bos(nrrd, n=args.grid, dots_per_side=100, points_per_dot=100, rays_per_point=args.rays, write_field=not os.path.exists(nrrd))
stride = max(1, call.num_particles // args.points)
sub = call.subset(np.arange(0, call.num_particles, stride)[:args.points]); sub.source_point_number = args.points
cache = os.path.join(d, f"parity_ref_{args.grid}_{args.points}_{args.rays}.npz")
refs = {}
if os.path.exists(cache):
    z = np.load(cache)
    for k in ("cpu", "cuda"):
        if k + "_img" in z.files:
            refs[k] = (z[k + "_img"], z[k + "_pos"], z[k + "_dir"])
else:
    from oracle.loader import load_ref_cpu
    from oracle import ref_cuda
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    cpu = load_ref_cpu()
    if cpu is not None:
        dd = os.path.join(d, "pp_cpu"); os.makedirs(dd, exist_ok=True)
        devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1); os.dup2(devnull, 1)
        try:
            refs["cpu"] = util.run_with_dumps(cpu, sub, dd)
        finally:
            os.dup2(saved, 1); os.close(devnull); os.close(saved)
    if ref_cuda.available():
        r = ref_cuda.run(sub, os.path.join(d, "pp_cuda"))
        refs["cuda"] = (r["image"], r["pos"], r["dir"])
        print("reference CUDA path: call %.1f s" % r["call_s"], flush=True)
    np.savez(cache, **{k + s: v[i] for k, v in refs.items() for i, s in enumerate(("_img", "_pos", "_dir"))})
if "cpu" in refs and "cuda" in refs:
    print("reference CUDA vs reference host build:", json.dumps(util.compare(refs["cpu"], refs["cuda"], 10.0)), flush=True)
lib = load_library(require_gpu=True)
for v in args.variants:
    name, _, envs = v.partition(":")
    pairs = [e.split("=", 1) for e in envs.split(",") if e]
    for k, val in pairs:
        os.environ[k] = val
    dd = os.path.join(d, "pp_gpu_" + name); os.makedirs(dd, exist_ok=True)
    got = util.run_with_dumps(lib, sub, dd)
    kern = lib.photon_b200_last_variant().decode()
    plain = sub.run(lib)
    for rk, ref in refs.items():
        m = util.compare(ref, got, 10.0)
        i0 = ref[0].astype(np.float64)
        m["image_rel_l2_nondump"] = float(np.linalg.norm(i0 - plain) / np.linalg.norm(i0))
        print("%-10s %-18s vs %-4s" % (name, kern, rk), json.dumps({k: (float("%.3g" % x) if isinstance(x, float) else x) for k, x in m.items()}), flush=True)
    for k, _ in pairs:
        os.environ.pop(k, None)
