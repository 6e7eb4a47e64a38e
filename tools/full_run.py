"""The whole BASELINE C3 job through ONE start_ray_tracing call, host buffers in, image out -- what
perform_ray_tracing_03.py does per image.  Wall clock of the call: cold (NRRD read from disk, field build,
upload) and warm (field cached in HBM, as for the second image of a BOS pair).
usage: python tools/full_run.py [--gpus N]      (N > 1: PHOTON_GPUS, the in-call split over the GPUs of the box)"""
import argparse, json, os, sys, tempfile, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from photon_b200 import scenes
from photon_b200.lib import load_library

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--grid", type=int, default=512)
ap.add_argument("--rays-per-point", type=int, default=1000)
args = ap.parse_args()
if args.gpus > 1:
    os.environ["PHOTON_GPUS"] = str(args.gpus)
d = os.path.join(tempfile.gettempdir(), "photon_bench"); os.makedirs(d, exist_ok=True)
nrrd = os.path.join(d, f"c3_blob_{args.grid}.nrrd")
call = scenes.c3_bos(nrrd, n=args.grid, dots_per_side=100, points_per_dot=100, rays_per_point=args.rays_per_point,
                     write_field=not os.path.exists(nrrd))
lib = load_library(require_gpu=True)
rays = call.num_particles * call.lightray_number_per_particle
# the N the library will really use: PHOTON_GPUS as the library reads it (set above from --gpus, or by the caller), capped by the box
gpus_used = max(1, min(int(os.environ.get("PHOTON_GPUS", "1")), lib.photon_b200_device_count()))
out = {"workload": f"C3 {args.grid}^3, {call.num_particles} source points x {call.lightray_number_per_particle} rays = {rays:.3e} rays, one start_ray_tracing call",
       "gpus": gpus_used, "gpus_requested": args.gpus}
for tag in ("cold", "warm"):
    t0 = time.perf_counter()
    img = call.run(lib)
    dt = time.perf_counter() - t0
    out[tag + "_seconds"] = dt
    out[tag + "_rays_per_s"] = rays / dt
    out["image_sum"] = float(img.sum(dtype=np.float64))
print(json.dumps(out))
