"""Times the stages of start_ray_tracing on one C3 batch (PHOTON_TIMING=1 prints them on stderr)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from photon_b200 import scenes
from photon_b200.lib import load_library
os.environ["PHOTON_TIMING"] = "1"
lib = load_library(require_gpu=True)
d = "/tmp/photon_bench"; os.makedirs(d, exist_ok=True)
nrrd = os.path.join(d, "c3_blob_512.nrrd")
call = scenes.c3_bos(nrrd, n=512, dots_per_side=100, points_per_dot=100, rays_per_point=1000, nx=1024, ny=1024,
                     write_field=not os.path.exists(nrrd), source_point_number=10000)
sub = call.subset(slice(0, 10000))
img = np.zeros(1024 * 1024, np.float32)
for i in range(4):
    t = time.perf_counter(); sub.run(lib, img); print("call", i, "%.1f ms" % ((time.perf_counter() - t) * 1e3), file=sys.stderr)
