"""Texture-unit characterisation probe: tex3D<float4> on a random-noise field at random
coordinates, saved with the texels for offline fitting of the unit's weight arithmetic.
usage: python tools/texprobe.py gpurun_out/texprobe.npz"""
import os, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from photon_b200.lib import load_library
from photon_b200.nrrd import write_nrrd
lib = load_library(require_gpu=True)
n = 24
rng = np.random.default_rng(5)
rho = (1.2 + 0.3 * rng.random((n, n, n))).astype(np.float32)
d = tempfile.mkdtemp()
path = os.path.join(d, "noise.nrrd")
write_nrrd(path, rho, (100.0, 100.0, 100.0), (0.0, 0.0, 750e3))
meta = np.zeros(16, np.float32)
texels = np.zeros((n, n, n, 4), np.float32)
assert lib.photon_b200_load_field(path.encode(), meta.ctypes.data, texels.ctypes.data, n ** 3) == 0
pts = (1.0 + rng.random((20000, 3)) * (n - 3)).astype(np.float32)
pts[:4000, 1:] = np.floor(pts[:4000, 1:]) + 0.5          # fraction 0 in y and z: a 1-D probe of the x weight
pts[:2000, 0] = np.floor(pts[:2000, 0]) + 0.5 + np.arange(2000, dtype=np.float32) / 2000.0   # x fraction sweeps 0..1
out = {}
for mode, name in ((0, "exact"), (2, "hw")):
    o = np.zeros((len(pts), 4), np.float32)
    assert lib.photon_b200_sample_field(path.encode(), mode, len(pts), pts.ctypes.data, o.ctypes.data) == 0
    out[name] = o
np.savez_compressed(sys.argv[1], texels=texels, pts=pts, **out)
print("saved", sys.argv[1], "max |hw - exact| / scale:", float((np.abs(out["hw"] - out["exact"]) / np.abs(texels).reshape(-1, 4).max(axis=0)).max()))
