#!/usr/bin/env python
"""Generates tests/golden/*.npz from the REFERENCE'S OWN SOURCES compiled for the host
(oracle/build_ref.sh -> oracle/_ref/libphoton_ref_cpu.so; needs /root/reference, so it runs
in the build container only).  The fixtures travel to the GPU box, the reference does not.

Each fixture holds, for one scene of tests/util.parity_cases (inputs are rebuilt
deterministically from photon_b200.scenes, nothing random):
  pos, dir   : float32 [n,3] per-ray sensor hit / post-volume direction dumps of the first
               GOLDEN_RAYS rays of every batch file (NaN = culled), as the reference wrote them
  tiles      : float64 [ny/16, nx/16] sums of the reference image over 16x16-pixel tiles
  image_sum  : float64 total radiance on the sensor
  nonzero    : number of non-zero pixels
usage:  python tests/golden/make_golden.py
"""
import ctypes
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import util  # noqa: E402

GOLDEN_RAYS = 4000


def summarise(image, pos, dr):
    ny, nx = image.shape
    tiles = image.astype(np.float64).reshape(ny // 16, 16, nx // 16, 16).sum(axis=(1, 3))
    return dict(pos=pos[:GOLDEN_RAYS].copy(), dir=dr[:GOLDEN_RAYS].copy(), tiles=tiles,
                image_sum=np.float64(image.astype(np.float64).sum()), nonzero=np.int64((image != 0).sum()),
                rays=np.int64(len(pos)), hits=np.int64(np.isfinite(pos[:, 0]).sum()))


def main():
    subprocess.run(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], check=True)
    ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libphoton_ref_cpu.so"))
    out = os.path.dirname(os.path.abspath(__file__))
    work = tempfile.mkdtemp(prefix="photon_golden_")
    for name, (call, _pitch) in util.parity_cases(work).items():
        d = os.path.join(work, name)
        os.makedirs(d)
        img, pos, dr = util.run_with_dumps(ref, call, d)
        np.savez_compressed(os.path.join(out, name + ".npz"), **summarise(img, pos, dr))
        print(name, "rays", len(pos), "hits", int(np.isfinite(pos[:, 0]).sum()), "sum", float(img.sum()))


if __name__ == "__main__":
    main()
