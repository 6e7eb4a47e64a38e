"""CPU: the oracle (oracle/photon_oracle.cpp) pinned against the reference.

1. tests/golden/*.npz are outputs of the REFERENCE'S OWN SOURCES compiled for the host
   (tests/golden/make_golden.py, oracle/build_ref.sh); the oracle must reproduce the per-ray
   dumps bit for bit and the image to summation-order accuracy.
2. Where /root/reference is present (the build container) the reference is run live on all rays
   of every scene and compared with the oracle the same way.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASE_NAMES = ["bos_rk4_thin_diffraction", "bos_euler_thin_diffraction", "bos_nogradient",
              "bos_rk4_multibatch_ragged", "piv_mie_thick_bilinear", "piv_diffuse_thin_bilinear",
              "piv_apparent", "piv_mie_thick_diffraction", "piv_chief_ray_only", "bos_rk45_thin_diffraction", "bos_ab4_thin_diffraction"]


@pytest.fixture(scope="module")
def cases(tmp_path_factory):
    return util.parity_cases(tmp_path_factory.mktemp("golden"))


@pytest.fixture(scope="module")
def reference_lib():
    so = os.path.join(ROOT, "oracle", "_ref", "libphoton_ref_cpu.so")
    if os.path.isdir("/root/reference/cuda_codes"):
        subprocess.run(["bash", os.path.join(ROOT, "oracle", "build_ref.sh")], check=True, capture_output=True)
    if not os.path.exists(so):
        pytest.skip("oracle/_ref/libphoton_ref_cpu.so not built (needs /root/reference)")
    return ctypes.CDLL(so)


@pytest.mark.parametrize("name", CASE_NAMES)
def test_oracle_reproduces_reference_golden(name, cases, oracle_lib, tmp_path):
    call, _ = cases[name]
    g = util.load_golden(name)
    img, pos, dr = util.run_with_dumps(oracle_lib, call, str(tmp_path))
    n = util.GOLDEN_RAYS
    assert len(pos) == int(g["rays"]) and int(np.isfinite(pos[:, 0]).sum()) == int(g["hits"])
    assert np.array_equal(pos[:n], g["pos"], equal_nan=True)          # bit-exact sensor hits
    assert np.array_equal(dr[:n], g["dir"], equal_nan=True)           # bit-exact post-volume directions
    tiles = util.tile_sums(img)
    assert np.linalg.norm(tiles - g["tiles"]) <= 2e-6 * np.linalg.norm(g["tiles"])   # float atomics order
    assert int((img != 0).sum()) == int(g["nonzero"])


@pytest.mark.parametrize("name", CASE_NAMES)
def test_oracle_matches_reference_sources_live(name, cases, oracle_lib, reference_lib, tmp_path, capfd):
    call, pitch = cases[name]
    (tmp_path / "o").mkdir()
    (tmp_path / "r").mkdir()
    a = util.run_with_dumps(oracle_lib, call, str(tmp_path / "o"))
    b = util.run_with_dumps(reference_lib, call, str(tmp_path / "r"))
    capfd.readouterr()                                                # the reference prints progress
    assert np.array_equal(a[1], b[1], equal_nan=True)
    assert np.array_equal(a[2], b[2], equal_nan=True)
    m = util.compare(a, b, pitch)
    assert m["image_rel_l2"] <= 5e-6, m
