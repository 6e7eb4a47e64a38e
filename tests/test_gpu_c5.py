"""BASELINE C5 ingredients at test size: three sequential optical elements, a density volume handed over in
memory ("mem:<name>"), and the integrator step scale of the step-size sweep."""
import os

import numpy as np
import pytest

from oracle.loader import load_oracle
from photon_b200 import scenes
from photon_b200.nrrd import read_nrrd, register_field
from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c5(tmp_path_factory, cuda_lib):
    d = tmp_path_factory.mktemp("c5")
    path = str(d / "blob.nrrd")
    call = scenes.c5_stress(path, n=48, nz=40, dots_per_side=5, points_per_dot=6, rays_per_point=64, nx=512, ny=512)
    call.source_point_number = call.num_particles
    return call, path, cuda_lib, d


def test_three_elements_match_oracle(c5):
    call, path, lib, d = c5
    os.makedirs(d / "o", exist_ok=True); os.makedirs(d / "g", exist_ok=True)
    ref = util.run_with_dumps(load_oracle(), call, str(d / "o"))
    got = util.run_with_dumps(lib, call, str(d / "g"))
    m = util.compare(ref, got, 10.0)
    assert m["hit_fraction"] > 0.1 and m["nan_mismatch"] <= 2, m
    # two thick lenses: the float32 ray-sphere conditioning of tests/util.py applies (twice)
    assert m["dpos_px_p999"] <= util.TOL_POS_PX and m["dpos_px"] <= 2 * util.THICK_LENS_MAX_PX, m
    assert m["ddir_rad"] <= util.TOL_DIR_RAD and m["image_rel_l2"] <= util.TOL_IMG_L2, m


def test_field_from_memory_equals_field_from_file(c5, monkeypatch):
    call, path, lib, d = c5
    monkeypatch.setenv("PHOTON_DETERMINISTIC", "1")
    from_file = call.run(lib)
    rho, spacings, origin = read_nrrd(path)
    keep = register_field(lib, "c5test", rho, spacings, origin)
    call.density_grad_filename = "mem:c5test"
    try:
        from_mem = call.run(lib)
    finally:
        call.density_grad_filename = path
        lib.photon_b200_unregister_field(b"c5test")
    assert from_file.sum() > 0 and keep is not None
    np.testing.assert_array_equal(from_file, from_mem)
    call.density_grad_filename = "mem:never-registered"
    try:
        untouched = call.run(lib, image=np.full(call.image_shape, 2.0, np.float32))
    finally:
        call.density_grad_filename = path
    assert (untouched == 2.0).all()                      # failure leaves the caller's image alone (SURVEY B-12)


def test_step_scale_halves_the_step(c5, monkeypatch):
    from photon_b200.scene import ResidentScene
    call, path, lib, d = c5
    steps = {}
    for k in ("1", "0.5"):
        monkeypatch.setenv("PHOTON_STEP_SCALE", k)
        s = ResidentScene(call)
        s.trace()
        steps[k] = s.ray_steps / s.rays
        img = s.download_image()
        assert np.isfinite(img).all() and img.sum() > 0
        s.close()
    assert 1.9 < steps["0.5"] / steps["1"] < 2.1, steps


@pytest.mark.gpu
@pytest.mark.parametrize("alg", [2, 1])
def test_wide_index_hot_loop_equals_the_32_bit_one(alg, cuda_lib, tmp_path, monkeypatch):
    """Fields of more than 2^32 voxels (BASELINE C5 at its full 2048^3 with PHOTON_FULL_DEPTH: 128 GiB of texels) run
    the 64-bit-voxel-index instantiation of the coherent hot loop.  An oracle run at that size is not affordable, so the
    instantiation is forced (PHOTON_INDEX64=1) on a field that fits both: same texels, same arithmetic, only the width
    of the index -- ray dumps and image must be bit-identical."""
    from photon_b200 import scenes
    call = scenes.c3_bos(str(tmp_path / "blob.nrrd"), n=64, dots_per_side=8, points_per_dot=8, rays_per_point=40, algorithm=alg)
    monkeypatch.setenv("PHOTON_CELLS", "0")
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    narrow = util.run_with_dumps(cuda_lib, call, str(tmp_path / "a"))
    v0 = cuda_lib.photon_b200_last_variant().decode()
    monkeypatch.setenv("PHOTON_INDEX64", "1")
    wide = util.run_with_dumps(cuda_lib, call, str(tmp_path / "b"))
    v1 = cuda_lib.photon_b200_last_variant().decode()
    assert v0.endswith("/coherent/dump") and v1.endswith("/wide/dump"), (v0, v1)
    assert np.array_equal(narrow[1], wide[1], equal_nan=True) and np.array_equal(narrow[2], wide[2], equal_nan=True)
    img0 = call.run(cuda_lib)
    assert cuda_lib.photon_b200_last_variant().decode().endswith("/wide/image")
    monkeypatch.delenv("PHOTON_INDEX64")
    img1 = call.run(cuda_lib)
    assert float(np.linalg.norm(img0.astype(np.float64) - img1) / np.linalg.norm(img1.astype(np.float64))) <= 1e-6
