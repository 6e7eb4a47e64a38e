"""BASELINE configs C1 / C2: the exact start_ray_tracing calls that the reference's UNCHANGED scripts
make for `sample_run_script.sh piv` and `sample_run_script.sh bos` (recorded by tools/stage_reference.py
+ tools/record_stub.c, stored by tests/golden/make_recorded.py), replayed at full size, against the
images the reference's own sources (host build, oracle/_ref) produced for those calls.

Tolerance: the north-star's image bound, <= 1e-4 relative L2 (both sides accumulate float32 in
arbitrary order).  PIV reference images are stored 4x4-binned (fixture size), compared binned.
"""
import os

import numpy as np
import pytest

from photon_b200 import recording
from tests import util

GOLDEN = util.GOLDEN_DIR


def binned(img, b=4):
    ny, nx = img.shape
    return img.astype(np.float64).reshape(ny // b, b, nx // b, b).sum(axis=(1, 3))


def rel_l2(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def bos_calls():
    c0, ex = recording.load_npz(os.path.join(GOLDEN, "recorded_bos.npz"))
    import copy
    c1 = copy.copy(c0)
    c1.simulate_density_gradients = bool(ex["call1_simulate_density_gradients"])
    c1.ray_tracing_algorithm = int(ex["call1_ray_tracing_algorithm"])
    assert str(ex["call1_density_grad_basename"]) == "sample-density.nrrd"
    c1.density_grad_filename = os.path.join(GOLDEN, "sample-density.nrrd")
    return (c0, c1), ex


def piv_calls():
    c0, ex = recording.load_npz(os.path.join(GOLDEN, "recorded_piv.npz"))
    import copy
    c1 = copy.copy(c0)
    for k in ("x", "y", "z", "radiance", "diameter_index"):
        setattr(c1, k, ex["call1_" + k])
    return (c0, c1), ex


def test_recorded_arguments_are_what_the_scripts_pass():
    """the values observed at the boundary."""
    (b0, b1), _ = bos_calls()
    assert (b0.num_particles, b0.lightray_number_per_particle, b0.source_point_number) == (120000, 500, 10000)
    assert not b0.simulate_density_gradients and b1.simulate_density_gradients and b1.ray_tracing_algorithm == 2
    assert b0.scattering_type == "diffuse" and b0.mie_irradiance.size == 0
    assert bool(b0.camera.implement_diffraction) and abs(b0.camera.diffraction_diameter - 3.0) < 1e-6
    assert abs(b0.ray_cone_pitch_ratio - 1e-4) < 1e-9 and abs(b0.lens_pitch - 13125.0) < 1e-3
    (p0, p1), _ = piv_calls()
    assert (p0.num_particles, p0.lightray_number_per_particle, p0.source_point_number) == (50000, 10000, 10000)
    assert p0.scattering_type == "mie" and p0.mie_irradiance.shape == (255, 27)
    assert not bool(p0.camera.implement_diffraction) and p0.ray_cone_pitch_ratio == 1.0
    assert not np.array_equal(p0.x, p1.x)                       # second frame: displaced particles
    e = p0.elements[0]
    assert e.element_type == b"l" and abs(e.element_properties.refractive_index - 1.476522) < 1e-5


def test_oracle_reproduces_reference_on_recorded_bos_call(oracle_lib):
    """CPU: the restatement on the full second BOS call (density field on, 6e7 rays)."""
    (_, c1), ex = bos_calls()
    img = c1.run(oracle_lib)
    assert rel_l2(img, ex["ref_image1"]) <= 5e-5        # float32 accumulation order of ~3e4 increments per pixel
    assert abs(float(img.astype(np.float64).sum()) / float(ex["ref_sum1"]) - 1) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("i", [0, 1])
def test_cuda_sample_bos_matches_reference(i, cuda_lib):
    calls, ex = bos_calls()
    img = calls[i].run(cuda_lib)
    m = rel_l2(img, ex["ref_image%d" % i])
    print("sample-data/bos call", i, "image rel L2 vs reference:", m)
    assert m <= util.TOL_IMG_L2
    assert abs(float(img.astype(np.float64).sum()) / float(ex["ref_sum%d" % i]) - 1) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("i", [0, 1])
def test_cuda_sample_piv_matches_reference(i, cuda_lib):
    calls, ex = piv_calls()
    img = calls[i].run(cuda_lib)
    m = rel_l2(binned(img), ex["ref_binned%d" % i])
    print("sample-data/piv frame", i, "binned image rel L2 vs reference:", m)
    assert m <= util.TOL_IMG_L2
    assert abs(float(img.astype(np.float64).sum()) / float(ex["ref_sum%d" % i]) - 1) < 1e-5


# ---- sample-data/bos/sample-parameters.mat: the shipped parameter file whose rays DO cross sample-density.nrrd -------
# Camera turned 150 degrees about y (camera <-> volume rotation of prt.cu:2061-2076 / 2100-2111 with a non-identity
# matrix), Euler (algorithm 1), thin lens, 512 x 512, 7225 regular dots x 100 rays.  The reference's unchanged scripts
# wrote per-ray dumps for it (save_lightrays = 1 in the file); every 180th ray of the second call is stored.
def rotated_calls():
    c0, ex = recording.load_npz(os.path.join(GOLDEN, "recorded_bos_rotated.npz"))
    import copy
    c1 = copy.copy(c0)
    c1.simulate_density_gradients = bool(ex["call1_simulate_density_gradients"])
    c1.ray_tracing_algorithm = int(ex["call1_ray_tracing_algorithm"])
    assert str(ex["call1_density_grad_basename"]) == "sample-density.nrrd"
    c1.density_grad_filename = os.path.join(GOLDEN, "sample-density.nrrd")
    return (c0, c1), ex


def test_rotated_recording_is_what_the_parameter_file_says():
    (c0, c1), ex = rotated_calls()
    assert (c0.num_particles, c0.lightray_number_per_particle) == (7225, 100)
    assert c1.ray_tracing_algorithm == 1 and c0.elements[0].element_type == b"t"
    assert c0.image_shape == (512, 512)
    R = np.array(list(c0.camera.rotation_matrix)).reshape(3, 3)
    assert abs(R[0, 0] - np.cos(np.deg2rad(150.0))) < 1e-6 and abs(abs(R[0, 2]) - 0.5) < 1e-6     # not the identity
    Ri = np.array(list(c0.camera.inverse_rotation_matrix)).reshape(3, 3)
    assert np.allclose(R @ Ri, np.eye(3), atol=1e-6)
    # the field moves the rays: the two reference images differ
    assert rel_l2(ex["ref_image1"], ex["ref_image0"]) > 1e-2


def _rotated_dumps(lib, call, tmp_path):
    call.source_point_number = call.num_particles
    img, pos, dr = util.run_with_dumps(lib, call, str(tmp_path))
    return img, pos, dr


@pytest.mark.parametrize("which", ["oracle", "hostcheck"])
def test_cpu_rotated_bos_matches_reference(which, oracle_lib, hostcheck_lib, tmp_path):
    """CPU: the restatement and the product's stage code compiled for the host, on the rotated-camera call through
    sample-density.nrrd, bit for bit on the rays the reference dumped."""
    (_, c1), ex = rotated_calls()
    img, pos, dr = _rotated_dumps(oracle_lib if which == "oracle" else hostcheck_lib, c1, tmp_path)
    idx = ex["ray_index"]
    assert len(pos) == 722500 and int(np.isfinite(pos[:, 0]).sum()) == int(ex["ref_hits"])
    assert np.array_equal(pos[idx], ex["ref_pos"], equal_nan=True)
    assert np.array_equal(dr[idx], ex["ref_dir"], equal_nan=True)
    assert rel_l2(img, ex["ref_image1"]) <= 5e-6


@pytest.mark.gpu
def test_cuda_rotated_bos_matches_reference(cuda_lib, tmp_path):
    """GPU: both calls of the rotated sample; the second marches through sample-density.nrrd with a non-identity
    camera <-> volume rotation.  North-star tolerances on the reference's own ray dumps and images."""
    (c0, c1), ex = rotated_calls()
    img0 = c0.run(cuda_lib)
    assert rel_l2(img0, ex["ref_image0"]) <= util.TOL_IMG_L2
    plain = c1.run(cuda_lib)                                            # the non-dump kernel
    assert rel_l2(plain, ex["ref_image1"]) <= util.TOL_IMG_L2
    img, pos, dr = _rotated_dumps(cuda_lib, c1, tmp_path)
    idx = ex["ray_index"]
    m = util.compare((ex["ref_image1"], ex["ref_pos"], ex["ref_dir"]), (img, pos[idx], dr[idx]), 10.0)
    print("rotated sample BOS vs reference:", m)
    util.assert_parity("bos_rotated_euler", m)
    assert abs(int(np.isfinite(pos[:, 0]).sum()) - int(ex["ref_hits"])) <= 2
