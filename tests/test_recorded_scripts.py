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
