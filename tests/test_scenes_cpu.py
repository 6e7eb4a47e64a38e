"""CPU: the scene builders of the larger BASELINE configurations at test size -- C4 (turbulent field, Mie table
from bhmie, thick lens, bilinear splat) and C5 (three sequential optical elements) -- through the product's stage
code compiled for the host (tests/hostcheck) against the oracle, bit for bit on the ray dumps."""
import numpy as np
import pytest

from photon_b200 import scenes
from tests import util


def _check(call, pitch, hostcheck_lib, oracle_lib, tmp_path):
    (tmp_path / "o").mkdir(); (tmp_path / "h").mkdir()
    call.source_point_number = call.num_particles
    ref = util.run_with_dumps(oracle_lib, call, str(tmp_path / "o"))
    got = util.run_with_dumps(hostcheck_lib, call, str(tmp_path / "h"))
    assert np.isfinite(ref[1][:, 0]).mean() > 0.05
    assert np.array_equal(ref[1], got[1], equal_nan=True)
    assert np.array_equal(ref[2], got[2], equal_nan=True)
    assert util.compare(ref, got, pitch)["image_rel_l2"] <= 1e-7
    return ref


def test_c4_scene(hostcheck_lib, oracle_lib, tmp_path):
    call = scenes.c4_piv(str(tmp_path / "turb.nrrd"), n=48, particles=300, rays_per_particle=40, nx=1024, ny=1024)
    assert call.mie_irradiance.shape == (255, 27) and call.simulate_density_gradients
    with_field = _check(call, 17.0, hostcheck_lib, oracle_lib, tmp_path)
    call.simulate_density_gradients = False
    (tmp_path / "n").mkdir()
    without = util.run_with_dumps(oracle_lib, call, str(tmp_path / "n"))
    both = np.isfinite(with_field[1][:, 0]) & np.isfinite(without[1][:, 0])
    shift = np.abs(with_field[1][both] - without[1][both])[:, :2].max()
    assert 1e-3 < shift < 100.0                      # the turbulent field moves the sensor hits by a fraction of a pixel


def test_c5_scene_three_elements(hostcheck_lib, oracle_lib, tmp_path):
    call = scenes.c5_stress(str(tmp_path / "blob.nrrd"), n=32, nz=24, dots_per_side=5, points_per_dot=4, rays_per_point=32,
                            nx=512, ny=512)
    assert len(call.elements) == 3 and list(call.element_system_index) == [3, 2, 1]
    _check(call, 10.0, hostcheck_lib, oracle_lib, tmp_path)


def test_turbulent_field_is_band_limited():
    f = scenes.turbulent_field(32, k_lo=2.0, k_hi=6.0, rms=0.05)
    assert f.shape == (32, 32, 32) and abs(f.std() - 0.05) < 1e-4 and abs(f.mean() - 1.225) < 1e-4
    F = np.abs(np.fft.fftn(f - f.mean()))
    k = np.sqrt(sum(np.meshgrid(*[np.fft.fftfreq(32, 1 / 32.0) ** 2] * 3, indexing="ij")))
    assert F[(k > 6.5)].max() < 1e-3 * F.max() and F[(k < 1.5)].max() < 1e-3 * F.max()
