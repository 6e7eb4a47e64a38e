"""GPU: size-independent properties of the CUDA path at sizes the CPU oracle would not finish
(256^3 field, millions of rays): shard additivity, determinism, linearity in radiance, analytic
deflection through a constant density gradient, no deflection through a uniform field, step counts."""
import os

import numpy as np
import pytest

from photon_b200 import scenes
from photon_b200.nrrd import write_nrrd

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def blob_call(tmp_path_factory, cuda_lib):
    d = tmp_path_factory.mktemp("prop")
    return scenes.c3_bos(str(d / "blob256.nrrd"), n=256, dots_per_side=40, points_per_dot=20, rays_per_point=200)


def _scene(call, monkeypatch=None, shard=(0, 1), **env):
    from photon_b200.scene import ResidentScene
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        return ResidentScene(call, shard=shard)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def test_ray_steps_and_hits_at_scale(blob_call):
    s = _scene(blob_call)
    s.trace()
    rays = blob_call.num_particles * blob_call.lightray_number_per_particle
    assert s.rays == rays and rays > 5e6
    # 256 planes -> 255 committed RK4 steps per ray; a ray whose last sub-position rounds onto the far
    # face discards that step (App. A-7), so a few rays in 10^4 commit 254
    assert rays * 254.99 <= s.ray_steps <= rays * 255
    assert s.sensor_hits == rays                         # f/11 cone of 1e-4 pitch: nothing is vignetted
    img = s.download_image()
    assert np.isfinite(img).all() and img.sum() > 0
    s.close()


def test_deterministic_mode_is_bit_exact_and_shard_additive(blob_call):
    full = _scene(blob_call, PHOTON_DETERMINISTIC=1)
    assert full.deterministic
    full.trace()
    q_full, scale = full.download_image_fixed()
    img_a = full.download_image()
    full.zero_image(); full.trace()
    img_b = full.download_image()
    np.testing.assert_array_equal(img_a, img_b)          # run to run
    full.close()
    parts = []
    for r in range(3):
        sh = _scene(blob_call, shard=(r, 3), PHOTON_DETERMINISTIC=1)
        sh.trace()
        q, sc = sh.download_image_fixed()
        assert sc == scale                               # every shard uses the scale of the whole call
        parts.append(q); sh.close()
    np.testing.assert_array_equal(parts[0] + parts[1] + parts[2], q_full)   # any sharding, exactly
    # and it agrees with the float-atomic image
    fl = _scene(blob_call); fl.trace(); img_f = fl.download_image(); fl.close()
    assert np.linalg.norm(img_a.astype(np.float64) - img_f) <= 1e-5 * np.linalg.norm(img_f)


def test_float_mode_shards_sum_to_the_full_image(blob_call):
    fl = _scene(blob_call); fl.trace(); full = fl.download_image().astype(np.float64); fl.close()
    acc = np.zeros_like(full)
    for r in range(2):
        sh = _scene(blob_call, shard=(r, 2)); sh.trace(); acc += sh.download_image(); sh.close()
    assert np.linalg.norm(acc - full) <= 1e-5 * np.linalg.norm(full)


def test_image_is_linear_in_source_radiance(blob_call):
    import copy
    a = _scene(blob_call, PHOTON_DETERMINISTIC=1); a.trace(); img1 = a.download_image(); a.close()
    c2 = copy.copy(blob_call); c2.radiance = blob_call.radiance * 2.0
    b = _scene(c2, PHOTON_DETERMINISTIC=1); b.trace(); img2 = b.download_image(); b.close()
    np.testing.assert_array_equal(img2, 2.0 * img1)


def test_philox_rng_is_reproducible_and_statistically_equivalent(blob_call):
    a = _scene(blob_call, PHOTON_RNG="philox", PHOTON_SEED=7, PHOTON_DETERMINISTIC=1); a.trace(); i1 = a.download_image(); a.close()
    b = _scene(blob_call, PHOTON_RNG="philox", PHOTON_SEED=7, PHOTON_DETERMINISTIC=1); b.trace(); i2 = b.download_image(); b.close()
    c = _scene(blob_call, PHOTON_RNG="philox", PHOTON_SEED=8, PHOTON_DETERMINISTIC=1); c.trace(); i3 = c.download_image(); c.close()
    r = _scene(blob_call, PHOTON_DETERMINISTIC=1); r.trace(); i0 = r.download_image(); r.close()
    np.testing.assert_array_equal(i1, i2)
    assert not np.array_equal(i1, i3)
    assert abs(float(i1.sum()) - float(i0.sum())) <= 1e-3 * float(i0.sum())


def _dump_positions(lib, call, tmp):
    from tests import util
    return util.run_with_dumps(lib, call, str(tmp))


def test_uniform_field_does_not_deflect(tmp_path, cuda_lib):
    n, extent = 64, 511 * 176.0
    sp = extent / (n - 1)
    write_nrrd(str(tmp_path / "uniform.nrrd"), np.full((n, n, n), 1.225, np.float32), (sp,) * 3,
               (-extent / 2, -extent / 2, 750e3 - extent / 2))
    call = scenes.c3_bos(str(tmp_path / "uniform.nrrd"), n=n, dots_per_side=12, points_per_dot=6, rays_per_point=50,
                         write_field=False)
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    on = _dump_positions(cuda_lib, call, tmp_path / "a")
    call.simulate_density_gradients = False
    off = _dump_positions(cuda_lib, call, tmp_path / "b")
    d = np.abs(on[1] - off[1])[:, :2].max() / 10.0
    assert d <= 2e-3, d                                   # pixels; only float32 noise of 63 straight steps


def test_constant_gradient_matches_paraxial_theory(tmp_path, cuda_lib):
    """Deflection through a slab of constant density gradient: eps = K * drho/dx * L / n0
    (nrrd_functions.py:60-82), seen on the sensor as M * Z_D * eps / pixel_pitch pixels."""
    n, extent = 64, 511 * 176.0
    sp = extent / (n - 1)
    drho = 2e-5                                           # kg/m^3 per um
    rho = scenes.constant_gradient_slab(n, sp, drho)
    write_nrrd(str(tmp_path / "slab.nrrd"), rho, (sp,) * 3, (-extent / 2, -extent / 2, 750e3 - extent / 2))
    call = scenes.c3_bos(str(tmp_path / "slab.nrrd"), n=n, dots_per_side=8, points_per_dot=4, rays_per_point=40,
                         write_field=False)
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    on = _dump_positions(cuda_lib, call, tmp_path / "a")
    call.simulate_density_gradients = False
    off = _dump_positions(cuda_lib, call, tmp_path / "b")
    shift_px = np.nanmean(on[1][:, 0] - off[1][:, 0]) / 10.0
    shift_y = np.nanmean(on[1][:, 1] - off[1][:, 1]) / 10.0
    K, L = 0.225e-3, extent
    n0 = 1 + K * 1.225
    eps = K * drho * L / n0                               # rad, towards +x
    z_d = 250e3                                           # volume centre to dot plane (rays enter at +250 mm ... )
    M = call.image_distance / 1e6
    expected = M * z_d * eps / 10.0
    assert abs(shift_y) < 0.02 * abs(expected) + 1e-3
    assert abs(abs(shift_px) - expected) <= 0.05 * expected, (shift_px, expected)
