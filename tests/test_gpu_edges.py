"""Edge cases at the start_ray_tracing boundary: empty input, rays that never meet the field, an
unreadable field file, vignetted rays, and the two noise modes.

Reference behaviour each one follows:
  * empty / ragged source arrays: the batch loop of prt.cu:3447-3560 runs ceil(n / source_point_number)
    times, zero times for n = 0, and the image comes back as it went in (prt.cu:3309);
  * a ray that misses the volume is returned unchanged (trace.h:1485-1496), so the result is the one of a
    call without density gradients;
  * NRRD failure: the reference exit()s (trace.h:1683); SURVEY App. B-12 -> message, image untouched;
  * vignetted rays are all-NaN (prt.cu:444-496, 575) and stay NaN in the dumps;
  * noise (prt.cu:1424-1434, trace.h:852-863): the reference seeds cuRAND from the clock, so only the
    statistics are pinned there; product and oracle share one Philox stream per ray, so here the
    comparison is per ray at the usual tolerances, plus the statistics.
"""
import copy
import os

import numpy as np
import pytest

from photon_b200 import scenes
from photon_b200.nrrd import write_nrrd
from tests import util

pytestmark = pytest.mark.gpu


def dump_dir(tmp_path, name):
    d = os.path.join(str(tmp_path), name)
    os.makedirs(d, exist_ok=True)
    return d


def small_bos(tmp_path, **kw):
    nrrd = os.path.join(str(tmp_path), "blob64.nrrd")
    args = dict(n=64, dots_per_side=8, points_per_dot=6, rays_per_point=40)
    args.update(kw)
    return scenes.c3_bos(nrrd, **args)


def test_empty_call_leaves_image_untouched(cuda_lib, tmp_path):
    call = small_bos(tmp_path)
    empty = call.subset(slice(0, 0))
    assert empty.num_particles == 0
    ny, nx = call.image_shape
    before = np.random.default_rng(3).random((ny, nx), dtype=np.float32)
    after = empty.run(cuda_lib, image=before)
    assert np.array_equal(after, before)
    nograd = copy.copy(empty)
    nograd.simulate_density_gradients = False
    assert np.array_equal(nograd.run(cuda_lib, image=before), before)


def test_single_point_single_ray(cuda_lib, oracle_lib, tmp_path):
    """Smallest non-empty input: one source point, one (chief) ray."""
    call = small_bos(tmp_path, rays_per_point=1).subset(slice(5, 6))
    ref = util.run_with_dumps(oracle_lib, call, dump_dir(tmp_path, "o1"))
    got = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "g1"))
    assert len(got[1]) == 1 and np.isfinite(got[1]).all()
    m = util.compare(ref, got, 10.0)
    assert m["dpos_px"] <= util.TOL_POS_PX and m["ddir_rad"] <= util.TOL_DIR_RAD and m["image_rel_l2"] <= util.TOL_IMG_L2, m


def test_rays_that_miss_the_field_are_unchanged(cuda_lib, oracle_lib, tmp_path):
    """A field box off to the side (x in [+200, +290] mm) is met by no ray: positions, directions and the
    image are those of the same call without density gradients, up to the float32 round trip of the
    volume-frame shift (z - (z_offset + 750e3) and back, prt.cu:2045,2119: half an ulp at 1e6 um)."""
    call = small_bos(tmp_path, write_field=False)
    n, extent = 32, 511 * 176.0
    sp = extent / (n - 1)
    path = os.path.join(str(tmp_path), "aside.nrrd")
    write_nrrd(path, scenes.gaussian_blob(n, sp, 15e3), (sp,) * 3, (200e3, -extent / 2, 750e3 - extent / 2))
    call.density_grad_filename = path
    plain = copy.copy(call)
    plain.simulate_density_gradients = False
    hit = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "a"))
    ref = util.run_with_dumps(cuda_lib, plain, dump_dir(tmp_path, "b"))
    orc = util.run_with_dumps(oracle_lib, call, dump_dir(tmp_path, "c"))
    assert np.isfinite(hit[1]).all()
    m = util.compare(ref, hit, 10.0)
    assert m["nan_mismatch"] == 0 and m["dpos_px"] <= 1e-3 and m["ddir_rad"] <= 1e-6, m
    m = util.compare(orc, hit, 10.0)
    assert m["nan_mismatch"] == 0 and m["dpos_px"] <= util.TOL_POS_PX and m["ddir_rad"] <= util.TOL_DIR_RAD, m
    assert m["image_rel_l2"] <= util.TOL_IMG_L2, m


def test_unreadable_field_reports_and_leaves_image(cuda_lib, tmp_path, capfd):
    import ctypes
    call = small_bos(tmp_path, write_field=False)
    call.density_grad_filename = os.path.join(str(tmp_path), "no_such_file.nrrd")
    ny, nx = call.image_shape
    before = np.full((ny, nx), 2.5, np.float32)
    after = call.run(cuda_lib, image=before)
    assert np.array_equal(after, before)
    cuda_lib.photon_b200_last_error.restype = ctypes.c_char_p
    msg = cuda_lib.photon_b200_last_error().decode()
    assert "no_such_file.nrrd" in msg or "nrrd" in msg.lower(), msg
    # a truncated file (header promises 64^3 floats, data is short) is refused the same way
    bad = os.path.join(str(tmp_path), "short.nrrd")
    with open(bad, "wb") as f:
        f.write(b"NRRD0005\ntype: float\ndimension: 3\nsizes: 64 64 64\nendian: little\nencoding: raw\n"
                b"spacings: 100 100 100\nspace origin: (0,0,0)\n\n" + b"\0" * 1000)
    call.density_grad_filename = bad
    assert np.array_equal(call.run(cuda_lib, image=before), before)
    # and the library is usable afterwards
    good = small_bos(tmp_path)
    assert good.run(cuda_lib).sum() > 0


def test_vignetted_rays_are_nan_in_both(cuda_lib, oracle_lib, tmp_path):
    """PIV aim radius up to the full pitch (ray_cone_pitch_ratio = 1): rays with r1 > 0.5
    miss the lens aperture and are culled -- the same rays on both sides."""
    call = scenes.c1_piv(particles=120, rays_per_particle=250, mie=False, lens_model="thin-lens")
    ref = util.run_with_dumps(oracle_lib, call, dump_dir(tmp_path, "o"))
    got = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "g"))
    nan_ref, nan_got = np.isnan(ref[1][:, 0]), np.isnan(got[1][:, 0])
    assert 0.2 < nan_got.mean() < 0.8, nan_got.mean()
    assert (nan_ref != nan_got).sum() <= 2
    assert np.isnan(got[1][nan_got]).all()                   # every component NaN, not just x


@pytest.mark.parametrize("scene", ["bos_diffraction", "piv_bilinear"])
def test_position_noise_matches_oracle_and_has_the_requested_spread(scene, cuda_lib, oracle_lib, tmp_path):
    """add_pos_noise: hit += sigma * pitch * N(0,1) per axis (prt.cu:1424-1434, 1777-1790)."""
    if scene == "bos_diffraction":
        call, pitch = small_bos(tmp_path), 10.0
    else:
        call, pitch = scenes.c1_piv(particles=150, rays_per_particle=200, mie=False, lens_model="thin-lens"), 17.0
    clean = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "clean"))
    sigma = 0.7
    call.add_pos_noise, call.pos_noise_std = True, sigma
    ref = util.run_with_dumps(oracle_lib, call, dump_dir(tmp_path, "o"))
    got = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "g"))
    m = util.compare(ref, got, pitch)
    assert m["nan_mismatch"] <= 2 and m["dpos_px"] <= util.TOL_POS_PX and m["image_rel_l2"] <= util.TOL_IMG_L2, m
    both = np.isfinite(clean[1][:, 0]) & np.isfinite(got[1][:, 0])
    d = (got[1][both] - clean[1][both])[:, :2] / pitch
    assert both.sum() > 5000
    se = sigma / np.sqrt(both.sum())
    assert np.abs(d.mean(axis=0)).max() < 5 * se, d.mean(axis=0)
    assert np.abs(d.std(axis=0) - sigma).max() < 0.03 * sigma, d.std(axis=0)
    assert abs(np.corrcoef(d[:, 0], d[:, 1])[0, 1]) < 0.05
    # a second run draws the same numbers (counter-based stream keyed by the ray index)
    again = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "g2"))
    assert np.array_equal(again[1], got[1], equal_nan=True)


def test_gradient_noise_matches_oracle(cuda_lib, oracle_lib, tmp_path):
    """add_ngrad_noise (Euler only, trace.h:852-863): g.xy += sigma * N(0,1) at every step."""
    call = small_bos(tmp_path, algorithm=1)
    clean = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "clean"))
    call.add_ngrad_noise, call.ngrad_noise_std = True, 2e-9         # 1/um; the blob's own gradient is ~1e-9
    ref = util.run_with_dumps(oracle_lib, call, dump_dir(tmp_path, "o"))
    got = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "g"))
    m = util.compare(ref, got, 10.0)
    assert m["nan_mismatch"] <= 2 and m["dpos_px"] <= util.TOL_POS_PX and m["ddir_rad"] <= util.TOL_DIR_RAD, m
    assert m["image_rel_l2"] <= util.TOL_IMG_L2, m
    both = np.isfinite(clean[2][:, 0]) & np.isfinite(got[2][:, 0])
    dd = (got[2][both] - clean[2][both])[:, :2]
    # a random walk of ~63 steps of 1427 um: std of the direction = sigma * step * sqrt(steps)
    expect = 2e-9 * (511 * 176.0 / 63) * np.sqrt(63)
    assert 0.7 * expect < dd.std() < 1.3 * expect, (dd.std(), expect)
    assert np.abs(dd[:, :2].mean(axis=0)).max() < 5 * expect / np.sqrt(both.sum())


def test_cell_table_agrees_with_corner_texels(cuda_lib, oracle_lib, tmp_path, monkeypatch):
    """The coherent RK4 loop reads each cell from the per-cell polynomial table; PHOTON_CELLS=0 and a
    table that does not fit (PHOTON_CELLS_MAX_GIB=0) read the corner texels.  All three against the oracle at the
    usual bounds; the two corner-texel routes are the same kernel and agree bit for bit."""
    call = small_bos(tmp_path, rays_per_point=60)
    ref = util.run_with_dumps(oracle_lib, call, dump_dir(tmp_path, "o"))
    table = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "t"))
    monkeypatch.setenv("PHOTON_CELLS", "0")
    corners = util.run_with_dumps(cuda_lib, call, dump_dir(tmp_path, "c"))
    monkeypatch.delenv("PHOTON_CELLS")
    second = tmp_path / "second"                              # its own file: the refusal is remembered with the cached field
    second.mkdir()
    other = small_bos(second, rays_per_point=60)
    monkeypatch.setenv("PHOTON_CELLS_MAX_GIB", "0")
    refused = util.run_with_dumps(cuda_lib, other, dump_dir(tmp_path, "r"))
    monkeypatch.delenv("PHOTON_CELLS_MAX_GIB")
    for name, got in (("table", table), ("corners", corners), ("refused", refused)):
        m = util.compare(ref, got, 10.0)
        assert m["nan_mismatch"] == 0 and m["dpos_px"] <= util.TOL_POS_PX and m["ddir_rad"] <= util.TOL_DIR_RAD, (name, m)
        assert m["image_rel_l2"] <= util.TOL_IMG_L2, (name, m)
    assert np.array_equal(corners[1], refused[1]) and np.array_equal(corners[2], refused[2])
    m = util.compare(corners, table, 10.0)
    assert m["dpos_px"] <= util.TOL_POS_PX and m["image_rel_l2"] <= util.TOL_IMG_L2, m
    assert not np.array_equal(corners[1], table[1])          # it really is a different evaluation order


# ---- what the library keeps between calls (field cache) and large splat footprints ------------------------------------
def test_field_of_a_live_scene_survives_other_loads_and_release(tmp_path, cuda_lib):
    """Cache entries are reference counted: loading another field on the same device, or asking the library to drop
    its cache, never frees the buffers a live scene reads.  Once the scene is gone the release returns the memory."""
    from photon_b200.scene import ResidentScene
    call = small_bos(tmp_path)
    cuda_lib.photon_b200_release_cache()
    assert cuda_lib.photon_b200_cache_bytes() == 0
    s = ResidentScene(call)
    s.trace(); first = s.download_image()
    held = cuda_lib.photon_b200_cache_bytes()
    assert held >= 16 * 64 ** 3
    # another file on the same device while the scene is alive
    other = os.path.join(str(tmp_path), "other.nrrd")
    write_nrrd(other, scenes.gaussian_blob(32, 511 * 176.0 / 31, 15e3), (511 * 176.0 / 31,) * 3, (-44968.0, -44968.0, 705032.0))
    meta = np.zeros(16, np.float32)
    assert cuda_lib.photon_b200_load_field(other.encode(), meta.ctypes.data, None, 0) == 0
    assert cuda_lib.photon_b200_release_cache() >= 16 * 32 ** 3          # the idle entry goes, the held one stays
    assert cuda_lib.photon_b200_cache_bytes() == held
    s.zero_image(); s.trace(); again = s.download_image()
    np.testing.assert_allclose(again, first, rtol=2e-5, atol=1e-6 * float(first.max()))
    s.close()
    assert cuda_lib.photon_b200_release_cache() == held
    assert cuda_lib.photon_b200_cache_bytes() == 0
    # and the next call simply loads it again
    assert call.run(cuda_lib).sum() > 0


@pytest.mark.parametrize("model,diam", [("apparent", 8.0), ("general", 9.0), ("apparent", 3.0)])
def test_large_diffraction_footprints_render(model, diam, cuda_lib, oracle_lib):
    """Footprints wider than the register path (more than 8 rows: D > ~3.5 px, or the apparent-image model whose
    radius is a full diameter) take the per-pixel path: nothing is dropped, the image still matches the oracle's
    (the reference has no bound on the diameter, prt.cu:1466-1540)."""
    c = scenes.c1_piv(particles=60, rays_per_particle=64, lens_model=model, nx=256, ny=256)
    c.camera.diffraction_diameter = diam
    c.camera.implement_diffraction = True
    ref = c.run(oracle_lib)
    got = c.run(cuda_lib)
    assert ref.sum() > 0
    rel = float(np.linalg.norm(got.astype(np.float64) - ref) / np.linalg.norm(ref.astype(np.float64)))
    print(model, diam, "image rel L2:", rel)
    assert rel <= util.TOL_IMG_L2


def test_save_lightrays_with_nothing_to_save_writes_empty_files(tmp_path, cuda_lib):
    """save_lightrays with num_lightrays_save = 0: the reference still writes its per-batch files, empty
    (prt.cu:3561-3611 with n = 0)."""
    c = small_bos(tmp_path, rays_per_point=8)
    c.source_point_number = 100
    d = dump_dir(tmp_path, "empty")
    c.save_lightrays = True
    c.num_lightrays_save = 0
    c.lightray_position_save_path = d
    c.lightray_direction_save_path = d
    img = c.run(cuda_lib)
    assert img.sum() > 0
    nb = -(-c.num_particles // 100)
    files = sorted(os.listdir(d))
    assert files == sorted(["pos_%04d.bin" % k for k in range(nb)] + ["dir_%04d.bin" % k for k in range(nb)])
    assert all(os.path.getsize(os.path.join(d, f)) == 0 for f in files)
