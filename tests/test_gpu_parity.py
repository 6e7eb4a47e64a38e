"""GPU parity: the CUDA library against the CPU oracle on identical inputs, through the C ABI
(start_ray_tracing with host buffers, ray dumps on).  Tolerances are the north-star's."""
import numpy as np
import pytest

from photon_b200 import scenes
from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cases(tmp_path_factory):
    return util.parity_cases(tmp_path_factory.mktemp("parity"))


CASE_NAMES = ["bos_rk4_thin_diffraction", "bos_euler_thin_diffraction", "bos_nogradient",
              "bos_rk4_multibatch_ragged", "piv_mie_thick_bilinear", "piv_diffuse_thin_bilinear",
              "piv_apparent", "piv_mie_thick_diffraction", "piv_chief_ray_only",
              "bos_rk45_thin_diffraction", "bos_ab4_thin_diffraction"]


@pytest.mark.parametrize("name", CASE_NAMES)
def test_cuda_matches_oracle(name, cases, cuda_lib, oracle_lib, tmp_path):
    call, pitch = cases[name]
    (tmp_path / "o").mkdir()
    (tmp_path / "g").mkdir()
    ref = util.run_with_dumps(oracle_lib, call, str(tmp_path / "o"))
    got = util.run_with_dumps(cuda_lib, call, str(tmp_path / "g"))
    m = util.compare(ref, got, pitch)
    print(name, m)
    util.assert_parity(name, m)


@pytest.mark.parametrize("name", CASE_NAMES)
def test_cuda_matches_reference_golden(name, cases, cuda_lib, tmp_path):
    """Against tests/golden/*.npz: outputs of the reference's own sources compiled for the host
    (tests/golden/make_golden.py), so this needs neither the oracle nor /root/reference."""
    call, pitch = cases[name]
    g = util.load_golden(name)
    img, pos, dr = util.run_with_dumps(cuda_lib, call, str(tmp_path))
    assert len(pos) == int(g["rays"])
    n = util.GOLDEN_RAYS
    m = util.compare((img, g["pos"], g["dir"]), (img, pos[:n], dr[:n]), pitch)
    m["hit_fraction"] = float(np.isfinite(pos[:, 0]).mean())
    util.assert_parity(name, m)
    assert abs(int(np.isfinite(pos[:, 0]).sum()) - int(g["hits"])) <= max(2, len(pos) // 5000)
    tiles = util.tile_sums(img)
    rel = np.linalg.norm(tiles - g["tiles"]) / np.linalg.norm(g["tiles"])
    assert rel <= util.TOL_IMG_L2, rel
    assert abs(float(img.astype(np.float64).sum()) - float(g["image_sum"])) <= 1e-5 * float(g["image_sum"])


def test_image_accumulates_onto_input(cases, cuda_lib):
    """The kernel adds onto whatever the caller passes in image_array (prt.cu:3309)."""
    call, _ = cases["bos_nogradient"]
    base = call.run(cuda_lib)
    pre = np.full(base.shape, 0.25, np.float32)
    again = call.run(cuda_lib, pre)
    np.testing.assert_allclose(again, base + 0.25, rtol=1e-5, atol=1e-6)


def test_texture_sampler_close_to_linear(cases, cuda_lib, monkeypatch, tmp_path):
    """Hardware tex3D (9-bit weights) against the software trilinear path."""
    call, pitch = cases["bos_rk4_thin_diffraction"]
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    lin = util.run_with_dumps(cuda_lib, call, str(tmp_path / "a"))
    monkeypatch.setenv("PHOTON_SAMPLER", "texture")
    tex = util.run_with_dumps(cuda_lib, call, str(tmp_path / "b"))
    m = util.compare(lin, tex, pitch)
    print("texture vs linear", m)
    assert m["dpos_px"] <= 5e-3 and m["ddir_rad"] <= 1e-5, m


def _intermediate(d, kind):
    import glob, os
    return np.concatenate([np.fromfile(f, np.float32).reshape(-1, 3)
                           for f in sorted(glob.glob(os.path.join(d, "intermediate_%s_*.bin" % kind)))])


@pytest.mark.parametrize("name", ["bos_rk4_thin_diffraction", "bos_euler_thin_diffraction"])
def test_generic_marcher_and_intermediate_dumps(name, cases, cuda_lib, oracle_lib, tmp_path):
    """With final-ray dumps only, the marchers run their hot loops (rk4_hot_loop / euler_hot_loop -- the code the
    benchmark times), which is what every other test in this file compares with the oracle.  Intermediate dumps
    (prt.cu:3613-3670) force every step through the generic path: the same scenes that way, the first positions
    and directions along each ray included, and the two marchers against each other."""
    call, pitch = cases[name]
    for k in "ogh":
        (tmp_path / k).mkdir()
    hot = util.run_with_dumps(cuda_lib, call, str(tmp_path / "h"))
    call.save_intermediate_ray_data = True
    call.num_intermediate_positions_save = 6
    try:
        ref = util.run_with_dumps(oracle_lib, call, str(tmp_path / "o"))
        got = util.run_with_dumps(cuda_lib, call, str(tmp_path / "g"))
    finally:
        call.save_intermediate_ray_data = False
        call.num_intermediate_positions_save = 0
    util.assert_parity(name + " generic", util.compare(ref, got, pitch))
    util.assert_parity(name + " hot vs generic", util.compare(got, hot, pitch))
    for kind, tol in (("pos", 0.25), ("dir", 1e-6)):                     # um (positions are ~1e5..1e6 um: 4 ulp), -
        a, b = _intermediate(str(tmp_path / "o"), kind), _intermediate(str(tmp_path / "g"), kind)
        assert a.shape == b.shape and len(a) == 6 * len(ref[1])
        assert np.array_equal(np.isnan(a), np.isnan(b))
        ok = np.isfinite(a)
        assert ok.mean() > 0.5 and np.abs(a[ok] - b[ok]).max() <= tol, (kind, np.abs(a[ok] - b[ok]).max())


# ---- the kernels the benchmark times --------------------------------------------------------------------------------
# Ray dumps run trace_rays_kernel<DUMP = true, 128 registers>; the benchmark and every production call run the
# non-dump instantiation (<false, 3, ...>, 80 registers, its own ptxas schedule).  Its only output is the image, so
# the image is what is compared with the oracle's, for every scene that marches through a field.
FIELD_CASES = ["bos_rk4_thin_diffraction", "bos_euler_thin_diffraction", "bos_rk4_multibatch_ragged",
               "bos_rk45_thin_diffraction", "bos_ab4_thin_diffraction"]


def _rel_l2(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


@pytest.mark.parametrize("name", FIELD_CASES)
def test_nondump_kernel_image_matches_oracle(name, cases, cuda_lib, oracle_lib):
    call, _ = cases[name]
    assert not call.save_lightrays
    ref = call.run(oracle_lib)
    got = call.run(cuda_lib)
    variant = cuda_lib.photon_b200_last_variant().decode()
    m = _rel_l2(got, ref)
    print(name, variant, "image rel L2 vs oracle:", m)
    assert variant.endswith("/image"), variant
    assert m <= (5e-3 if name in util.AB4_CASES else util.TOL_IMG_L2), (name, variant, m)


GATHERERS = {"cells": {}, "coherent": {"PHOTON_CELLS": "0"}, "fanout": {"PHOTON_HOT": "4"}}


@pytest.mark.parametrize("dump", [True, False], ids=["dump", "image"])
@pytest.mark.parametrize("gatherer", list(GATHERERS))
def test_every_rk4_gatherer_dump_and_image(gatherer, dump, cases, cuda_lib, oracle_lib, tmp_path, monkeypatch):
    """The three RK4 hot loops the driver can pick for a linear field -- per-cell polynomial table, coherent corner
    texels (what a coherent scene runs when the table does not fit: C3 at 1024^3, all of C5), fan-out corner texels
    (PIV through a field) -- each through its dump kernel (per-ray parity) and its non-dump kernel (image parity).
    photon_b200_last_variant proves which kernel ran."""
    call, pitch = cases["bos_rk4_thin_diffraction"]
    for k, v in GATHERERS[gatherer].items():
        monkeypatch.setenv(k, v)
    (tmp_path / "o").mkdir(); (tmp_path / "g").mkdir()
    if dump:
        ref = util.run_with_dumps(oracle_lib, call, str(tmp_path / "o"))
        got = util.run_with_dumps(cuda_lib, call, str(tmp_path / "g"))
        variant = cuda_lib.photon_b200_last_variant().decode()
        m = util.compare(ref, got, pitch)
        print(gatherer, variant, m)
        assert variant == "rk4/%s/dump" % gatherer, variant
        util.assert_parity("bos_rk4_thin_diffraction", m)
    else:
        ref = call.run(oracle_lib)
        got = call.run(cuda_lib)
        variant = cuda_lib.photon_b200_last_variant().decode()
        m = _rel_l2(got, ref)
        print(gatherer, variant, "image rel L2 vs oracle:", m)
        assert variant == "rk4/%s/image" % gatherer, variant
        assert m <= util.TOL_IMG_L2, (variant, m)


def test_c4_shaped_scene_matches_oracle(cuda_lib, oracle_lib, tmp_path):
    """BASELINE C4's combination at test size: Mie table from bhmie, thick lens, bilinear splat, turbulent field,
    rays that fan out over many cells (cone ratio 1) -> the fan-out gatherer, dump and non-dump kernels."""
    from photon_b200 import scenes
    call = scenes.c4_piv(str(tmp_path / "turb.nrrd"), n=48, particles=300, rays_per_particle=200, nx=1024, ny=1024)
    call.source_point_number = call.num_particles
    (tmp_path / "o").mkdir(); (tmp_path / "g").mkdir()
    ref = util.run_with_dumps(oracle_lib, call, str(tmp_path / "o"))
    got = util.run_with_dumps(cuda_lib, call, str(tmp_path / "g"))
    variant = cuda_lib.photon_b200_last_variant().decode()
    m = util.compare(ref, got, 17.0)
    print("C4-shaped", variant, m)
    assert variant == "rk4/fanout/dump", variant
    util.assert_parity("piv_mie_thick_bilinear", m)                      # thick lens: 1e-3 px at the 99.9th percentile
    plain = call.run(cuda_lib)
    variant = cuda_lib.photon_b200_last_variant().decode()
    assert variant == "rk4/fanout/image", variant
    assert _rel_l2(plain, ref[0]) <= util.TOL_IMG_L2
    # the field matters in this scene: without it the hits move
    call.simulate_density_gradients = False
    (tmp_path / "n").mkdir()
    nof = util.run_with_dumps(oracle_lib, call, str(tmp_path / "n"))
    both = np.isfinite(ref[1][:, 0]) & np.isfinite(nof[1][:, 0])
    assert np.abs(ref[1][both] - nof[1][both])[:, :2].max() > 1e-3


NOFIELD_CASES = ["piv_mie_thick_bilinear", "piv_diffuse_thin_bilinear", "piv_apparent", "piv_mie_thick_diffraction",
                 "bos_nogradient", "piv_chief_ray_only"]


@pytest.mark.parametrize("name", NOFIELD_CASES)
def test_queued_nofield_kernel_matches_fused_kernel_and_oracle(name, cases, cuda_lib, oracle_lib, monkeypatch):
    """Scenes without a field run trace_nofield_kernel on the plain image launch: rays that pass the first aperture
    are queued in shared memory and walk the lens system, the sensor and the splat in full warps.  Every ray goes
    through the same functions as in the fused kernel (PHOTON_QUEUE=0), so the images may differ by float summation
    order only, the same rays must reach the sensor, and the image must sit within the north-star bound of the
    oracle's."""
    from photon_b200.scene import ResidentScene
    call, _ = cases[name]
    assert not call.save_lightrays and not call.simulate_density_gradients
    ref = call.run(oracle_lib)
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("PHOTON_QUEUE", mode)
        sc = ResidentScene(call)
        sc.trace()
        out[mode] = (np.array(sc.download_image(), np.float64), sc.sensor_hits, cuda_lib.photon_b200_last_variant().decode())
        sc.close()
    (img_q, hits_q, var_q), (img_f, hits_f, var_f) = out["1"], out["0"]
    assert var_q == "nofield/queue/image" and var_f == "nofield/none/image", (var_q, var_f)
    assert hits_q == hits_f and hits_q > 0
    d = _rel_l2(img_q, img_f)
    print(name, "queued vs fused:", d, "hits", hits_q, "vs oracle:", _rel_l2(img_q, ref))
    assert d <= 2e-6, (name, d)
    assert _rel_l2(img_q, ref) <= util.TOL_IMG_L2


def test_queued_nofield_kernel_ragged_counts_and_position_noise(cuda_lib, monkeypatch):
    """Ray counts that fill the queue unevenly (rpp = 1, 255, 256, 257, 1000: partial chunks, drains of fewer than 256
    rays, several drains per point) and the position-noise stream, which is keyed by the ray's own number: the queued
    kernel must give every ray the draws the fused kernel gives it."""
    from photon_b200.scene import ResidentScene
    for rpp, noise in ((1, 0.0), (255, 0.0), (256, 0.0), (257, 0.3), (1000, 0.3)):
        call = scenes.c1_piv(particles=64, rays_per_particle=rpp)
        if noise:
            call.add_pos_noise = True
            call.pos_noise_std = noise
        imgs = {}
        for mode in ("1", "0"):
            monkeypatch.setenv("PHOTON_QUEUE", mode)
            sc = ResidentScene(call)
            sc.trace()
            imgs[mode] = (np.array(sc.download_image(), np.float64), sc.sensor_hits)
            sc.close()
        assert imgs["1"][1] == imgs["0"][1], (rpp, imgs["1"][1], imgs["0"][1])
        if imgs["0"][1]:
            assert _rel_l2(imgs["1"][0], imgs["0"][0]) <= 2e-6, rpp
