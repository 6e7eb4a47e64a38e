"""CPU: the product's per-ray stage code (photon_b200/csrc/*.cuh compiled for the host by
tests/hostcheck) against the oracle.  Without FMA contraction the two must agree bit for bit
on ray dumps; this catches arithmetic slips before any GPU time is spent."""
import numpy as np
import pytest

from tests import util


@pytest.fixture(scope="module")
def cases(tmp_path_factory):
    return util.parity_cases(tmp_path_factory.mktemp("hostcheck"))


@pytest.mark.parametrize("name", ["bos_rk4_thin_diffraction", "bos_euler_thin_diffraction", "bos_nogradient",
                                  "bos_rk4_multibatch_ragged", "piv_mie_thick_bilinear",
                                  "piv_diffuse_thin_bilinear", "piv_apparent", "piv_mie_thick_diffraction",
                                  "piv_chief_ray_only", "bos_rk45_thin_diffraction", "bos_ab4_thin_diffraction"])
def test_stage_code_bit_exact_vs_oracle(name, cases, hostcheck_lib, oracle_lib, tmp_path):
    call, pitch = cases[name]
    (tmp_path / "o").mkdir()
    (tmp_path / "h").mkdir()
    ref = util.run_with_dumps(oracle_lib, call, str(tmp_path / "o"))
    got = util.run_with_dumps(hostcheck_lib, call, str(tmp_path / "h"))
    assert np.array_equal(ref[1], got[1], equal_nan=True)
    assert np.array_equal(ref[2], got[2], equal_nan=True)
    m = util.compare(ref, got, pitch)
    assert m["image_rel_l2"] <= 1e-7, m


def test_thick_lens_conditioning(cases, oracle_lib, tmp_path):
    """Measures how far a ONE-ulp change of the generated tangents (what a different libm does to
    prt.cu:128-130,228) moves the sensor hits in the reference's float32 arithmetic.  This is the
    justification for tests/util.THICK_LENS_MAX_PX: thick-lens scenes cannot be reproduced to
    1e-3 px on every ray by ANY implementation whose tanf/atanf differ in the last bit."""
    import ctypes
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = os.path.join(root, "tests", "hostcheck", "hostcheck.cpp")
    out = str(tmp_path / "libhostcheck_perturbed.so")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([gxx, "-O2", "-std=c++17", "-DPH_NO_FMA", "-DPH_PERTURB_GEN", "-ffp-contract=off", "-fPIC", "-shared",
                    "-x", "c++", "-o", out, src, "-lz"], check=True, capture_output=True)
    pert = ctypes.CDLL(out)
    res = {}
    for name in ("piv_mie_thick_bilinear", "piv_diffuse_thin_bilinear", "bos_rk4_thin_diffraction"):
        call, pitch = cases[name]
        d0 = tmp_path / (name + "_o"); d0.mkdir()
        d1 = tmp_path / (name + "_p"); d1.mkdir()
        ref = util.run_with_dumps(oracle_lib, call, str(d0))
        got = util.run_with_dumps(pert, call, str(d1))
        res[name] = util.compare(ref, got, pitch)
    print({k: (v["dpos_px"], v["dpos_px_p999"]) for k, v in res.items()})
    assert res["piv_diffuse_thin_bilinear"]["dpos_px"] <= util.TOL_POS_PX       # thin lens / BOS: well conditioned
    assert res["bos_rk4_thin_diffraction"]["dpos_px"] <= util.TOL_POS_PX
    thick = res["piv_mie_thick_bilinear"]
    assert thick["dpos_px_p999"] <= util.TOL_POS_PX and thick["dpos_px"] <= util.THICK_LENS_MAX_PX


@pytest.mark.parametrize("model,diam", [("apparent", 8.0), ("general", 9.0)])
def test_large_footprints_on_the_host(model, diam, hostcheck_lib, oracle_lib):
    """Footprints of more than 8 rows leave the register path of erf_splat (ray_optics.cuh) for one pair of erf per
    pixel; the increments stay the oracle's (same expression, same order)."""
    from photon_b200 import scenes
    c = scenes.c1_piv(particles=60, rays_per_particle=64, lens_model=model, nx=256, ny=256)
    c.camera.diffraction_diameter = diam
    c.camera.implement_diffraction = True
    ref = c.run(oracle_lib)
    got = c.run(hostcheck_lib)
    assert ref.sum() > 0
    assert float(np.linalg.norm(got.astype(np.float64) - ref) / np.linalg.norm(ref.astype(np.float64))) <= 1e-7


@pytest.mark.parametrize("kind", ["thick", "thin", "three_elements"])
def test_first_aperture_test_agrees_with_the_optical_train(kind, hostcheck_lib):
    """trace_nofield_kernel (scenes without a field) queues only the rays that pass first_aperture_culls
    (ray_optics.cuh), which restates the first cull of refract_element on its own.  On the host, over every ray of a
    scene: a ray it culls never yields a sensor sample, it does cull about half of all rays (the lens samples cover
    twice the aperture's radius), and for a thin lens -- whose aperture test is the ONLY way to die in the train -- no
    ray it lets through dies there: the two decisions are the same decision."""
    import ctypes as C
    from photon_b200 import scenes
    if kind == "thin":
        call = scenes.c1_piv(particles=150, rays_per_particle=400, mie=False, lens_model="thin-lens")
    else:
        call = scenes.c1_piv(particles=150, rays_per_particle=400)
        if kind == "three_elements":
            scenes.three_element_system(call)
    cnt = {n: C.c_longlong.in_dll(hostcheck_lib, "g_hostcheck_" + n) for n in ("culled", "cull_violations", "uncull_dead_in_train")}
    for c in cnt.values():
        c.value = 0
    img = call.run(hostcheck_lib)
    rays = call.num_particles * call.lightray_number_per_particle
    assert img.sum() > 0
    assert cnt["cull_violations"].value == 0
    assert 0.3 * rays < cnt["culled"].value < 0.7 * rays, (cnt["culled"].value, rays)
    if kind == "thin":
        assert cnt["uncull_dead_in_train"].value == 0
