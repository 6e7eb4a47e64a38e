"""source-point generation through the C ABI against the reference's own generator functions
(tests/golden/source_points_golden.npz, made by make_source_golden.py with numpy's stream pinned to
RandomState(SEED)).  Host code: runs without a GPU."""
import copy
import os

import numpy as np

from photon_b200 import source_points as sp

SEED = 4242
OPTICAL_SYSTEM = {"design": {"optical_element": [{"optical_element": [{
    "element_properties": {"refractive_index": 1.5},
    "element_geometry": {"front_surface_radius": 1e5, "back_surface_radius": -1e5, "vertex_distance": 1e4}}]}]}}


def bos_parameters(kind, n):
    return {"lens_design": {"focal_length": 105e3, "object_distance": 1e6},
            "camera_design": {"implement_diffraction": True, "diffraction_diameter": 3.0, "pixel_pitch": 10.0,
                              "x_camera_angle": 0.0, "y_camera_angle": 0.0},
            "bos_pattern": {"X_Min": -2e4, "X_Max": 2e4, "Y_Min": -1.5e4, "Y_Max": 1.5e4, "grid_point_number": max(n, 2),
                            "grid_point_diameter": 300.0, "particle_number_per_grid_point": 50,
                            "dot_overlap": kind == "overlap", "dot_distribution": "regular" if kind == "regular" else "random",
                            "dot_spacing": 12.0}}


def piv_parameters(n):
    return {"lens_design": {"focal_length": 105e3, "object_distance": 7e5},
            "camera_design": {"x_camera_angle": 0.02, "y_camera_angle": -0.035, "pixel_pitch": 17.0},
            "particle_field": {"data_directory": "", "data_filename_prefix": "", "particle_number": n, "gaussian_beam_fwhm": 7.5e3,
                               "perform_mie_scattering": False, "load_particle_data": False, "lightray_number_per_particle": 100,
                               "Z_Min": -5e3, "Z_Max": 5e3}}


G = None


def golden():
    global G
    if G is None:
        G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "source_points_golden.npz"))
    return G


def test_sunflower_matches_reference():
    x, y = sp.calculate_sunflower_coordinates(300.0, 100, np.random.RandomState(SEED))
    g = golden()
    assert x.shape == g["sun_x"].shape
    np.testing.assert_allclose(x, g["sun_x"], rtol=0, atol=1e-10)
    np.testing.assert_allclose(y, g["sun_y"], rtol=0, atol=1e-10)
    assert x[-1] == 0.0 and y[-1] == 0.0


def test_non_overlapping_dots_match_reference():
    c = sp.create_non_overlapping_dot_coordinates(bos_parameters("random", 400), np.random.RandomState(SEED))
    np.testing.assert_array_equal(c, golden()["dots"])           # same draws, same accept/reject decisions
    d = np.sqrt(((c[:, None, :] - c[None, :, :]) ** 2).sum(-1)) + np.eye(len(c)) * 1e9
    assert d.min() > 1.5 * 300.0


def test_bos_lightfield_matches_reference():
    g = golden()
    for tag, n in (("regular", 0), ("random", 150), ("overlap", 200)):
        src, gx, gy = sp.generate_bos_lightfield_data(copy.deepcopy(bos_parameters(tag, n)), OPTICAL_SYSTEM, np.random.RandomState(SEED))
        np.testing.assert_allclose(gx, g["bos_%s_gx" % tag], rtol=0, atol=1e-9)
        np.testing.assert_allclose(gy, g["bos_%s_gy" % tag], rtol=0, atol=1e-9)
        for k in ("x", "y", "z"):                                # the ABI arrays are float32 (pr03.py:1733-1736)
            ref = g["bos_%s_%s" % (tag, k)].astype(np.float32)
            assert src[k].shape == ref.shape, (tag, k)
            np.testing.assert_allclose(src[k], ref, rtol=0, atol=np.abs(ref).max() * 2e-7)
        np.testing.assert_array_equal(src["radiance"], g["bos_%s_radiance" % tag])
        assert src["z_offset"] == float(g["bos_%s_z_offset" % tag])
        assert src["x"].dtype == np.float32 and src["radiance"].dtype == np.float64 and src["diameter_index"].dtype == np.int32


def test_piv_particles_match_reference():
    g = golden()
    src = sp.generate_piv_lightfield_data(piv_parameters(5000), OPTICAL_SYSTEM, g["piv_rotation"], np.random.RandomState(SEED))
    for k in ("x", "y", "z"):
        ref = np.asarray(g["piv_" + k]).reshape(-1).astype(np.float32)
        np.testing.assert_allclose(src[k], ref, rtol=0, atol=np.abs(ref).max() * 2e-7)
    np.testing.assert_allclose(src["radiance"], g["piv_radiance"], rtol=1e-13)
    assert src["z_offset"] == float(g["piv_z_offset"])
