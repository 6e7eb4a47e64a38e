"""Parameter-file front end (photon_b200/run_simulation.py) against the reference:

  * the call assembled from the reference's own sample ``.mat`` files equals, byte for byte, the call the
    reference's UNCHANGED scripts made for those files (tests/golden/recorded_*.npz, recorded at the C boundary);
  * lens solution and particle diameter distribution against the reference's own function bodies
    (tests/golden/frontend_golden.npz, made by tests/golden/make_frontend_golden.py);
  * GPU: ``run_simulation_02`` on the BOS sample file end to end (both calls, density field on the second),
    raw images against the images the reference's sources produced, 16-bit images against the numpy
    restatement of pr03.py:2189-2243.
"""
import copy
import os

import numpy as np
import pytest

from photon_b200 import recording
from photon_b200 import run_simulation as rs
from photon_b200 import source_points as sp
from tests import util

G = util.GOLDEN_DIR


def parameters(kind):
    return rs.loadmat(os.path.join(G, "sample-parameters-%s.mat" % kind))


def calibration_parameters():
    """The PIV sample's lens and camera (tilted by a degree or two) with a 9 x 7 calibration grid on 3 planes."""
    P = parameters("piv")
    P["simulation_type"] = "cal"
    P["camera_design"]["x_camera_angle"], P["camera_design"]["y_camera_angle"] = 0.02, -0.035
    P["calibration_grid"] = {"generate_calibration_grid_images": 1, "calibration_plane_number": 3, "calibration_plane_spacing": 2.5e3,
                             "grid_point_diameter": 1500.0, "x_grid_point_spacing": 8e3, "y_grid_point_spacing": 9e3,
                             "x_grid_point_number": 9, "y_grid_point_number": 7, "particle_number_per_grid_point": 160,
                             "lightray_number_per_particle": 50, "lightray_process_number": 1e6}
    return P


def target_image():
    """A 60 x 48 RGBA float image with a few grey dots, as mpimg.imread returns a PNG."""
    rng = np.random.RandomState(11)
    img = np.zeros((48, 60, 4), np.float32)
    for r, c in rng.randint(3, 45, size=(25, 2)):
        img[r - 1:r + 2, c - 1:c + 2, :3] = rng.rand()
    img[:, :, 3] = 1.0
    return img


def recorded_source(kind, P, optical_system, rec):
    block = P["bos_pattern" if kind == "bos" else "particle_field"]
    return {"x": rec.x, "y": rec.y, "z": rec.z, "radiance": rec.radiance, "diameter_index": rec.diameter_index,
            "lightray_number_per_particle": block["lightray_number_per_particle"], "source_point_number": 10000,
            "z_offset": sp.object_plane(P, optical_system) - P["lens_design"]["object_distance"],
            "object_distance": P["lens_design"]["object_distance"]}


def test_loadmat_gives_nested_dicts():
    P = parameters("piv")
    assert P["simulation_type"] == "piv" and P["lens_design"]["focal_length"] == 105000.0
    assert list(P["particle_field"]["frame_vector"]) == [1, 2]
    assert isinstance(P["density_gradients"], dict) and P["camera_design"]["x_pixel_number"] == 1024


@pytest.mark.parametrize("kind", ["bos", "piv"])
def test_assembled_call_equals_the_recorded_reference_call(kind):
    P = parameters(kind)
    optical_system = rs.create_camera_optical_system(P)
    rec, _ = recording.load_npz(os.path.join(G, "recorded_%s.npz" % kind))
    P["output_data"]["lightray_positions_filepath"] = P["output_data"]["lightray_directions_filepath"] = ""
    if kind == "piv":
        scattering, kind_str = rs.create_mie_scattering_data(P, np.random.RandomState(1)), "mie"
    else:
        scattering, kind_str = float("nan"), "diffuse"
    call = rs.prepare_call(P, optical_system, P["camera_design"]["pixel_gain"], scattering, kind_str,
                           recorded_source(kind, P, optical_system, rec))
    for k in rec.SCALARS:
        a, b = getattr(call, k), getattr(rec, k)
        if isinstance(b, float):
            assert np.float32(a) == np.float32(b), (k, a, b)         # the boundary carries float32
        else:
            assert a == b, (k, a, b)
    assert bytes(call.camera) == bytes(rec.camera)
    assert len(call.elements) == 1 and bytes(call.elements[0]) == bytes(rec.elements[0])
    assert np.array_equal(call.element_center, rec.element_center)
    assert np.array_equal(call.element_plane_parameters, rec.element_plane_parameters)
    assert np.array_equal(call.element_system_index, rec.element_system_index)
    if kind == "piv":
        assert np.array_equal(call.mie_angles, rec.mie_angles) and call.mie_irradiance.shape == (255, 27)
        assert np.abs(call.mie_irradiance / rec.mie_irradiance - 1).max() < 1e-6
        assert np.array_equal(call.mie_beam, rec.mie_beam)
        assert np.array_equal(call.mie_inverse_rotation, rec.mie_inverse_rotation)


def test_lens_solution_and_diameter_distribution_match_reference_functions():
    g = np.load(os.path.join(G, "frontend_golden.npz"))
    for kind in ("piv", "bos"):
        for model in ("general", "thin-lens"):
            P = parameters(kind)
            P["lens_design"]["lens_model"] = model
            lens = rs.create_camera_optical_system(P)["design"]["optical_element"][0]["optical_element"][0]
            geo, prop = lens["element_geometry"], lens["element_properties"]
            got = np.array([geo["pitch"], geo["front_surface_radius"], geo["back_surface_radius"], geo["vertex_distance"],
                            prop["refractive_index"], prop["thin_lens_focal_length"]])
            assert np.array_equal(got, g["%s_%s_lens" % (kind, model)]), (kind, model, got)
    P = parameters("piv")
    d, pdf = rs.calculate_particle_diameter_distribution(P)
    assert np.allclose(d, g["diameter_vector"], rtol=1e-13, atol=0) and np.allclose(pdf, g["diameter_pdf"], rtol=1e-12, atol=0)
    assert abs(pdf.sum() - 1) < 1e-12 and d.size == 27
    P["particle_field"]["particle_number"] = 5000
    idx = rs.calculate_particle_diameter_indices(P, g["diameter_pdf"], g["diameter_vector"], np.random.RandomState(int(g["seed"])))
    assert np.array_equal(idx, g["diameter_indices"])
    assert idx.max() <= 25                                   # the last bin is never assigned (rs02.py:624)


@pytest.mark.parametrize("plane", [0, 2])
def test_calibration_planes_match_reference_function(plane):
    """generate_calibration_lightfield_data against the reference's own function body (rs02.py:1057-1249)."""
    g = np.load(os.path.join(G, "frontend_golden.npz"))
    P = calibration_parameters()
    src = sp.generate_calibration_lightfield_data(P, rs.create_camera_optical_system(P), plane,
                                                  np.random.RandomState(int(g["seed"]) + plane))
    for k in ("x", "y", "z", "radiance", "diameter_index"):
        want = g["cal%d_%s" % (plane, k)]
        assert src[k].shape == want.shape, (k, src[k].shape, want.shape)
        assert np.allclose(src[k], want, rtol=0, atol=1e-6 if k in "xyz" else 0), (k, np.abs(src[k] - want).max())
    assert src["z_offset"] == float(g["cal%d_z_offset" % plane])
    assert src["x"].size > 9 * 7 * 160                         # the grid plus the two origin markers


def test_image_target_matches_reference_function():
    """generate_bos_image_lightfield_data against the reference's function body with imread stubbed (rs02.py:1554-1697)."""
    g = np.load(os.path.join(G, "frontend_golden.npz"))
    P = calibration_parameters()
    src = sp.generate_bos_image_lightfield_data(P, rs.create_camera_optical_system(P), target_image())
    for k in ("x", "y", "z", "radiance", "diameter_index"):
        want = g["img_%s" % k]
        assert src[k].shape == want.shape and np.allclose(src[k], want, rtol=0, atol=1e-6 if k in "xyz" else 0), k
    assert src["source_point_number"] == int(g["img_source_point_number"]) == src["x"].size < 10000


def test_unknown_simulation_type_is_refused():
    P = parameters("bos")
    P["simulation_type"] = "holography"
    with pytest.raises(ValueError):
        rs.run_simulation_02(P, write_images=False)


@pytest.mark.gpu
def test_calibration_run_end_to_end(cuda_lib, tmp_path):
    """'cal' branch: three planes, each image shows the dot grid (about 9 x 7 + 2 bright blobs) and the planes differ."""
    P = calibration_parameters()
    P["output_data"]["image_directory"] = str(tmp_path)
    P["output_data"]["save_lightrays"] = 0
    results = rs.run_simulation_02(P, rng=np.random.RandomState(5))
    assert [r[0] for r in results] == ["calibration_image_plane_%04d" % i for i in (1, 2, 3)]
    sums = []
    for name, I, I_raw in results:
        assert I.dtype == np.uint16 and I.max() == 65535 and I_raw.sum() > 0
        lit = (I_raw > 0.05 * I_raw.max())
        assert 0.001 < lit.mean() < 0.2, lit.mean()              # dots, not a flood
        sums.append(float(I_raw.astype(np.float64).sum()))
    assert not np.array_equal(results[0][2], results[2][2])


@pytest.mark.gpu
def test_bos_sample_file_end_to_end(cuda_lib, tmp_path, monkeypatch):
    """`sample_run_script.sh bos` through the front end: the recorded dot pattern stands in for the random
    one (the reference draws it from numpy's global generator), everything else comes from the .mat file."""
    from oracle import postprocess
    P = parameters("bos")
    P["density_gradients"]["density_gradient_filename"] = os.path.join(G, "sample-density.nrrd")
    P["output_data"]["image_directory"] = str(tmp_path)
    rec, ex = recording.load_npz(os.path.join(G, "recorded_bos.npz"))

    def recorded_dots(simulation_parameters, optical_system, rng=None):
        return recorded_source("bos", simulation_parameters, optical_system, rec), np.zeros(1), np.zeros(1)

    monkeypatch.setattr(sp, "generate_bos_lightfield_data", recorded_dots)
    results = rs.run_simulation_02(copy.deepcopy(P))
    assert [r[0] for r in results] == ["bos_pattern_image_1", "bos_pattern_image_2"]
    for i, (name, I, I_raw) in enumerate(results):
        ref = ex["ref_image%d" % i]
        m = float(np.linalg.norm(I_raw.astype(np.float64) - ref) / np.linalg.norm(ref.astype(np.float64)))
        assert m <= util.TOL_IMG_L2, (name, m)
        on_disk = np.fromfile(os.path.join(str(tmp_path), "raw", name + ".bin"), np.float32).reshape(I_raw.shape)
        assert np.array_equal(on_disk, I_raw)
        want = postprocess.postprocess_u16(I_raw, P["camera_design"]["pixel_gain"], int(P["camera_design"]["pixel_bit_depth"]))
        assert I.dtype == np.uint16 and np.array_equal(I, want), name
        assert I.max() == 65535
