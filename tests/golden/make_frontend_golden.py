"""Generates tests/golden/frontend_golden.npz from the reference's own function bodies in
python_codes/run_simulation_02.py (taken unmodified with ast; numpy's global generator replaced by a seeded
RandomState): the single-lens optical system solved from the two sample parameter files, and the log-normal
particle diameter distribution / index draw of the PIV sample, and two planes of the calibration target
(generate_calibration_lightfield_data) for the grid in ``calibration_parameters()``.
Run in the build container only:  python tests/golden/make_frontend_golden.py"""
import ast
import os
import sys

import numpy as np
import scipy.special as scispec

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from photon_b200.run_simulation import loadmat  # noqa: E402

PATH = "/root/reference/python_codes/run_simulation_02.py"
NAMES = {"create_single_lens_optical_system", "create_camera_optical_system", "log_normal_pdf", "inverse_log_normal_pdf",
         "log_normal_cdf", "calculate_log_normal_pdf_extrema", "calculate_particle_diameter_distribution",
         "calculate_particle_diameter_indices", "calculate_sunflower_coordinates", "rotate_coordinates",
         "calculate_rotation_matrix", "generate_calibration_lightfield_data", "generate_bos_image_lightfield_data"}
SEED = 77


class PinnedRandom:
    def __init__(self, seed):
        self.rs = np.random.RandomState(seed)

    def rand(self, *a):
        return self.rs.rand(*a)


class NumpyShim:
    NAN = float("nan")
    infty = float("inf")

    def __init__(self, seed):
        self.random = PinnedRandom(seed)

    def __getattr__(self, k):
        return getattr(np, k)


tree = ast.parse(open(PATH).read())
tree.body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in NAMES]
ns = {"np": NumpyShim(SEED), "scispec": scispec, "print": lambda *a, **k: None}
exec(compile(tree, PATH, "exec"), ns)

out = {}
for kind in ("piv", "bos"):
    P = loadmat(os.path.join(HERE, "sample-parameters-%s.mat" % kind))
    for model in ("general", "thin-lens"):
        P["lens_design"]["lens_model"] = model
        lens = ns["create_camera_optical_system"](P)["design"]["optical_element"][0]["optical_element"][0]
        g, p = lens["element_geometry"], lens["element_properties"]
        out["%s_%s_lens" % (kind, model)] = np.array([g["pitch"], g["front_surface_radius"], g["back_surface_radius"],
                                                      g["vertex_distance"], p["refractive_index"], p["thin_lens_focal_length"]])
P = loadmat(os.path.join(HERE, "sample-parameters-piv.mat"))
d, pdf = ns["calculate_particle_diameter_distribution"](P)
out["diameter_vector"], out["diameter_pdf"] = d, pdf
P["particle_field"]["particle_number"] = 5000
out["diameter_indices"] = ns["calculate_particle_diameter_indices"](P, pdf, d)
from tests.test_run_simulation import calibration_parameters  # noqa: E402
for plane in (0, 2):
    ns["np"].random = PinnedRandom(SEED + plane)
    C = calibration_parameters()
    src = ns["generate_calibration_lightfield_data"](C, ns["create_camera_optical_system"](C), plane)
    for k in ("x", "y", "z", "radiance", "diameter_index"):
        out["cal%d_%s" % (plane, k)] = np.asarray(src[k], np.float64).reshape(-1)      # np.matrix rows -> flat
    out["cal%d_z_offset" % plane] = src["z_offset"]
from tests.test_run_simulation import target_image  # noqa: E402


class ImreadStub:                                   # the reference reads a PNG it does not ship
    @staticmethod
    def imread(filename):
        return target_image()


ns["mpimg"] = ImreadStub
C = calibration_parameters()
src = ns["generate_bos_image_lightfield_data"](C, ns["create_camera_optical_system"](C))
for k in ("x", "y", "z", "radiance", "diameter_index"):
    out["img_%s" % k] = np.asarray(src[k], np.float64).reshape(-1)
out["img_source_point_number"] = src["source_point_number"]
out["seed"] = SEED
np.savez_compressed(os.path.join(HERE, "frontend_golden.npz"), **out)
print({k: np.asarray(v).shape for k, v in out.items()})
