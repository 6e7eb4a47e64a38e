"""Generates tests/golden/source_points_golden.npz from the reference's own generator functions in
python_codes/run_simulation_02.py (unmodified bodies taken with ast; numpy's global generator replaced by
a seeded RandomState and its clock re-seed ``np.random.seed()`` made a no-op, so the stream is pinned).
Run in the build container only:  python tests/golden/make_source_golden.py"""
import ast
import copy
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.test_source_points import bos_parameters, piv_parameters, OPTICAL_SYSTEM, SEED  # noqa: E402

PATH = "/root/reference/python_codes/run_simulation_02.py"
NAMES = {"calculate_sunflower_coordinates", "create_non_overlapping_dot_coordinates", "generate_bos_lightfield_data",
         "load_lightfield_data", "rotate_coordinates", "calculate_rotation_matrix"}


class PinnedRandom:
    def __init__(self, seed):
        self.rs = np.random.RandomState(seed)

    def seed(self, *a):
        pass

    def rand(self, *a):
        return self.rs.rand(*a)


class NumpyPinned:
    def __init__(self, seed):
        self.random = PinnedRandom(seed)

    def __getattr__(self, k):
        return getattr(np, k)


def load(seed):
    tree = ast.parse(open(PATH).read())
    tree.body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in NAMES]
    ns = {"np": NumpyPinned(seed), "print": lambda *a, **k: None}
    exec(compile(tree, PATH, "exec"), ns)
    return ns


out = {}
ns = load(SEED)
x, y = ns["calculate_sunflower_coordinates"](300.0, 100)
out["sun_x"], out["sun_y"] = np.asarray(x, np.float64).reshape(-1), np.asarray(y, np.float64).reshape(-1)
ns = load(SEED)
out["dots"] = ns["create_non_overlapping_dot_coordinates"](bos_parameters("random", 400))
for tag, sp in (("regular", bos_parameters("regular", 0)), ("random", bos_parameters("random", 150)), ("overlap", bos_parameters("overlap", 200))):
    ns = load(SEED)
    src, gx, gy = ns["generate_bos_lightfield_data"](copy.deepcopy(sp), OPTICAL_SYSTEM)
    for k in ("x", "y", "z", "radiance", "diameter_index"):
        out["bos_%s_%s" % (tag, k)] = np.asarray(src[k], np.float64)
    out["bos_%s_gx" % tag], out["bos_%s_gy" % tag] = gx, gy
    out["bos_%s_z_offset" % tag] = src["z_offset"]
ns = load(SEED)
src = ns["load_lightfield_data"](piv_parameters(5000), OPTICAL_SYSTEM, None, 1, {})
for k in ("x", "y", "z", "radiance"):
    out["piv_" + k] = np.asarray(src[k], np.float64)
out["piv_z_offset"] = src["z_offset"]
out["piv_rotation"] = ns["calculate_rotation_matrix"](0.02, -0.035, 0.0)
np.savez_compressed(os.path.join(HERE, "source_points_golden.npz"), **out)
print({k: np.shape(v) for k, v in out.items()})
