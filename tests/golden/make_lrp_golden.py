"""Generates tests/golden/lrp_golden.npz by running the reference's own
python_codes/light_ray_processing.py (imported from /root/reference) on seeded dump files.
Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_lrp_golden.py"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import ast  # noqa: E402
# The module's top level sets up matplotlib through the author's private modules (modify_plt_settings,
# loadmat_functions: not in the repository), so the function definitions are taken from the unmodified
# source with ast and executed on their own; only numpy and os are needed by the ones used here.
_src = open("/root/reference/python_codes/light_ray_processing.py").read()
_tree = ast.parse(_src)
_tree.body = [n for n in _tree.body if isinstance(n, ast.FunctionDef)]
class _NumpyCompat:
    """numpy as the reference (written for NumPy 1.x) calls it: reshape(a=..., newshape=...)."""
    def __getattr__(self, k):
        return getattr(np, k)

    @staticmethod
    def reshape(a, newshape):
        return np.reshape(a, newshape)


_ns = {"np": _NumpyCompat(), "os": os}
exec(compile(_tree, "/root/reference/python_codes/light_ray_processing.py", "exec"), _ns)
ref = type("ref", (), {k: staticmethod(v) for k, v in _ns.items() if callable(v) and not k.startswith("_")})

from tests.test_dot_stats import _write_dumps  # noqa: E402

seed, n, rpd = 20251104, 4000, 250
d = tempfile.mkdtemp()
_write_dumps(d, np.random.default_rng(seed), n)
pos1, pos2, dir1, dir2 = ref.load_light_ray_data(d)
d_pos, d_dir = ref.calculate_lightray_deflections(pos1, pos2, dir1, dir2)
dp = ref.calculate_dot_average(d_pos, rpd)
dd = ref.calculate_dot_average(d_dir, rpd)
pix = ref.convert_pos_to_pix(pos1, {"x": -100.0, "y": 50.0}, 17.0)
out = {"seed": seed, "n": n, "rays_per_dot": rpd, "pix_x": pix["x"]}
for c in "xyz":
    out["pos1_" + c] = pos1[c]; out["dir2_" + c] = dir2[c]
    out["dpos_dot_" + c] = dp[c]; out["ddir_dot_" + c] = dd[c]
np.savez_compressed(os.path.join(HERE, "lrp_golden.npz"), **out)
print("wrote lrp_golden.npz", {k: np.shape(v) for k, v in out.items()})
