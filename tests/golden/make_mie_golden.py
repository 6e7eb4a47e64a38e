"""Generates tests/golden/bhmie_golden.npz from the reference's own python_codes/bhmie.py and
run_simulation_02.py::calculate_mie_scattering_intensity (unmodified function bodies, taken from the
source files with ast).  Run in the build container only:  python tests/golden/make_mie_golden.py

Why ast: bhmie.py starts with ``from numpy import *``; from NumPy 2 on that also imports numpy's
``max`` (an array reduction whose second argument is an axis) over the builtin the code means at
bhmie.py:50 ``max(xstop, ymod)``.  The functions are executed in a namespace that holds numpy's names
as NumPy 1.x exported them (builtin max/min/abs/round/sum kept)."""
import ast
import builtins
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/python_codes"


def functions_of(path, names, ns):
    tree = ast.parse(open(path).read())
    tree.body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    exec(compile(tree, path, "exec"), ns)
    return ns


ns = {k: getattr(np, k) for k in np.__all__ if hasattr(np, k)}
for k in ("max", "min", "abs", "round", "sum", "any", "all", "pow", "bool", "int", "float", "complex"):
    ns[k] = getattr(builtins, k)
ns["abs"] = np.abs          # bhmie.py uses abs on complex arrays and scalars: numpy's abs handles both
functions_of(os.path.join(REF, "bhmie.py"), {"bhmie"}, ns)
ref_bhmie = ns["bhmie"]
ns2 = {"np": np, "bhmie": ref_bhmie, "abs": np.abs}
functions_of(os.path.join(REF, "run_simulation_02.py"), {"calculate_mie_scattering_intensity"}, ns2)
ref_table = ns2["calculate_mie_scattering_intensity"]

cases = [(3.0, 1.33, 128), (0.5, 1.5, 16), (25.0, 1.46 + 0.01j, 200), (11.81, 1.0904, 128), (0.05, 1.6, 2), (120.0, 1.2, 300)]
out = {"cases_x": np.array([c[0] for c in cases]), "cases_m": np.array([complex(c[1]) for c in cases]),
       "cases_nang": np.array([c[2] for c in cases])}
for i, (x, m, nang) in enumerate(cases):
    s1, s2, qext, qsca, qback, gsca, theta = ref_bhmie(x, m, nang)
    out["s1_%d" % i] = s1; out["s2_%d" % i] = s2; out["q_%d" % i] = np.array([qext, qsca, qback, gsca]); out["theta_%d" % i] = theta
# the table of a PIV run: 27 diameters, 128 angles per quadrant (255 rows), olive oil in air, 532 nm
params = {"particle_field": {"medium_refractive_index": 1.0, "particle_refractive_index": 1.46,
                             "mie_scattering_angle_number": 128, "beam_wavelength": 0.532}}
diam = np.linspace(0.1, 2.0, 27)
ang, irr = ref_table(params, diam)
out["table_diam"] = diam; out["table_angle"] = ang; out["table_irradiance"] = irr
np.savez_compressed(os.path.join(HERE, "bhmie_golden.npz"), **out)
print("wrote bhmie_golden.npz", irr.shape, ang.shape)
