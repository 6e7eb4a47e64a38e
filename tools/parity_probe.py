#!/usr/bin/env python
"""GPU: parity metrics of one golden scene against the oracle, for the library PHOTON_B200_LIB selects.
usage: python tools/parity_probe.py [case ...]"""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.loader import load_oracle
from photon_b200.lib import load_library
from tests import util

d = tempfile.mkdtemp()
cases = util.parity_cases(d)
lib, orc = load_library(require_gpu=True), load_oracle()
for name in sys.argv[1:] or ["bos_rk4_thin_diffraction"]:
    call, pitch = cases[name]
    for k in "og":
        os.makedirs(os.path.join(d, name, k), exist_ok=True)
    ref = util.run_with_dumps(orc, call, os.path.join(d, name, "o"))
    got = util.run_with_dumps(lib, call, os.path.join(d, name, "g"))
    m = util.compare(ref, got, pitch)
    print(os.environ.get("PHOTON_B200_LIB", "default").split("/")[-1], name,
          " ".join("%s=%.3g" % (k, m[k]) for k in ("dpos_px", "dpos_px_p99", "ddir_rad", "image_rel_l2")), flush=True)
