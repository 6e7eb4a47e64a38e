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
                                  "piv_chief_ray_only"])
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
