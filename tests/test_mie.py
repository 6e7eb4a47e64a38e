"""the Mie table (bhmie) through the C ABI against the reference's own bhmie.py /
calculate_mie_scattering_intensity outputs (tests/golden/bhmie_golden.npz, made by make_mie_golden.py).
Host code: runs without a GPU."""
import os

import numpy as np
import pytest

from photon_b200 import mie
from photon_b200.lib import PhotonLibraryError

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bhmie_golden.npz"))
RTOL = 1e-11        # same recurrences in double; only complex division / abs differ in the last bits


@pytest.mark.parametrize("i", range(len(G["cases_x"])))
def test_bhmie_matches_reference(i):
    x, m, nang = float(G["cases_x"][i]), complex(G["cases_m"][i]), int(G["cases_nang"][i])
    s1, s2, qext, qsca, qback, gsca, theta = mie.bhmie(x, m, nang)
    scale1, scale2 = np.abs(G["s1_%d" % i]).max(), np.abs(G["s2_%d" % i]).max()
    assert s1.shape == (2 * nang - 1,)
    assert np.abs(s1 - G["s1_%d" % i]).max() <= RTOL * scale1
    assert np.abs(s2 - G["s2_%d" % i]).max() <= RTOL * scale2
    np.testing.assert_allclose([qext, qsca, qback, gsca], G["q_%d" % i], rtol=1e-10)
    np.testing.assert_array_equal(theta, G["theta_%d" % i])


def test_table_matches_reference():
    params = {"particle_field": {"medium_refractive_index": 1.0, "particle_refractive_index": 1.46,
                                 "mie_scattering_angle_number": 128, "beam_wavelength": 0.532}}
    ang, irr = mie.calculate_mie_scattering_intensity(params, G["table_diam"])
    np.testing.assert_allclose(ang, G["table_angle"], rtol=1e-15)
    assert irr.shape == G["table_irradiance"].shape == (255, 27)
    np.testing.assert_allclose(irr, G["table_irradiance"], rtol=1e-9)


def test_physical_properties():
    # optical theorem side: non-absorbing sphere scatters what it extinguishes; Rayleigh limit ~ x^4
    s1, s2, qext, qsca, qback, gsca, _ = mie.bhmie(5.0, 1.5, 64)
    assert abs(qext - qsca) < 1e-10 * qext
    q1 = mie.bhmie(0.01, 1.5, 8)[3]; q2 = mie.bhmie(0.02, 1.5, 8)[3]
    assert abs(q2 / q1 - 16.0) < 0.02
    assert -1.0 < gsca < 1.0


def test_argument_errors():
    with pytest.raises(PhotonLibraryError):
        mie.bhmie(1.0, 1.5, 2000)            # bhmie.py:21 'nang > mxnang'
    with pytest.raises(PhotonLibraryError):
        mie.bhmie(-1.0, 1.5, 16)
