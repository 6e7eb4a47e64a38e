"""Mie scattering table: host mirror of python_codes/bhmie.py and of
run_simulation_02.py:641-696 (calculate_mie_scattering_intensity) over the C ABI
(photon_b200_bhmie / photon_b200_mie_irradiance_table, include/photon_b200.h).  Same names,
argument meaning and return values as the reference."""
import ctypes as C

import numpy as np

from .lib import PhotonLibraryError, load_library


def _lib():
    lib = load_library(require_gpu=False)
    lib.photon_b200_bhmie.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.photon_b200_bhmie.restype = C.c_int
    lib.photon_b200_mie_irradiance_table.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int,
                                                     C.c_void_p, C.c_void_p]
    lib.photon_b200_mie_irradiance_table.restype = C.c_int
    return lib


def bhmie(x, refrel, nang):
    """bhmie.py:3-173 -> (s1, s2, qext, qsca, qback, gsca, theta); s1, s2 complex128[2*nang-1]."""
    lib = _lib()
    nang = int(nang)
    refrel = complex(refrel)
    s1 = np.zeros(2 * nang - 1, np.complex128)
    s2 = np.zeros(2 * nang - 1, np.complex128)
    q = np.zeros(4, np.float64)
    if lib.photon_b200_bhmie(float(x), refrel.real, refrel.imag, nang, s1.ctypes.data, s2.ctypes.data, q.ctypes.data) != 0:
        raise PhotonLibraryError(lib.photon_b200_last_error().decode())
    theta = np.linspace(0, 1, s1.size)                     # bhmie.py:171 (sic: not an angle)
    return s1, s2, q[0], q[1], q[2], q[3], theta


def calculate_mie_scattering_intensity(simulation_parameters, particle_diameter_vector):
    """rs02.py:641-696 -> [scattering_angle (rad, 2*nang-1), scattering_irradiance (2*nang-1, n_diameters)]."""
    pf = simulation_parameters["particle_field"]
    sizes = np.ascontiguousarray(np.asarray(particle_diameter_vector, np.float64).reshape(-1))
    nang = int(pf["mie_scattering_angle_number"])
    ang = np.zeros(2 * nang - 1, np.float64)
    irr = np.zeros((2 * nang - 1, sizes.size), np.float64)
    lib = _lib()
    if lib.photon_b200_mie_irradiance_table(sizes.ctypes.data, sizes.size, float(pf["medium_refractive_index"]),
                                            float(pf["particle_refractive_index"]), float(pf["beam_wavelength"]), nang,
                                            ang.ctypes.data, irr.ctypes.data) != 0:
        raise PhotonLibraryError(lib.photon_b200_last_error().decode())
    return [ang, irr]
