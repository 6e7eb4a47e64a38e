"""CPU checker (test infrastructure only): restatement of the image post-processing that follows the
native call in python_codes/perform_ray_tracing_03.py:2189-2243.  Parity unpinned by reference tests
(the code is inline in a 400-line function); restated line by line.

dtype note: the image is the float32 array handed to ctypes; every in-place operation keeps float32, and
under the NumPy 1.x value-based casting the reference was written for, ``(2**bits - 1) * I / max`` stays
float32 as well (NumPy 2 would promote if pixel_bit_depth arrived as a float64 scalar).  float32 here."""
import numpy as np


def postprocess_u16(I_raw, pixel_gain, pixel_bit_depth, intensity_rescaling=True):
    I = np.array(I_raw, np.float32, copy=True)                       # :2216
    I[I < 0.0] = 0.0                                                  # :2220
    if intensity_rescaling:                                           # :2228
        I[~np.isfinite(I)] = 0.0                                      # :2234
        I *= np.float32(10 ** (float(pixel_gain) / 20.0))             # :2237
        mx = np.max(I)
        if mx > 0.0:                                                  # :2241
            I = (np.float32(2 ** int(pixel_bit_depth) - 1) * I) / mx  # :2242
        I = np.round(I)                                               # :2245
        I *= np.float32((2 ** 16 - 1.0) / (2 ** int(pixel_bit_depth) - 1.0))   # :2249
    return I.astype(np.int64).astype(np.uint16)                       # :2252 np.uint16(I)
