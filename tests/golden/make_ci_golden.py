#!/usr/bin/env python
"""Generates tests/golden/ci_golden.npz: known answers of CubicInterpolationCUDA's OWN CPU reference
(oracle/_ref/libphoton_ref_ci.so = examples/referenceCubicTexture3D/cubicPrefilter3D.hpp, cubicFilter3D_kernel.hpp,
code/internal/bspline_kernel.cu, cubicPrefilter_kernel.cu compiled for the host by oracle/build_ref_ci.sh).

  python tests/golden/make_ci_golden.py          (needs /root/reference: build container only)

Stored: a seeded random density volume rho[20][22][24], the n-1 volume handed to CI (0.225e-3 * rho in float32),
CI's B-spline coefficients of it, 600 sample coordinates (all 64 taps inside the volume) with CI's interpolated
values, and CI's weights / first-derivative weights / piecewise B-spline at 65 fractions.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
HERE = os.path.dirname(os.path.abspath(__file__))


def load_ci():
    so = os.path.join(ROOT, "oracle", "_ref", "libphoton_ref_ci.so")
    if os.path.isdir("/root/reference"):
        subprocess.run(["bash", os.path.join(ROOT, "oracle", "build_ref_ci.sh")], check=True, capture_output=True)
    lib = C.CDLL(so)
    lib.photon_ci_bspline.restype = C.c_float
    lib.photon_ci_bspline.argtypes = [C.c_float]
    lib.photon_ci_bspline_1st_derivative.restype = C.c_float
    lib.photon_ci_bspline_1st_derivative.argtypes = [C.c_float]
    lib.photon_ci_weights.argtypes = [C.c_float, C.c_void_p, C.c_void_p]
    lib.photon_ci_prefilter3d.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint]
    lib.photon_ci_interpolate.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_void_p, C.c_void_p]
    return lib


def ci_answers(lib, nm1, pts, fractions):
    d, h, w = nm1.shape
    coef = np.ascontiguousarray(nm1, np.float32).copy()
    lib.photon_ci_prefilter3d(coef.ctypes.data, w, h, d)
    pts = np.ascontiguousarray(pts, np.float32)
    val = np.zeros(len(pts), np.float32)
    lib.photon_ci_interpolate(coef.ctypes.data, w, h, d, len(pts), pts.ctypes.data, val.ctypes.data)
    wts = np.zeros((len(fractions), 4), np.float32)
    dws = np.zeros((len(fractions), 4), np.float32)
    for k, t in enumerate(fractions):
        lib.photon_ci_weights(float(t), wts[k].ctypes.data, dws[k].ctypes.data)
    ts = np.linspace(-2.5, 2.5, 101).astype(np.float32)
    b = np.array([lib.photon_ci_bspline(float(t)) for t in ts], np.float32)
    db = np.array([lib.photon_ci_bspline_1st_derivative(float(t)) for t in ts], np.float32)
    return coef, val, wts, dws, ts, b, db


def inputs():
    rng = np.random.default_rng(20121)
    rho = (1.2 + 0.3 * rng.random((20, 22, 24))).astype(np.float32)
    nm1 = (np.float32(0.225e-3) * rho).astype(np.float32)
    # coordinates in CI's / tex3D's convention (texel centres at +0.5); base texel 1 .. N-3 keeps all 64 taps inside
    lo, hi = 1.5, np.array([24, 22, 20], np.float64) - 2.5
    pts = (lo + rng.random((600, 3)) * (hi - lo)).astype(np.float32)
    fractions = np.linspace(0.0, 1.0, 65).astype(np.float32)
    return rho, nm1, pts, fractions


if __name__ == "__main__":
    lib = load_ci()
    rho, nm1, pts, fractions = inputs()
    coef, val, wts, dws, ts, b, db = ci_answers(lib, nm1, pts, fractions)
    np.savez_compressed(os.path.join(HERE, "ci_golden.npz"), rho=rho, nm1=nm1, coef=coef, pts=pts, val=val,
                        fractions=fractions, weights=wts, dweights=dws, ts=ts, bspline=b, dbspline=db)
    print("ci_golden.npz: coef", coef.shape, "values", val[:3], "weights", wts[32], dws[32])
