"""Cubic mode (PHOTON_INTERPOLATION=cubic): tricubic B-spline interpolation of n-1 with analytic
gradient on a scalar coefficient volume.  The reference's own cubic branch is unreachable through the
ABI (prt.cu:3330 hard-codes interpolation_scheme = 1), so there is no traced reference output.  PINNED at
the level the reference offers: CubicInterpolationCUDA ships a plain-C++ CPU implementation of its
algorithm (examples/referenceCubicTexture3D/cubicPrefilter3D.hpp:51-150, cubicFilter3D_kernel.hpp:46-93;
weights and derivative weights code/internal/bspline_kernel.cu:50-104).  oracle/build_ref_ci.sh compiles
it from where it lies; tests/golden/ci_golden.npz holds its answers (tests/golden/make_ci_golden.py).
  KAT  prefilter bit for bit, weights and derivative weights bit for bit, 64-tap value to float rounding
       (CI sums the taps in z, y, x order, the product contracts x -> y -> z), gradient against the sum
       built from CI's derivative weights -- oracle and host-compiled stage code on the CPU, the CUDA
       prefilter and sampler on the GPU;
  CPU  the oracle's restatement of CubicInterpolationCUDA's published algorithm has the properties the
       algorithm promises (knot reproduction, exactness on polynomials up to degree 3 away from the
       faces, gradient = derivative of the value), agrees with the linear mode to O(h^2) on a smooth
       field, and the product's stage code compiled for the host is bit-exact against it;
  GPU  the CUDA prefilter is bit-exact against the oracle's, single samples agree to rounding, and
       whole traces meet the north-star tolerances against the oracle in the same mode.
"""
import ctypes as C
import os

import numpy as np
import pytest

from photon_b200 import nrrd, scenes
from tests import util


@pytest.fixture()
def cubic_env():
    os.environ["PHOTON_INTERPOLATION"] = "cubic"
    yield
    os.environ.pop("PHOTON_INTERPOLATION", None)


def oracle_samples(lib, path, cubic, pts):
    pts = np.ascontiguousarray(pts, np.float32)
    out = np.zeros((len(pts), 4), np.float32)
    lib.photon_oracle_sample_field.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    assert lib.photon_oracle_sample_field(path.encode(), int(cubic), len(pts), pts.ctypes.data, out.ctypes.data) == 0
    return out


def oracle_coefficients(lib, path, shape):
    out = np.zeros(shape, np.float32)
    lib.photon_oracle_load_coefficients.argtypes = [C.c_char_p, C.c_void_p, C.c_longlong]
    assert lib.photon_oracle_load_coefficients(path.encode(), out.ctypes.data, out.size) == 0
    return out


K = np.float32(0.225e-3)


def write_field(path, rho, spacing=100.0):
    n = rho.shape[0]
    nrrd.write_nrrd(path, rho.astype(np.float32), (spacing,) * 3, (0.0, 0.0, 750e3))
    return n


def test_knot_reproduction_and_prefilter_inverse(oracle_lib, tmp_path):
    """The spline through the prefiltered coefficients interpolates the samples (CI's defining property)."""
    rng = np.random.default_rng(7)
    n = 24
    rho = 1.2 + 0.3 * rng.random((n, n, n))
    path = str(tmp_path / "rand.nrrd")
    write_field(path, rho)
    kk, jj, ii = np.meshgrid(np.arange(2, n - 2), np.arange(2, n - 2), np.arange(2, n - 2), indexing="ij")
    pts = np.stack([ii.ravel() + 0.5, jj.ravel() + 0.5, kk.ravel() + 0.5], axis=1)      # texel centres
    got = oracle_samples(oracle_lib, path, 1, pts)[:, 3]
    want = (K * rho.astype(np.float32))[kk.ravel(), jj.ravel(), ii.ravel()]
    assert np.max(np.abs(got - want)) <= 2e-6 * np.max(np.abs(want))


def test_exact_on_cubic_polynomial_with_analytic_gradient(oracle_lib, tmp_path):
    n, h = 32, 100.0
    c = np.arange(n, dtype=np.float64)
    z, y, x = np.meshgrid(c, c, c, indexing="ij")
    f = lambda x, y, z: 1.0 + 0.02 * x + 0.01 * y * z / n + 1e-3 * x * x - 2e-5 * x * x * x + 3e-4 * y * y + 1e-5 * z ** 3
    path = str(tmp_path / "poly.nrrd")
    write_field(path, f(x, y, z), h)
    rng = np.random.default_rng(3)
    p = rng.uniform(8.0, n - 9.0, (2000, 3))                 # grid units, away from the clamped faces
    got = oracle_samples(oracle_lib, path, 1, p + 0.5)
    px, py, pz = p[:, 0], p[:, 1], p[:, 2]
    val = K * f(px, py, pz)
    gx = K * (0.02 + 2e-3 * px - 6e-5 * px * px) / h
    gy = K * (0.01 * pz / n + 6e-4 * py) / h
    gz = K * (0.01 * py / n + 3e-5 * pz * pz) / h
    scale = np.max(np.abs(val))
    assert np.max(np.abs(got[:, 3] - val)) <= 5e-6 * scale
    for g, w in ((got[:, 0], gx), (got[:, 1], gy), (got[:, 2], gz)):
        assert np.max(np.abs(g - w)) <= 2e-4 * np.max(np.abs(w))


def test_cubic_agrees_with_linear_on_a_smooth_field(oracle_lib, tmp_path):
    path = str(tmp_path / "blob.nrrd")
    n = 64
    spacing = 511 * 176.0 / (n - 1)
    nrrd.write_nrrd(path, scenes.gaussian_blob(n, spacing, 15e3), (spacing,) * 3, (0.0, 0.0, 750e3))
    rng = np.random.default_rng(5)
    p = rng.uniform(4.0, n - 4.0, (4000, 3)).astype(np.float32)
    lin = oracle_samples(oracle_lib, path, 0, p)
    cub = oracle_samples(oracle_lib, path, 1, p)
    assert np.max(np.abs(lin[:, 3] - cub[:, 3])) <= 5e-3 * np.max(np.abs(lin[:, 3] - lin[:, 3].min()))     # linear is O(h^2)
    g = np.max(np.abs(lin[:, :3]))
    assert np.max(np.abs(lin[:, :3] - cub[:, :3])) <= 0.03 * g                           # O(h^2) on 64^3


@pytest.mark.parametrize("alg", [2, 1])
def test_stage_code_bit_exact_vs_oracle_cubic(alg, cubic_env, hostcheck_lib, oracle_lib, tmp_path):
    c = scenes.c3_bos(str(tmp_path / "b.nrrd"), n=48, dots_per_side=8, points_per_dot=8, rays_per_point=16, algorithm=alg)
    (tmp_path / "o").mkdir(); (tmp_path / "h").mkdir()
    ref = util.run_with_dumps(oracle_lib, c, str(tmp_path / "o"))
    got = util.run_with_dumps(hostcheck_lib, c, str(tmp_path / "h"))
    assert np.array_equal(ref[1], got[1], equal_nan=True)
    assert np.array_equal(ref[2], got[2], equal_nan=True)
    os.environ.pop("PHOTON_INTERPOLATION")
    (tmp_path / "l").mkdir()
    lin = util.run_with_dumps(oracle_lib, c, str(tmp_path / "l"))
    m = util.compare(lin, ref, 10.0)
    assert 0 < m["dpos_px"] < 0.05, m                                                   # a different interpolant, same physics


# ------------------------------------------------------- known answers of CI's own CPU reference
def ci_golden():
    return np.load(os.path.join(util.GOLDEN_DIR, "ci_golden.npz"))


def ci_gradient(g):
    """Value and gradient (per texel) at g['pts'] as the 64-term sums of CI's weights and first-derivative weights
    (bspline_kernel.cu:83-104) over CI's coefficients, in float64: the known answer for the analytic gradient."""
    coef = g["coef"].astype(np.float64)
    pts = g["pts"].astype(np.float32)
    b = pts - np.float32(0.5)
    f = np.floor(b)
    fr = (b - f).astype(np.float32)
    idx = f.astype(np.int64)
    frs, W, DW = g["fractions"], g["weights"].astype(np.float64), g["dweights"].astype(np.float64)

    def wts(t):                                     # CI's weights are cubic polynomials: evaluate them as CI does, in float32
        one = np.float32(1.0) - t; sq = t * t; osq = one * one
        w = np.stack([np.float32(1 / 6.0) * osq * one, np.float32(2 / 3.0) - np.float32(0.5) * sq * (np.float32(2.0) - t),
                      np.float32(2 / 3.0) - np.float32(0.5) * osq * (np.float32(2.0) - one), np.float32(1 / 6.0) * sq * t], -1)
        d = np.stack([np.float32(-0.5) * sq + t - np.float32(0.5), np.float32(1.5) * sq - np.float32(2.0) * t,
                      np.float32(-1.5) * sq + t + np.float32(0.5), np.float32(0.5) * sq], -1)
        return w.astype(np.float64), d.astype(np.float64)
    # the closed forms above ARE CI's (checked against its exported functions on the stored fractions)
    w0, d0 = wts(frs)
    assert np.array_equal(w0.astype(np.float32), W.astype(np.float32)) and np.array_equal(d0.astype(np.float32), DW.astype(np.float32))
    wx, dx = wts(fr[:, 0]); wy, dy = wts(fr[:, 1]); wz, dz = wts(fr[:, 2])
    val = np.zeros(len(pts)); gx = np.zeros(len(pts)); gy = np.zeros(len(pts)); gz = np.zeros(len(pts))
    for c in range(4):
        for bb in range(4):
            for a in range(4):
                t = coef[idx[:, 2] - 1 + c, idx[:, 1] - 1 + bb, idx[:, 0] - 1 + a]
                val += wx[:, a] * wy[:, bb] * wz[:, c] * t
                gx += dx[:, a] * wy[:, bb] * wz[:, c] * t
                gy += wx[:, a] * dy[:, bb] * wz[:, c] * t
                gz += wx[:, a] * wy[:, bb] * dz[:, c] * t
    return val, gx, gy, gz


def _write_ci_volume(tmp_path, g, spacing=100.0):
    path = str(tmp_path / "ci_volume.nrrd")
    nrrd.write_nrrd(path, g["rho"], (spacing,) * 3, (0.0, 0.0, 750e3))
    return path


def _check_against_ci(coefs, samples, g, spacing=100.0):
    assert np.array_equal(coefs, g["coef"])                                      # the prefilter, bit for bit
    scale = float(np.abs(g["val"]).max())
    assert np.max(np.abs(samples[:, 3] - g["val"])) <= 2e-6 * scale              # CI's own 64-tap sum (float32, z-y-x order)
    val, gx, gy, gz = ci_gradient(g)
    assert np.max(np.abs(samples[:, 3] - val)) <= 1e-6 * scale
    for k, want in enumerate((gx, gy, gz)):
        want = want / spacing                                                    # per texel -> per micrometre
        assert np.max(np.abs(samples[:, k] - want)) <= 2e-5 * np.max(np.abs(want)), k


def test_ci_reference_live_matches_golden():
    """Where /root/reference is mounted: rebuild CI's CPU reference from its own sources and regenerate the answers."""
    if not os.path.isdir("/root/reference/cuda_codes/CubicInterpolationCUDA"):
        pytest.skip("reference not mounted (GPU box)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_ci_golden", os.path.join(util.GOLDEN_DIR, "make_ci_golden.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    lib = mod.load_ci()
    rho, nm1, pts, fractions = mod.inputs()
    coef, val, wts, dws, ts, b, db = mod.ci_answers(lib, nm1, pts, fractions)
    g = ci_golden()
    assert np.array_equal(rho, g["rho"]) and np.array_equal(coef, g["coef"]) and np.array_equal(val, g["val"])
    assert np.array_equal(wts, g["weights"]) and np.array_equal(dws, g["dweights"])
    # CI's piecewise B-spline and its derivative agree with its own branch-free weights: w_a(f) = B(a - 1 - f) ...
    for a in range(4):
        for f, w, d in zip(fractions[::8], wts[::8], dws[::8]):
            assert abs(lib.photon_ci_bspline(float(f) - (a - 1)) - w[a]) <= 2e-7
            if 0.0 < f < 1.0:
                assert abs(-lib.photon_ci_bspline_1st_derivative(float(a - 1) - float(f)) - d[a]) <= 4e-7


def test_oracle_cubic_against_ci_known_answers(oracle_lib, tmp_path):
    g = ci_golden()
    path = _write_ci_volume(tmp_path, g)
    _check_against_ci(oracle_coefficients(oracle_lib, path, g["coef"].shape), oracle_samples(oracle_lib, path, 1, g["pts"]), g)


def test_weights_bit_exact_against_ci(oracle_lib):
    """bspline_w of the product / oracle = CI's bspline_weights and bspline_weights_1st_derivative, bit for bit."""
    g = ci_golden()
    oracle_lib.photon_oracle_bspline_weights.argtypes = [C.c_float, C.c_void_p, C.c_void_p]
    w = np.zeros(4, np.float32); d = np.zeros(4, np.float32)
    for t, cw, cd in zip(g["fractions"], g["weights"], g["dweights"]):
        oracle_lib.photon_oracle_bspline_weights(float(t), w.ctypes.data, d.ctypes.data)
        assert np.array_equal(w, cw) and np.array_equal(d, cd), t


# ------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_cuda_cubic_against_ci_known_answers(cuda_lib, tmp_path):
    g = ci_golden()
    path = _write_ci_volume(tmp_path, g)
    got = np.zeros(g["coef"].shape, np.float32)
    cuda_lib.photon_b200_load_coefficients.argtypes = [C.c_char_p, C.c_void_p, C.c_longlong]
    assert cuda_lib.photon_b200_load_coefficients(path.encode(), got.ctypes.data, got.size) == 0
    _check_against_ci(got, cuda_samples(cuda_lib, path, 1, g["pts"]), g)


def cuda_samples(lib, path, cubic, pts):
    pts = np.ascontiguousarray(pts, np.float32)
    out = np.zeros((len(pts), 4), np.float32)
    lib.photon_b200_sample_field.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    assert lib.photon_b200_sample_field(path.encode(), int(cubic), len(pts), pts.ctypes.data, out.ctypes.data) == 0
    return out


@pytest.mark.gpu
def test_cuda_prefilter_bit_exact_and_samples(cuda_lib, oracle_lib, tmp_path):
    path = str(tmp_path / "blob.nrrd")
    n = 64
    spacing = 511 * 176.0 / (n - 1)
    rng = np.random.default_rng(11)
    rho = scenes.gaussian_blob(n, spacing, 15e3) + 0.01 * rng.random((n, n, n)).astype(np.float32)
    nrrd.write_nrrd(path, rho, (spacing,) * 3, (0.0, 0.0, 750e3))
    want = oracle_coefficients(oracle_lib, path, (n, n, n))
    got = np.zeros((n, n, n), np.float32)
    cuda_lib.photon_b200_load_coefficients.argtypes = [C.c_char_p, C.c_void_p, C.c_longlong]
    assert cuda_lib.photon_b200_load_coefficients(path.encode(), got.ctypes.data, got.size) == 0
    assert np.array_equal(want, got)                                                     # same recursion, no contraction
    p = rng.uniform(-0.5, n + 0.5, (20000, 3)).astype(np.float32)                        # includes the clamped faces
    for cubic in (0, 1):
        a = oracle_samples(oracle_lib, path, cubic, p)
        b = cuda_samples(cuda_lib, path, cubic, p)
        assert np.max(np.abs(a[:, 3] - b[:, 3])) <= 1e-6 * np.max(np.abs(a[:, 3]))
        assert np.max(np.abs(a[:, :3] - b[:, :3])) <= 2e-5 * np.max(np.abs(a[:, :3]))    # FMA vs separate roundings


@pytest.mark.gpu
@pytest.mark.parametrize("alg", [2, 1])
def test_cuda_matches_oracle_cubic(alg, cubic_env, cuda_lib, oracle_lib, tmp_path):
    c = scenes.c3_bos(str(tmp_path / "b.nrrd"), n=64, dots_per_side=10, points_per_dot=10, rays_per_point=20, algorithm=alg)
    (tmp_path / "o").mkdir(); (tmp_path / "g").mkdir()
    ref = util.run_with_dumps(oracle_lib, c, str(tmp_path / "o"))
    got = util.run_with_dumps(cuda_lib, c, str(tmp_path / "g"))
    m = util.compare(ref, got, 10.0)
    print("cubic alg", alg, m)
    util.assert_parity("bos_cubic", m)


@pytest.mark.gpu
def test_tma_brick_loop_is_bit_identical_to_the_table_loop(cubic_env, cuda_lib, oracle_lib, tmp_path):
    """PHOTON_CUBIC_BRICK=1: the cubic hot loop on per-warp bricks of the quad-row table staged in shared memory by TMA
    (CubicGatherBrick, ray_stages.cuh).  Same taps and the same contraction as the table loop, so per-ray dumps and the
    non-dump image must be IDENTICAL, the variant name proves which instantiation ran, no refill may time out, and the
    result sits within the north-star bounds of the oracle like the table loop's."""
    c = scenes.c3_bos(str(tmp_path / "b.nrrd"), n=64, dots_per_side=10, points_per_dot=10, rays_per_point=70, algorithm=2)
    for d in ("o", "t", "b"):
        (tmp_path / d).mkdir()
    ref = util.run_with_dumps(oracle_lib, c, str(tmp_path / "o"))
    tab = util.run_with_dumps(cuda_lib, c, str(tmp_path / "t"))
    assert cuda_lib.photon_b200_last_variant().decode() == "rk4/cubic/dump"
    img_tab = np.array(c.run(cuda_lib), np.float64)
    os.environ["PHOTON_CUBIC_BRICK"] = "1"
    try:
        t0 = cuda_lib.photon_b200_brick_timeouts()
        got = util.run_with_dumps(cuda_lib, c, str(tmp_path / "b"))
        assert cuda_lib.photon_b200_last_variant().decode() == "rk4/cubic-brick/dump"
        img_got = np.array(c.run(cuda_lib), np.float64)
        assert cuda_lib.photon_b200_last_variant().decode() == "rk4/cubic-brick/image"
        assert cuda_lib.photon_b200_brick_timeouts() == t0 >= 0
    finally:
        os.environ.pop("PHOTON_CUBIC_BRICK")
    for a, b in zip(tab[1:], got[1:]):                       # per-ray positions and directions: the same bits (NaN = NaN)
        assert np.array_equal(np.asarray(a), np.asarray(b), equal_nan=True)
    rel = np.linalg.norm(img_got - img_tab) / np.linalg.norm(img_tab)
    assert rel <= 1e-6, rel                                                # float atomics: summation order only
    util.assert_parity("bos_cubic", util.compare(ref, got, 10.0))
