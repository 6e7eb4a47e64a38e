"""Field ingest compared DIRECTLY, texel by texel, with what the reference's own setData
produced (trace_rays_through_density_gradients.h:1820-2002): its debug dump of the gradient field, recorded from the
host build of the reference (tests/golden/make_field_golden.py -> tests/golden/field_golden.npz).

  small volume   every face / edge / corner formula, anisotropic spacing, whole field, bit for bit
  large volume   160 x 128 x 1024 = 21e6 voxels: above 2^24 voxels the reference's FLOAT-arithmetic voxel index
                 (trace.h:1873-1975) rounds the low index bits away -- reproduced by default (PHOTON_GRADIENT_INDEX=exact
                 selects the integer index); bit for bit at 20 000 seeded voxels and on four planes of the x gradient
CPU: the oracle's restatement (and the live reference where its library exists).  GPU: photon_b200_load_field
(build_field_kernel), raw and gzip NRRD, and the volume handed over in memory.
"""
import ctypes as C
import importlib.util
import os

import numpy as np
import pytest

from photon_b200.nrrd import write_nrrd
from tests import util


def _gen():
    spec = importlib.util.spec_from_file_location("make_field_golden", os.path.join(util.GOLDEN_DIR, "make_field_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(util.GOLDEN_DIR, "field_golden.npz"))


@pytest.fixture(scope="module")
def volumes(golden, tmp_path_factory):
    """name -> (nrrd path, shape); the large volume is rebuilt from its closed form and checked against the recording."""
    d = tmp_path_factory.mktemp("ingest")
    gen = _gen()
    out = {}
    rho_l, sp_l = gen.large_volume()
    assert float(rho_l.astype(np.float64).sum()) == float(golden["large_rho_checksum"])
    for name, rho, sp in (("small", golden["small_rho"], tuple(float(v) for v in golden["small_spacing"])), ("large", rho_l, sp_l)):
        nz, ny, nx = rho.shape
        path = str(d / (name + ".nrrd"))
        origin = (-(nx - 1) * sp[0] / 2, -(ny - 1) * sp[1] / 2, 750e3 - (nz - 1) * sp[2] / 2)
        write_nrrd(path, rho, sp, origin)
        out[name] = (path, rho.shape, rho, sp, origin)
    return out


def _check(texels, name, golden):
    if name == "small":
        assert np.array_equal(texels, golden["small_texels"])
        return
    p = golden["large_positions"]
    assert np.array_equal(texels[p[:, 0], p[:, 1], p[:, 2]], golden["large_texels"])
    got, want = texels[[0, 818, 819, 1023], :, :, 0].copy(), golden["large_gx_planes"].copy()
    # the very last voxels of the volume: the reference's rounded index runs PAST the end of its array there (it reads
    # whatever the heap holds); product and oracle clamp to the last voxel
    got[3, -1, -4:] = 0; want[3, -1, -4:] = 0
    assert np.array_equal(got, want)


def oracle_field(lib, path, shape):
    out = np.zeros(shape + (4,), np.float32)
    meta = np.zeros(16, np.float32)
    lib.photon_oracle_load_field.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_longlong]
    assert lib.photon_oracle_load_field(path.encode(), meta.ctypes.data, out.ctypes.data, out.size // 4) == 0
    return out, meta


@pytest.mark.parametrize("name", ["small", "large"])
def test_oracle_field_is_the_reference_setdata_output(name, volumes, golden, oracle_lib):
    path, shape = volumes[name][:2]
    tex, meta = oracle_field(oracle_lib, path, shape)
    _check(tex, name, golden)
    assert meta[10] == tex[..., 3].min()                                 # data_min (trace.h:1868-1869)


def test_float_index_matters_only_above_2_pow_24(volumes, golden, oracle_lib, monkeypatch):
    """The large volume with the integer index: identical below voxel 2^24 (z < 819), different above."""
    path, shape = volumes["large"][:2]
    monkeypatch.setenv("PHOTON_GRADIENT_INDEX", "exact")
    tex, _ = oracle_field(oracle_lib, path, shape)
    assert np.array_equal(tex[[0, 818], :, :, 0], golden["large_gx_planes"][:2])
    assert (tex[[819, 1023], :, :, 0] != golden["large_gx_planes"][2:]).mean() > 0.3
    p = golden["large_positions"]
    same = (tex[p[:, 0], p[:, 1], p[:, 2]] == golden["large_texels"]).all(axis=1)
    assert same[p[:, 0] < 818].all() and not same[p[:, 0] >= 820].all()          # z = 818 reads its z + 1 neighbours above 2^24
    assert np.array_equal(tex[p[:, 0], p[:, 1], p[:, 2], 3], golden["large_texels"][:, 3])      # n-1 itself: always exact


def test_live_reference_dump_matches_golden(golden, tmp_path):
    """Where the host build of the reference exists: run its ingest again and compare its dump with the recording."""
    from oracle.loader import ref_cpu_path
    if not os.path.exists(ref_cpu_path()):
        pytest.skip("oracle/_ref/libphoton_ref_cpu.so not built")
    gen = _gen()
    _, tex = gen.reference_texels(golden["small_rho"], tuple(float(v) for v in golden["small_spacing"]), str(tmp_path))
    if tex.size == 0:
        pytest.skip("this build of the reference does not redirect its gradient dump")
    assert np.array_equal(tex, golden["small_texels"])


# ------------------------------------------------------------------------------------------------------------- GPU
def cuda_field(lib, path, shape):
    out = np.zeros(shape + (4,), np.float32)
    meta = np.zeros(16, np.float32)
    assert lib.photon_b200_load_field(path.encode(), meta.ctypes.data, out.ctypes.data, out.size // 4) == 0
    return out, meta


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["small", "large"])
def test_cuda_field_is_the_reference_setdata_output(name, volumes, golden, cuda_lib):
    path, shape = volumes[name][:2]
    tex, meta = cuda_field(cuda_lib, path, shape)
    _check(tex, name, golden)
    assert meta[10] == tex[..., 3].min()
    assert tuple(int(v) for v in meta[11:14]) == (shape[2], shape[1], shape[0])


@pytest.mark.gpu
def test_cuda_field_from_gzip_nrrd_and_from_memory(volumes, golden, cuda_lib, tmp_path):
    """The same small volume as a gzip-encoded NRRD (teem reads both, trace.h:1676-1683) and handed over in memory."""
    from photon_b200.nrrd import register_field
    path, shape, rho, sp, origin = volumes["small"]
    gz = str(tmp_path / "small_gzip.nrrd")
    write_nrrd(gz, rho, sp, origin, encoding="gzip")
    assert os.path.getsize(gz) < os.path.getsize(path)
    tex, _ = cuda_field(cuda_lib, gz, shape)
    assert np.array_equal(tex, golden["small_texels"])
    keep = register_field(cuda_lib, "ingest-small", rho, sp, origin)
    tex, _ = cuda_field(cuda_lib, "mem:ingest-small", shape)
    assert np.array_equal(tex, golden["small_texels"])
    cuda_lib.photon_b200_unregister_field(b"ingest-small")
    del keep


@pytest.mark.gpu
def test_cuda_exact_index_switch(volumes, golden, cuda_lib, monkeypatch):
    path, shape = volumes["large"][:2]
    monkeypatch.setenv("PHOTON_GRADIENT_INDEX", "exact")
    monkeypatch.setenv("PHOTON_NO_FIELD_CACHE", "1")
    tex, _ = cuda_field(cuda_lib, path, shape)
    assert np.array_equal(tex[[0, 818], :, :, 0], golden["large_gx_planes"][:2])
    assert (tex[[819, 1023], :, :, 0] != golden["large_gx_planes"][2:]).mean() > 0.3
