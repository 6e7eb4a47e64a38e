"""CPU: the C-ABI library loads (no GPU needed for that) and exports every symbol that include/*.h
declares; the ctypes mirror has the reference's struct sizes.  No compute calls."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    names = []
    for h in ("photon_abi.h", "photon_b200.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        text = re.sub(r"//[^\n]*", "", text)
        for m in re.finditer(r"\b(start_ray_tracing|photon_[a-z0-9_]+)\s*\(", text):
            names.append(m.group(1))
    return sorted(set(names))


def test_headers_declare_the_reference_entry_point():
    assert "start_ray_tracing" in declared_functions()
    assert len(declared_functions()) >= 25


def test_library_exports_every_declared_symbol():
    from photon_b200.lib import library_path
    path = library_path()
    assert os.path.exists(path), "run __graft_entry__.build() first"
    lib = ctypes.CDLL(path)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing


def test_no_gpu_is_reported_not_emulated():
    """Without a CUDA device the library says so; there is no CPU path behind the ABI."""
    from photon_b200.lib import PhotonLibraryError, load_library
    lib = load_library(require_gpu=False)
    assert lib.photon_b200_abi_version() >= 1
    if lib.photon_b200_device_count() == 0:
        with pytest.raises(PhotonLibraryError):
            load_library(require_gpu=True)


def test_ctypes_mirror_matches_reference_layout():
    """Sizes/offsets of perform_ray_tracing_03.py:1651-1852, also static_asserted in C."""
    from photon_b200 import abi
    assert ctypes.sizeof(abi.scattering_data_struct) == 72
    assert ctypes.sizeof(abi.lightfield_source_struct) == 64
    assert ctypes.sizeof(abi.element_data_struct) == 120
    assert ctypes.sizeof(abi.camera_design_struct) == 112
    assert abi.lightfield_source_struct.z_offset.offset == 52
    assert abi.element_data_struct.element_type.offset == 80
    assert abi.camera_design_struct.rotation_matrix.offset == 40
    assert abi.camera_design_struct.inverse_rotation_matrix.offset == 76


def test_abi_header_compiles_as_c():
    """photon_abi.h is a C header (the reference's prt.h is included from C++ and bound from ctypes)."""
    import subprocess
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "t.c")
        open(src, "w").write('#include "photon_b200.h"\nint main(void) { return (int)sizeof(element_data_t) - 120; }\n')
        gcc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
        subprocess.run([gcc, "-std=c11", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", src, "-o",
                        os.path.join(d, "t.o")], check=True)
