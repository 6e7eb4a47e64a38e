"""NRRD0005 writer/reader for density volumes (raw or gzip, float32, x fastest).

Writes the same header fields pynrrd wrote for ``sample-data/bos/sample-density.nrrd``
(type, dimension, space, sizes, endian, encoding, spacings, space origin), which are the
only ones the tracer reads (``trace_rays_through_density_gradients.h:1687-1712``).
"""
import gzip

import numpy as np


def write_nrrd(path, rho_zyx, spacings, origin, encoding="raw"):
    """``rho_zyx``: float32 array indexed [z][y][x]; spacings/origin in micrometres (x, y, z)."""
    rho = np.ascontiguousarray(rho_zyx, np.float32)
    nz, ny, nx = rho.shape
    spacings = [float(v) for v in spacings]      # numpy scalars would print as "np.float64(...)"
    origin = [float(v) for v in origin]
    hdr = (
        "NRRD0005\n"
        "# written by photon_b200.nrrd\n"
        "type: float\n"
        "dimension: 3\n"
        "space: 3D-left-handed\n"
        f"sizes: {nx} {ny} {nz}\n"
        "endian: little\n"
        f"encoding: {encoding}\n"
        f"spacings: {spacings[0]!r} {spacings[1]!r} {spacings[2]!r}\n"
        f"space origin: ({origin[0]!r},{origin[1]!r},{origin[2]!r})\n"
        "\n"
    )
    with open(path, "wb") as f:
        f.write(hdr.encode("ascii"))
        if encoding == "raw":
            rho.tofile(f)
        elif encoding == "gzip":
            f.write(gzip.compress(rho.tobytes(), compresslevel=1))
        else:
            raise ValueError(encoding)


def read_nrrd(path):
    """Returns (rho[z][y][x] float32, spacings(3), origin(3))."""
    with open(path, "rb") as f:
        fields = {}
        magic = f.readline()
        if not magic.startswith(b"NRRD"):
            raise ValueError("not a NRRD file")
        while True:
            line = f.readline()
            if not line or line.strip() == b"":
                break
            line = line.decode("ascii", "replace").strip()
            if line.startswith("#") or ":" not in line:
                continue
            k, v = line.split(":", 1)
            fields[k.strip()] = v.strip()
        data = f.read()
    nx, ny, nz = (int(t) for t in fields["sizes"].split())
    if fields.get("encoding", "raw") in ("gzip", "gz"):
        data = gzip.decompress(data)
    dt = {"float": np.float32, "double": np.float64}[fields.get("type", "float")]
    rho = np.frombuffer(data, dt, count=nx * ny * nz).reshape(nz, ny, nx).astype(np.float32)
    spacings = tuple(float(t) for t in fields["spacings"].split())
    origin = tuple(float(t) for t in fields["space origin"].strip("()").split(","))
    return rho, spacings, origin


def register_field(lib, name, rho, spacings, origin):
    """Hand a density volume to the library in memory (photon_b200_register_field); afterwards
    ``density_grad_filename = "mem:<name>"`` selects it.  ``rho``: float32 [z][y][x] numpy array (kept
    alive by the returned handle) or a ``(device_pointer, (nz, ny, nx))`` pair for an array already on a GPU."""
    import ctypes as C
    if isinstance(rho, tuple):
        ptr, (nz, ny, nx) = rho
        keep = None
    else:
        keep = np.ascontiguousarray(rho, np.float32)
        nz, ny, nx = keep.shape
        ptr = keep.ctypes.data
    sp = (C.c_double * 3)(*[float(v) for v in spacings])
    og = (C.c_double * 3)(*[float(v) for v in origin])
    if lib.photon_b200_register_field(name.encode(), C.c_void_p(int(ptr)), int(nx), int(ny), int(nz), sp, og) != 0:
        raise RuntimeError(lib.photon_b200_last_error().decode())
    return keep
