"""Recordings of ``start_ray_tracing`` calls made by the reference's unchanged scripts.

``tools/stage_reference.py`` runs ``sample_run_script.sh piv|bos`` (pr03.py:1914-1938 is the call
site) against ``tools/record_stub.c``, which writes every argument to a directory.  This module
turns such a directory into a :class:`photon_b200.abi.RayTracingCall` and stores / loads calls as
single ``.npz`` files (the committed fixtures under ``tests/golden/recorded_*.npz``), so the exact
inputs of BASELINE configs C1 and C2 can be replayed where /root/reference does not exist.
"""
import ctypes as C
import os

import numpy as np

from .abi import RayTracingCall, camera_design_struct, element_data_struct

_INT = {"lightray_number_per_particle", "num_lightrays_save", "ray_tracing_algorithm",
        "num_intermediate_positions_save", "source_point_number"}
_BOOL = {"simulate_density_gradients", "save_lightrays", "add_pos_noise", "add_ngrad_noise",
         "save_intermediate_ray_data"}
_STR = {"scattering_type", "density_grad_filename", "lightray_position_save_path", "lightray_direction_save_path"}


def _scalars(path):
    out = {}
    with open(path) as f:
        for line in f:
            k, _, v = line.rstrip("\n").partition(" ")
            out[k] = v
    return out


def load_call(directory):
    """One ``call_NN`` directory written by tools/record_stub.c -> RayTracingCall."""
    sc = _scalars(os.path.join(directory, "scalars.txt"))
    assert (int(sc["sizeof_scattering"]), int(sc["sizeof_source"]), int(sc["sizeof_element"]),
            int(sc["sizeof_camera"])) == (72, 64, 120, 112), sc

    def arr(name, dtype):
        return np.fromfile(os.path.join(directory, name + ".bin"), dtype)

    c = RayTracingCall()
    for k in RayTracingCall.SCALARS:
        if k not in sc:
            continue
        v = sc[k]
        setattr(c, k, v if k in _STR else bool(int(v)) if k in _BOOL else int(v) if k in _INT else float(v))
    c.x, c.y, c.z = arr("x", np.float32), arr("y", np.float32), arr("z", np.float32)
    c.radiance, c.diameter_index = arr("radiance", np.float64), arr("diameter_index", np.int32)
    c.mie_inverse_rotation, c.mie_beam = arr("mie_inverse_rotation", np.float32), arr("mie_beam", np.float32)
    na, nd = int(sc["num_angles"]), int(sc["num_diameters"])
    c.mie_angles = arr("mie_angles", np.float32)
    irr = arr("mie_irradiance", np.float32)
    c.mie_irradiance = irr.reshape(na, nd) if irr.size else np.zeros((0, 0), np.float32)
    ne = int(sc["num_elements"])
    c.element_center = arr("element_center", np.float64).reshape(ne, 3)
    c.element_plane_parameters = arr("element_plane_parameters", np.float64).reshape(ne, 4)
    c.element_system_index = arr("element_system_index", np.int32)
    raw = arr("elements", np.uint8).tobytes()
    c.elements = [element_data_struct.from_buffer_copy(raw[i * 120:(i + 1) * 120]) for i in range(ne)]
    c.camera = camera_design_struct.from_buffer_copy(arr("camera", np.uint8).tobytes())
    return c


def load_recording(root):
    """All calls of one script run, in call order."""
    names = sorted(d for d in os.listdir(root) if d.startswith("call_"))
    return [load_call(os.path.join(root, d)) for d in names]


def save_npz(path, call, **extra):
    """Store one call (plus optional arrays such as reference outputs) as a compressed npz."""
    d = {k: np.asarray(getattr(call, k)) for k in RayTracingCall.SCALARS}
    for k in ("x", "y", "z", "radiance", "diameter_index", "mie_inverse_rotation", "mie_beam", "mie_angles",
              "mie_irradiance", "element_center", "element_plane_parameters", "element_system_index"):
        d[k] = np.asarray(getattr(call, k))
    d["elements_raw"] = np.frombuffer(b"".join(bytes(e) for e in call.elements), np.uint8)
    d["camera_raw"] = np.frombuffer(bytes(call.camera), np.uint8)
    d.update(extra)
    np.savez_compressed(path, **d)


def load_npz(path):
    """-> (RayTracingCall, dict of the extra arrays stored beside it)."""
    z = np.load(path)
    c = RayTracingCall()
    known = set()
    for k in RayTracingCall.SCALARS:
        v = z[k]
        known.add(k)
        setattr(c, k, str(v) if k in _STR else bool(v) if k in _BOOL else int(v) if k in _INT else float(v))
    for k in ("x", "y", "z", "radiance", "diameter_index", "mie_inverse_rotation", "mie_beam", "mie_angles",
              "mie_irradiance", "element_center", "element_plane_parameters", "element_system_index"):
        setattr(c, k, z[k])
        known.add(k)
    raw = z["elements_raw"].tobytes()
    c.elements = [element_data_struct.from_buffer_copy(raw[i:i + 120]) for i in range(0, len(raw), 120)]
    c.camera = camera_design_struct.from_buffer_copy(z["camera_raw"].tobytes())
    known |= {"elements_raw", "camera_raw"}
    assert C.sizeof(camera_design_struct) == 112
    return c, {k: z[k] for k in z.files if k not in known}
