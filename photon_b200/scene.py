"""Resident-scene wrapper over the extended C ABI (include/photon_b200.h).

``trace_image(call)`` is the drop-in path (one ``start_ray_tracing`` call with host
buffers, exactly what ``perform_ray_tracing_03.py:1929`` does).  ``ResidentScene`` keeps the
scene in HBM so that the tracing kernel can be re-run and timed on its own.
"""
import ctypes as C

import numpy as np

from .abi import RayTracingCall
from .lib import PhotonLibraryError, load_library


def trace_image(call: RayTracingCall, image=None):
    """Run one ``start_ray_tracing`` call of the CUDA library; returns the (ny, nx) image."""
    lib = load_library(require_gpu=True)
    return call.run(lib, image)


class ResidentScene:
    def __init__(self, call: RayTracingCall, shard=(0, 1), device=-1):
        self.lib = load_library(require_gpu=True)
        self.call = call
        ny, nx = call.image_shape
        dummy = np.zeros(1, np.float32)
        args, keep = call.marshal(dummy)
        # photon_scene_create takes the start_ray_tracing arguments minus image / dump options
        (lens_pitch, image_distance, sc, sct, src, rpp, wl, fnum, ne, ec, elems, ep, esi, cam, _img,
         sim, fname, _save, _pp, _dp, _ns, alg, apn, pstd, agn, gstd, cone, _si, _ni) = args
        fn = self.lib.photon_scene_create
        fn.argtypes = [C.c_float, C.c_float, C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_float, C.c_float,
                       C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                       C.c_bool, C.c_char_p, C.c_int, C.c_bool, C.c_float, C.c_bool, C.c_float, C.c_float,
                       C.c_int, C.c_int, C.c_int]
        self.handle = fn(lens_pitch, image_distance, C.cast(sc, C.c_void_p), sct, C.cast(src, C.c_void_p), rpp, wl,
                         fnum, ne, ec, C.cast(elems, C.c_void_p), ep, esi, C.cast(cam, C.c_void_p), sim, fname, alg,
                         apn, pstd, agn, gstd, cone, int(shard[0]), int(shard[1]), int(device))
        del keep
        if not self.handle:
            raise PhotonLibraryError(self.lib.photon_b200_last_error().decode() or "photon_scene_create failed")
        self.shape = (ny, nx)

    # ------------------------------------------------------------------
    def _ck(self, rc, what):
        if rc != 0:
            raise PhotonLibraryError(f"{what}: {self.lib.photon_b200_last_error().decode()}")

    def zero_image(self):
        self._ck(self.lib.photon_scene_zero_image(self.handle), "zero_image")

    def upload_image(self, img):
        img = np.ascontiguousarray(img, np.float32)
        self._ck(self.lib.photon_scene_upload_image(self.handle, img.ctypes.data), "upload_image")

    def download_image(self):
        out = np.empty(self.shape, np.float32)
        self._ck(self.lib.photon_scene_download_image(self.handle, out.ctypes.data), "download_image")
        return out

    def export_u16(self, pixel_gain=None, pixel_bit_depth=None, intensity_rescaling=True, image_noise=0.0, seed=0):
        """The 16-bit image the reference writes to TIFF (pr03.py:2189-2243), computed on the device.
        Defaults for gain / bit depth come from the call's camera_design."""
        cam = self.call.camera
        g = cam.pixel_gain if pixel_gain is None else pixel_gain
        b = cam.pixel_bit_depth if pixel_bit_depth is None else pixel_bit_depth
        out = np.empty(self.shape, np.uint16)
        self._ck(self.lib.photon_scene_export_u16(self.handle, float(g), int(b), int(bool(intensity_rescaling)),
                                                  float(image_noise), int(seed) & 0xffffffff, out.ctypes.data), "export_u16")
        return out

    @property
    def deterministic(self):
        return bool(self.lib.photon_scene_deterministic(self.handle))

    def download_image_fixed(self):
        """Deterministic mode: (int64 image in units of 1/scale, scale)."""
        out = np.empty(self.shape, np.int64)
        scale = C.c_double(0.0)
        self._ck(self.lib.photon_scene_download_image_fixed(self.handle, out.ctypes.data, C.byref(scale)), "download_image_fixed")
        return out, scale.value

    def trace(self, sync=True):
        self._ck(self.lib.photon_scene_trace(self.handle), "trace")
        if sync:
            self.sync()

    def trace_range(self, begin, count, sync=True):
        self._ck(self.lib.photon_scene_trace_range(self.handle, int(begin), int(count)), "trace_range")
        if sync:
            self.sync()

    def sync(self):
        self._ck(self.lib.photon_scene_sync(self.handle), "sync")

    def trace_batch(self, batch, n_save, n_inter=0):
        """Returns (final_pos, final_dir[, inter_pos, inter_dir]) as float32 arrays."""
        pos = np.empty((n_save, 3), np.float32)
        dr = np.empty((n_save, 3), np.float32)
        ip = idr = None
        if n_inter > 0:
            ip = np.empty((n_save, n_inter, 3), np.float32)
            idr = np.empty((n_save, n_inter, 3), np.float32)
        self._ck(self.lib.photon_scene_trace_batch(self.handle, batch, n_save, pos.ctypes.data, dr.ctypes.data, n_inter,
                                                   ip.ctypes.data if ip is not None else None,
                                                   idr.ctypes.data if idr is not None else None), "trace_batch")
        return (pos, dr) if n_inter <= 0 else (pos, dr, ip, idr)

    # dot statistics: light_ray_processing.py's dump consumer on the device (include/photon_b200.h)
    def enable_dot_stats(self, rays_per_dot):
        self._ck(self.lib.photon_scene_enable_dot_stats(self.handle, int(rays_per_dot)), "enable_dot_stats")
        self._rays_per_dot = int(rays_per_dot)

    def reset_dot_stats(self):
        self._ck(self.lib.photon_scene_reset_dot_stats(self.handle), "reset_dot_stats")

    def download_dot_stats(self):
        """float64 [n_dots, 8]: sums of sensor x, y, z; sums of arccos(direction cosine) x, y, z; rays with a
        position; rays with a direction."""
        n = int(self.lib.photon_scene_num_dots(self.handle))
        out = np.zeros((n, 8), np.float64)
        self._ck(self.lib.photon_scene_download_dot_stats(self.handle, out.ctypes.data, n), "download_dot_stats")
        return out

    def reset_counters(self):
        self._ck(self.lib.photon_scene_reset_counters(self.handle), "reset_counters")

    @property
    def ray_steps(self):
        return int(self.lib.photon_scene_ray_steps(self.handle))

    @property
    def rays(self):
        return int(self.lib.photon_scene_rays(self.handle))

    @property
    def sensor_hits(self):
        return int(self.lib.photon_scene_sensor_hits(self.handle))

    @property
    def last_trace_ms(self):
        return float(self.lib.photon_scene_last_trace_ms(self.handle))

    @property
    def kernel_launches(self):
        return int(self.lib.photon_scene_kernel_launches(self.handle))

    @property
    def num_points(self):
        return int(self.lib.photon_scene_num_points(self.handle))

    @property
    def image_device_ptr(self):
        return int(self.lib.photon_scene_image_device_ptr(self.handle))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.photon_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
