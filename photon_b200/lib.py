"""Loads the CUDA library.  There is no CPU fallback: a missing library or a machine
without a CUDA device is an error the caller sees."""
import ctypes as C
import os

from . import build as _build
from .abi import bind_start_ray_tracing

_lib = None


class PhotonLibraryError(RuntimeError):
    pass


def library_path():
    return os.environ.get("PHOTON_B200_LIB") or _build.LIB


def _point_at_bundled_nccl():
    """PHOTON_REDUCE=nccl makes the library dlopen NCCL.  A Python process that also imports torch must end up with ONE
    libnccl.so.2: point the library at the file PyTorch bundles (the nvidia-nccl wheel), found without importing torch."""
    if os.environ.get("PHOTON_NCCL_LIB"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else []):
            f = os.path.join(d, "lib", "libnccl.so.2")
            if os.path.exists(f):
                os.environ["PHOTON_NCCL_LIB"] = f
                return
    except Exception:
        pass


def load_library(require_gpu=False):
    """dlopen libparallel_ray_tracing.so (prebuilt in-tree) and declare the extended ABI."""
    global _lib
    if _lib is None:
        path = library_path()
        if not os.path.exists(path):
            raise PhotonLibraryError(
                f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  photon_b200 has no CPU fallback.")
        _point_at_bundled_nccl()
        lib = C.CDLL(path)
        bind_start_ray_tracing(lib)
        lib.photon_b200_abi_version.restype = C.c_int
        lib.photon_b200_device_count.restype = C.c_int
        lib.photon_b200_last_error.restype = C.c_char_p
        lib.photon_scene_create.restype = C.c_void_p
        lib.photon_scene_destroy.argtypes = [C.c_void_p]
        lib.photon_scene_destroy.restype = None
        for name in ("photon_scene_zero_image", "photon_scene_trace", "photon_scene_sync",
                     "photon_scene_reset_counters", "photon_scene_kernel_launches", "photon_scene_num_points"):
            f = getattr(lib, name)
            f.argtypes = [C.c_void_p]
            f.restype = C.c_int
        for name in ("photon_scene_upload_image", "photon_scene_download_image"):
            f = getattr(lib, name)
            f.argtypes = [C.c_void_p, C.c_void_p]
            f.restype = C.c_int
        lib.photon_scene_deterministic.argtypes = [C.c_void_p]
        lib.photon_scene_deterministic.restype = C.c_int
        lib.photon_scene_download_image_fixed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.photon_scene_download_image_fixed.restype = C.c_int
        lib.photon_scene_image_device_ptr.argtypes = [C.c_void_p]
        lib.photon_scene_image_device_ptr.restype = C.c_void_p
        for name in ("photon_scene_ray_steps", "photon_scene_rays", "photon_scene_sensor_hits"):
            f = getattr(lib, name)
            f.argtypes = [C.c_void_p]
            f.restype = C.c_longlong
        lib.photon_scene_last_trace_ms.argtypes = [C.c_void_p]
        lib.photon_scene_last_trace_ms.restype = C.c_float
        lib.photon_scene_trace_batch.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p,
                                                 C.c_int, C.c_void_p, C.c_void_p]
        lib.photon_scene_trace_batch.restype = C.c_int
        lib.photon_b200_load_field.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_longlong]
        lib.photon_b200_load_field.restype = C.c_int
        lib.photon_scene_trace_range.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.photon_scene_trace_range.restype = C.c_int
        lib.photon_b200_register_field.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        lib.photon_b200_register_field.restype = C.c_int
        lib.photon_b200_unregister_field.argtypes = [C.c_char_p]
        lib.photon_b200_unregister_field.restype = C.c_int
        lib.photon_scene_export_u16.argtypes = [C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_float, C.c_uint, C.c_void_p]
        lib.photon_scene_export_u16.restype = C.c_int
        lib.photon_scene_enable_dot_stats.argtypes = [C.c_void_p, C.c_longlong]
        lib.photon_scene_enable_dot_stats.restype = C.c_int
        lib.photon_scene_num_dots.argtypes = [C.c_void_p]
        lib.photon_scene_num_dots.restype = C.c_longlong
        lib.photon_scene_reset_dot_stats.argtypes = [C.c_void_p]
        lib.photon_scene_reset_dot_stats.restype = C.c_int
        lib.photon_scene_download_dot_stats.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong]
        lib.photon_scene_download_dot_stats.restype = C.c_int
        lib.photon_b200_release_cache.argtypes = []
        lib.photon_b200_release_cache.restype = C.c_longlong
        lib.photon_b200_cache_bytes.argtypes = []
        lib.photon_b200_cache_bytes.restype = C.c_longlong
        lib.photon_b200_last_variant.argtypes = []
        lib.photon_b200_last_variant.restype = C.c_char_p
        lib.photon_b200_brick_timeouts.argtypes = []
        lib.photon_b200_brick_timeouts.restype = C.c_int
        lib.photon_b200_sample_field.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        lib.photon_b200_sample_field.restype = C.c_int
        lib.photon_b200_measure_fp32_tflops.argtypes = [C.c_int]
        lib.photon_b200_measure_fp32_tflops.restype = C.c_double
        lib.photon_b200_measure_fp32x2_tflops.argtypes = [C.c_int]
        lib.photon_b200_measure_fp32x2_tflops.restype = C.c_double
        lib.photon_b200_export_u16_static.argtypes = [C.c_void_p, C.c_longlong, C.c_float, C.c_int, C.c_int, C.c_float, C.c_uint,
                                                      C.c_void_p, C.c_void_p]
        lib.photon_b200_export_u16_static.restype = C.c_int
        _lib = lib
    if require_gpu and _lib.photon_b200_device_count() < 1:
        raise PhotonLibraryError("no CUDA device visible; photon_b200 has no CPU fallback")
    return _lib
