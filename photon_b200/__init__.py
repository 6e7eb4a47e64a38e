"""photon_b200 -- B200-native drop-in for photon's ray-tracing core.

The product is ``photon_b200/lib/libparallel_ray_tracing.so`` (hand-written CUDA for
sm_100a behind photon's C ABI, see include/photon_abi.h and include/photon_b200.h).
This package is the thin host side: ctypes mirror of the ABI, scene builders, NRRD I/O
and the staging harness that runs the reference's unchanged Python scripts on it.
"""
from .abi import RayTracingCall  # noqa: F401

__all__ = ["RayTracingCall"]
