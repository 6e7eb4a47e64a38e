"""Source-point sharding and the single exchange step of the path.

Rays are independent; only the sensor image is shared.  Every rank traces a contiguous range of
the source points with the density field replicated on its GPU, then the partial images are summed
onto one rank with ONE reduce (NCCL on GPUs, gloo in the CPU tests).  The range formula is the one
the library applies for PHOTON_SHARD=r/n and photon_scene_create(..., shard_index, shard_count)
(photon_b200/csrc/driver.cu, scene_fill).
"""
import numpy as np


def shard_range(num_points, rank, world):
    """[begin, end) of rank's contiguous source-point shard."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad shard {rank}/{world}")
    return (num_points * rank) // world, (num_points * (rank + 1)) // world


def shard_call(call, rank, world):
    """The RayTracingCall restricted to rank's shard (what PHOTON_SHARD does inside the library)."""
    b, e = shard_range(call.num_particles, rank, world)
    return call.subset(slice(b, e))


def reduce_image(image, dst=0, group=None):
    """Sum the ranks' partial images onto rank ``dst`` with one torch.distributed reduce.
    ``image``: numpy array (CPU/gloo) or torch tensor (CUDA/NCCL).  float32 images are summed in
    float32; int64 fixed-point images (deterministic mode) are summed exactly.  Returns the reduced
    array on ``dst`` (other ranks get their input back)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return image
    is_np = isinstance(image, np.ndarray)
    t = torch.from_numpy(np.ascontiguousarray(image)) if is_np else image
    if dist.get_backend(group) == "nccl" and not t.is_cuda:
        t = t.cuda()
    dist.reduce(t, dst=dst, op=dist.ReduceOp.SUM, group=group)
    if is_np:
        return t.cpu().numpy()
    return t
