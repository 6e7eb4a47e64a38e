"""Several GPUs inside one start_ray_tracing call (PHOTON_GPUS): source points split into contiguous
shards, field replicated, partial images combined onto device 0 by the library's own kernel over NVLink peer memory
(default), by one ncclReduce (PHOTON_REDUCE=nccl) or through the host (PHOTON_REDUCE=host).  Needs two devices; skipped
on a one-GPU box."""
import os

import numpy as np
import pytest

from photon_b200 import scenes

pytestmark = pytest.mark.gpu


def _run(call, lib, base, **env):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        img = base.copy()
        return call.run(lib, image=img)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.fixture(scope="module")
def setup(tmp_path_factory, cuda_lib):
    if cuda_lib.photon_b200_device_count() < 2:
        pytest.skip("needs two GPUs")
    d = tmp_path_factory.mktemp("multi")
    call = scenes.c3_bos(str(d / "blob.nrrd"), n=64, dots_per_side=12, points_per_dot=8, rays_per_point=128, nx=512, ny=512)
    base = np.random.default_rng(5).uniform(0, 3, call.image_shape).astype(np.float32)     # accumulate-onto input
    return call, cuda_lib, base


REDUCE = ["peer", "nccl", "host"]


@pytest.mark.parametrize("reduce", REDUCE)
def test_split_call_matches_one_gpu(setup, reduce):
    call, lib, base = setup
    one = _run(call, lib, base, PHOTON_GPUS=1)
    two = _run(call, lib, base, PHOTON_GPUS=2, PHOTON_REDUCE=reduce)
    assert (one - base).sum() > 0
    rel = np.linalg.norm(one.astype(np.float64) - two) / np.linalg.norm(one.astype(np.float64) - base)
    assert rel < 1e-5, rel                       # float atomics: order of the additions only


@pytest.mark.parametrize("reduce", REDUCE)
def test_split_call_is_bit_exact_in_deterministic_mode(setup, reduce):
    call, lib, base = setup
    zero = np.zeros_like(base)
    one = _run(call, lib, zero, PHOTON_GPUS=1, PHOTON_DETERMINISTIC=1)
    two = _run(call, lib, zero, PHOTON_GPUS=2, PHOTON_DETERMINISTIC=1, PHOTON_REDUCE=reduce)
    np.testing.assert_array_equal(one, two)


def test_peer_reduce_gives_the_same_image_run_after_run(setup):
    """The peer-memory reduce adds the shard images in device order: with the shards fixed, two calls differ only by
    the float atomics inside each shard -- and not at all in deterministic mode (covered above).  Here: the default
    route is the peer kernel and it agrees with the NCCL route to summation order."""
    call, lib, base = setup
    a = _run(call, lib, base, PHOTON_GPUS=2)
    b = _run(call, lib, base, PHOTON_GPUS=2, PHOTON_REDUCE="nccl")
    rel = np.linalg.norm(a.astype(np.float64) - b) / np.linalg.norm(a.astype(np.float64) - base)
    assert rel < 1e-5, rel
