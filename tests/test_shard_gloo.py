"""CPU, world_size 2, gloo: the N>1 host logic (contiguous source-point shards + ONE reduce of the
partial sensor images).  The tracer behind it here is the CPU oracle (no GPU in this container);
on the GPU box bench.py runs the same logic over NCCL with the CUDA library."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, workdir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="2")
    import torch.distributed as dist
    from oracle.loader import load_oracle
    from photon_b200 import scenes
    from photon_b200.shard import reduce_image, shard_call, shard_range
    dist.init_process_group("gloo", rank=rank, world_size=world)
    call = scenes.c3_bos(os.path.join(workdir, "blob32.nrrd"), n=32, dots_per_side=6, points_per_dot=5,
                         rays_per_point=16, write_field=(rank == 0))
    dist.barrier()
    lib = load_oracle(build=False)
    part = shard_call(call, rank, world).run(lib)
    b, e = shard_range(call.num_particles, rank, world)
    total = reduce_image(part.copy(), dst=0)
    np.save(os.path.join(workdir, f"part{rank}.npy"), part)
    if rank == 0:
        np.save(os.path.join(workdir, "reduced.npy"), total)
        np.save(os.path.join(workdir, "full.npy"), call.run(lib))
    open(os.path.join(workdir, f"range{rank}.txt"), "w").write(f"{b} {e} {call.num_particles}")
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_reduce_to_the_full_image(tmp_path, oracle_lib):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0 = [int(v) for v in open(tmp_path / "range0.txt").read().split()]
    r1 = [int(v) for v in open(tmp_path / "range1.txt").read().split()]
    assert r0[0] == 0 and r0[1] == r1[0] and r1[1] == r1[2]                 # contiguous, complete, disjoint
    full = np.load(tmp_path / "full.npy").astype(np.float64)
    red = np.load(tmp_path / "reduced.npy").astype(np.float64)
    p0, p1 = np.load(tmp_path / "part0.npy"), np.load(tmp_path / "part1.npy")
    assert p0.sum() > 0 and p1.sum() > 0
    np.testing.assert_array_equal(red, (p0 + p1).astype(np.float64))        # the reduce is a plain sum
    assert np.linalg.norm(red - full) <= 1e-6 * np.linalg.norm(full)


def test_shard_ranges_cover_every_point_once():
    from photon_b200.shard import shard_range
    for n in (0, 1, 7, 10000, 1200000):
        for w in (1, 2, 3, 8):
            edges = [shard_range(n, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(w - 1))
            assert max(e - b for b, e in edges) - min(e - b for b, e in edges) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
