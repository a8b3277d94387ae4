"""Host-side logic of the multi-GPU path on CPU: sample sharding and the film reduce over a
world_size-2 gloo group (the GPU path runs the same functions over NCCL)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import fs2_b200
from fs2_b200 import multi_gpu


def test_shard_samples_partitions_exactly():
    for total in (0, 1, 7, 64, 257):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                first, n, stride = multi_gpu.shard_samples(total, r, world)
                seen += [first + k * stride for k in range(n)]
            assert sorted(seen) == list(range(total))
            sizes = [multi_gpu.shard_samples(total, r, world)[1] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
            if total:
                # rank 0 owns the last sample: its aux buffers are the image's (last sample wins)
                first, n, stride = multi_gpu.shard_samples(total, 0, world)
                assert first + (n - 1) * stride == total - 1


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_px, total = 500, 10
    rng = np.random.RandomState(123)
    samples = rng.rand(total, n_px, 3).astype(np.float32)  # every rank knows all samples
    first, n, stride = multi_gpu.shard_samples(total, rank, world)
    film = torch.zeros(n_px, 4)
    for k in range(n):
        film[:, :3] += torch.from_numpy(samples[first + k * stride])
        film[:, 3] += 1
    flat = film.reshape(-1)
    multi_gpu.reduce_film(flat, 0)
    if rank == 0:
        np.save(os.path.join(out_dir, "film.npy"), flat.reshape(n_px, 4).numpy())
        np.save(os.path.join(out_dir, "want.npy"), samples.sum(0))
    dist.barrier()
    dist.destroy_process_group()


def test_film_reduce_world2_gloo(tmp_path):
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    film = np.load(tmp_path / "film.npy")
    want = np.load(tmp_path / "want.npy")
    assert (film[:, 3] == 10).all()
    assert np.allclose(film[:, :3], want, rtol=1e-5, atol=1e-5)
    mean = film[:, :3] / film[:, 3:4]
    assert np.allclose(mean, want / 10, rtol=1e-5, atol=1e-6)


def _tile_worker(rank, world, port, out_dir):
    """Tile sharding: every pixel's guide-plane record (normal, depth) lives on its owner rank and
    is zero elsewhere; the gather is an integer sum of the float words, so the bit pattern —
    a -0.0 normal component included — arrives intact (a float sum turns -0.0 + 0.0 into +0.0;
    seen on the 4K config-4 frame)."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w, h = 64, 24
    rng = np.random.RandomState(5)
    full = rng.randn(h, w, 4).astype(np.float32)
    full[::3, ::5, 1] = -0.0
    ys, xs = np.mgrid[0:h, 0:w]
    mine = multi_gpu.tile_owner(xs, ys, world) == rank
    aux = np.where(mine[..., None], full, np.float32(0.0)).astype(np.float32)
    words = torch.from_numpy(aux.view(np.int32).reshape(-1).copy())
    multi_gpu.reduce_film(words, 0)
    if rank == 0:
        np.save(os.path.join(out_dir, "aux.npy"), words.numpy().view(np.float32).reshape(h, w, 4))
        np.save(os.path.join(out_dir, "full.npy"), full)
    dist.barrier()
    dist.destroy_process_group()


def test_tile_gather_keeps_bit_patterns_world2_gloo(tmp_path):
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_tile_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(tmp_path / "aux.npy")
    want = np.load(tmp_path / "full.npy")
    assert (got.view(np.uint32) == want.view(np.uint32)).all()
    assert (np.signbit(want) & (want == 0)).any()  # the fixture really holds -0.0
