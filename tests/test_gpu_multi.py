"""Multi-GPU paths against the single-GPU film of the same seed (the N-GPU sum
must equal the 1-GPU result run with the same per-sample seeds within 1 ulp x spp).

  * tile sharding emulated on one GPU (fs_set_tile_shard, ranks rendered one after the other
    into one film) — runs everywhere;
  * the single-process multi-device context (fs_create_multi: host threads + the peer-memory
    combine kernel), sample and tile sharding, through the C ABI, the C++ host mirror
    (SetTraceDevices / TraceSamples) and fs2_render --gpus — needs >= 2 visible devices;
  * the multi-process path (one process per GPU, NCCL reduce ordered on the context's stream,
    multi_gpu.render_sharded) — needs >= 2 visible devices."""
import os
import subprocess

import numpy as np
import pytest
import torch

import fs2_b200
from fs2_b200 import abi, build, host, multi_gpu
from refdata import golden

pytestmark = pytest.mark.gpu
N_DEV = torch.cuda.device_count() if torch.cuda.is_available() else 0
need2 = pytest.mark.skipif(N_DEV < 2, reason="needs two CUDA devices")


def _film(ctx, cam, spp, seed, flags=0):
    ctx.reset()
    ctx.trace(cam, spp, seed=seed, flags=flags)
    f = ctx.read_film(aux=True)
    return {k: v.copy() for k, v in f.items()}


def _same_image(a, b, spp):
    """Counts exact, aux planes bit-identical, radiance within spp ulp of the larger value."""
    assert (a["count"] == spp).all() and (b["count"] == spp).all()
    assert (a["mat_id"] == b["mat_id"]).all()
    assert (a["normal"].view(np.uint32) == b["normal"].view(np.uint32)).all()
    assert (a["depth"].view(np.uint32) == b["depth"].view(np.uint32)).all()
    x, y = a["mean"].astype(np.float64), b["mean"].astype(np.float64)
    scale = np.maximum(np.maximum(np.abs(x), np.abs(y)), 1e-30)
    ulps = np.abs(x - y) / scale / 2.0 ** -23
    assert float(ulps.max()) <= spp, float(ulps.max())
    return float(ulps.max())


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("size", [(96, 64), (50, 37)])
def test_tile_shards_partition_the_image(world, size):
    """Every rank's tiles, rendered one after the other into one film on one GPU, are the whole
    image: each pixel belongs to exactly one rank ((tx + ty) mod world) and gets all its samples."""
    g = golden("mesh_tiny")
    w, h = size
    spp = 3
    with fs2_b200.Context(device=0) as c:
        c.upload(g.scene)
        c.resize(w, h)
        whole = _film(c, g.cam, spp, 31)
        c.reset()
        for r in range(world):
            c.set_tile_shard(r, world)
            c.trace(g.cam, spp, seed=31, flags=abi.FS_TRACE_NO_RESOLVE)
            part = c.save_film().reshape(h, w, 4)[..., 3]
            ys, xs = np.mgrid[0:h, 0:w]
            owned = multi_gpu.tile_owner(xs, ys, world) <= r
            assert (part[owned] == spp).all() and (part[~owned] == 0).all()
        c.set_tile_shard(0, 1)
        c.resolve()
        parts = {k: v.copy() for k, v in c.read_film(aux=True).items()}
    _same_image(parts, whole, spp)


@need2
@pytest.mark.parametrize("name", ["mesh_tiny", "glass_small", "many303"])
@pytest.mark.parametrize("mode", ["samples", "tiles", "auto_low_spp"])
def test_multi_device_context_equals_single_device(name, mode):
    g = golden(name)
    w, h = 112, 72
    spp = 1 if mode == "auto_low_spp" else 7
    flags = {"samples": abi.FS_TRACE_SHARD_SAMPLES, "tiles": abi.FS_TRACE_SHARD_TILES, "auto_low_spp": 0}[mode]
    with fs2_b200.Context(device=0) as c:
        c.upload(g.scene)
        c.resize(w, h)
        one = _film(c, g.cam, spp, 17)
    devices = list(range(min(N_DEV, 4)))
    with fs2_b200.Context(devices=devices, stats=True) as m:
        m.upload(g.scene)
        m.resize(w, h)
        many = _film(m, g.cam, spp, 17, flags)
        st = m.stats()
        assert st["n_paths"] == w * h * spp  # counters of the whole job, summed over devices
        # accumulation across calls: 2 more samples on top, still the single-device sample set
        m.trace(g.cam, 2, first_sample=spp, seed=17, flags=flags)
        more = {k: v.copy() for k, v in m.read_film(aux=True).items()}
    _same_image(many, one, spp)
    with fs2_b200.Context(device=0) as c:
        c.upload(g.scene)
        c.resize(w, h)
        one9 = _film(c, g.cam, spp + 2, 17)
    _same_image(more, one9, spp + 2)


@need2
def test_host_mirror_and_cli_use_every_requested_gpu(tmp_path):
    """base::SetTraceDevices + TraceSamples, and fs2_render --gpus 2 against --gpus 1."""
    import gen_scenes
    scene = gen_scenes.generate(str(tmp_path), ["mesh_tiny"])["mesh_tiny"]
    hs = host.HostScene(scene)
    cam = hs.camera()
    w, h, spp = 96, 64, 6
    films = []
    for devices in ([0], [0, 1]):
        host.set_trace_devices(devices)
        hs.lib.fs2h_set_seed(5)
        frame = host.HostFrame(w, h)
        frame.trace(hs, cam, spp)
        assert hs.lib.fs2h_device_count() == len(devices)
        films.append((frame.mean().copy(), frame.count().copy()))
        frame.close()
    host.set_trace_devices([0])
    (m1, c1), (m2, c2) = films
    assert (c1 == spp).all() and (c2 == spp).all()
    assert np.allclose(m1, m2, rtol=spp * 2.0 ** -23, atol=1e-30)
    exe = build.build_cli()
    out = []
    for gpus in ("1", "2"):
        pfm = str(tmp_path / ("g%s.pfm" % gpus))
        r = subprocess.run([exe, "-f", scene, "-w", "80", "-h", "48", "--spp", "5", "--seed", "9", "--gpus", gpus,
                            "--out", str(tmp_path / ("g%s.ppm" % gpus)), "--pfm", pfm], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        assert ("on %s GPU" % gpus) in r.stdout
        out.append(np.frombuffer(open(pfm, "rb").read()[len(b"PF\n80 48\n-1.0\n"):], np.float32))
    assert np.allclose(out[0], out[1], rtol=5 * 2.0 ** -23, atol=1e-30)


def _nccl_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    g = golden("mesh_tiny")
    w, h = 112, 72
    ctx = fs2_b200.Context(device=rank)
    ctx.upload(g.scene)
    ctx.resize(w, h)
    for mode, spp, tiles in (("samples", 9, False), ("tiles", 9, True), ("auto", 1, None)):
        # several images back to back: a reduce that is not ordered with the next reset, or a
        # resolve that does not wait for the reduce, shows as wrong counts or a torn image
        for rep in range(3):
            multi_gpu.render_sharded(ctx, g.cam, spp, 40 + rep, rank, world, tiles=tiles)
        if rank == 0:
            f = ctx.read_film(aux=True)
            np.savez(os.path.join(out_dir, mode + ".npz"), **f)
        dist.barrier()
    ctx.close()
    dist.destroy_process_group()


@need2
def test_nccl_sharded_render_equals_single_gpu(tmp_path):
    import torch.multiprocessing as mp
    port = 29600 + (os.getpid() % 2000)
    mp.spawn(_nccl_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    g = golden("mesh_tiny")
    with fs2_b200.Context(device=0) as c:
        c.upload(g.scene)
        c.resize(112, 72)
        for mode, spp in (("samples", 9), ("tiles", 9), ("auto", 1)):
            one = _film(c, g.cam, spp, 42)
            many = dict(np.load(tmp_path / (mode + ".npz")))
            _same_image(many, one, spp)
