#!/usr/bin/env python3
"""compute-sanitizer target: a small pass through every device entry point (scene upload with a
GPU-built octree, first hit, the default wavefront trace and the megakernel, film read-back,
post-processing, known-answer selftest) at sizes memcheck finishes in seconds.

    compute-sanitizer --tool memcheck python scripts/sanitize_small.py
"""
import ctypes as C
import os
import sys
import tempfile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))

import numpy as np  # noqa: E402

import fs2_b200  # noqa: E402
import gen_scenes  # noqa: E402
from fs2_b200 import abi, engine, host  # noqa: E402

names = sys.argv[1:] or ["demo_dof", "mesh_tiny", "glass_small", "many303"]
tmp = tempfile.mkdtemp(prefix="fs2_sanitize_")
paths = gen_scenes.generate(tmp, names)
lib = host.load_host_library()
ctx = fs2_b200.Context(device=0)
for k, name in enumerate(names):
    lib.fs2h_set_gpu_build(0 if k % 2 else -1)  # alternate host / GPU octree build
    hs = host.HostScene(paths[name])
    cam = hs.camera()
    ctx.upload(hs.desc())
    w, h = 50 + 7 * k, 30 + 3 * k  # not multiples of the 16x8 tile
    hit = ctx.first_hit(cam, w, h)
    ctx.resize(w, h)
    ctx.trace(cam, 3, seed=1)
    ctx.trace(cam, 2, first_sample=3, seed=1, flags=abi.FS_TRACE_MEGAKERNEL)
    film = ctx.read_film(aux=True)
    assert (film["count"] == 5).all() and np.isfinite(film["mean"]).all()
    post = ctx.post_film(abi.post_params(denoise_passes=2, bloom_strength=0.3, bloom_radius=3,
                                         flare_strength=0.1, flare_ghosts=2))
    assert np.isfinite(post["rgb"]).all()
    print("%s %dx%d: hits %d, mean %.4f" % (name, w, h, int((hit["object"] != 0xFFFFFFFF).sum()),
                                            float(film["mean"].mean())))
lib.fs2h_set_gpu_build(-1)
words = np.zeros(48, np.uint32)
assert engine.load_library().fs_device_selftest(0, words.ctypes.data_as(C.c_void_p), 48) == 0
ctx.close()
print("sanitize_small: done")
