#!/usr/bin/env python3
"""Quick device-side timing probe: python scripts/probe.py SCENE_NAME W H SPP [stats] [reps]
SCENE_NAME is a key of scenes/gen_scenes.py (generated into scenes/_cache and loaded through
the C++ host library)."""
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))
import fs2_b200
from fs2_b200 import host
import gen_scenes

name, w, h, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
stats = len(sys.argv) > 5 and sys.argv[5] == "stats"
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 3
path = gen_scenes.generate(os.path.join(REPO, "scenes", "_cache"), [name])[name]
t0 = time.time()
hs = host.HostScene(path)
desc = hs.desc()
print("host load+build %.2fs (%.1f MB)" % (time.time() - t0, hs.nbytes() / 1e6))
cam = hs.camera()
ctx = fs2_b200.Context(0, stats=stats)
ctx.upload(desc)
ctx.resize(w, h)
ctx.trace(cam, 1, seed=1)
for rep in range(reps):
    ctx.reset()
    ctx.reset_stats()
    ctx.trace(cam, spp, seed=2 + rep)
    st = ctx.stats()
    ms = st["last_kernel_ms"]
    print("rep %d: %.2f ms kernel (%.2f total)  %.2f Mpaths/s  %.2f Mrays/s  seg/path %.2f" %
          (rep, ms, st["last_trace_ms"], st["n_paths"] / ms / 1e3, st["n_segments"] / ms / 1e3,
           st["n_segments"] / st["n_paths"]),
          {k: v for k, v in st.items() if k.startswith("n_")} if stats else "")
