#!/usr/bin/env python3
"""Quick device-side timing probe: python scripts/probe.py BLOB W H SPP [stats]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fs2_b200
from fs2_b200 import abi

blob, w, h, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
stats = len(sys.argv) > 5 and sys.argv[5] == "stats"
scene = abi.FlatScene.load(blob)
cam = abi.camera_from_file(blob + ".camera")
ctx = fs2_b200.Context(0, stats=stats)
t0 = time.time()
ctx.upload(scene)
print("upload %.3fs (%.1f MB)" % (time.time() - t0, scene.nbytes() / 1e6))
ctx.resize(w, h)
ctx.trace(cam, 1, seed=1)
for rep in range(3):
    ctx.reset()
    ctx.reset_stats()
    t0 = time.time()
    ctx.trace(cam, spp, seed=2 + rep)
    wall = time.time() - t0
    st = ctx.stats()
    ms = st["last_trace_ms"]
    print("rep %d: %.2f ms device (%.2f wall)  %.2f Mpaths/s  %.2f Mrays/s  seg/path %.2f" %
          (rep, ms, wall * 1e3, st["n_paths"] / ms / 1e3, st["n_segments"] / ms / 1e3,
           st["n_segments"] / st["n_paths"]), {k: v for k, v in st.items() if k.startswith("n_")})
fh = ctx.first_hit(cam, w, h)
print("first-hit coverage %.3f" % (fh["object"] != 0xFFFFFFFF).mean())
