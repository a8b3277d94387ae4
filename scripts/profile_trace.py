#!/usr/bin/env python3
"""Short profiling driver (ncu target): uploads the bench scene once and traces a few samples
per pixel with one kernel variant.  bench.py is too long to sit under ncu's replay.

    python scripts/profile_trace.py [variant] [spp] [width] [height] [scene]
"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))

import fs2_b200  # noqa: E402
import gen_scenes  # noqa: E402
from fs2_b200 import abi, host  # noqa: E402

variant = sys.argv[1] if len(sys.argv) > 1 else "wavefront"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
width = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
height = int(sys.argv[4]) if len(sys.argv) > 4 else 1080
scene = sys.argv[5] if len(sys.argv) > 5 else "mesh100k"
flags = {"phase": abi.FS_TRACE_MEGAKERNEL, "wavefront": abi.FS_TRACE_DEFAULT, "lbvh_lane": abi.FS_TRACE_LBVH_LANE,
         "coop": abi.FS_TRACE_COOP, "per_thread": abi.FS_TRACE_PER_THREAD}[variant]
cache = os.path.join(REPO, "scenes", "_cache")
path = gen_scenes.generate(cache, [scene])[scene]
hs = host.HostScene(path)
ctx = fs2_b200.Context(device=0)
ctx.upload(hs.desc())
ctx.resize(width, height)
for i in range(2):
    ctx.reset()
    ctx.trace(hs.camera(), spp, seed=5 + i, flags=flags | abi.FS_TRACE_NO_RESOLVE)
    st = ctx.stats()
    print("%s %dx%d %d spp: %.2f ms (kernel %.2f ms), %.1f Mpaths/s, %d launches" %
          (variant, width, height, spp, st["last_trace_ms"], st["last_kernel_ms"],
           width * height * spp / st["last_trace_ms"] / 1e3, st["launches"]))
ctx.close()
