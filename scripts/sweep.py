#!/usr/bin/env python3
"""BASELINE.json config 5: resolution x spp sweep of configs 2 and 3 on one GPU (the 1/2/4/8-GPU
axis is bench.py under torchrun).  Prints one JSON line per point.

    python scripts/sweep.py [scene ...]        # default: mesh100k glass
"""
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))

import fs2_b200  # noqa: E402
import gen_scenes  # noqa: E402
from fs2_b200 import abi, host  # noqa: E402

RES = [(1280, 720), (1920, 1080), (2560, 1440), (3840, 2160)]
SPP = [16, 64, 256, 1024]
scenes = sys.argv[1:] or ["mesh100k", "glass"]
cache = os.path.join(REPO, "scenes", "_cache")
for scene in scenes:
    hs = host.HostScene(gen_scenes.generate(cache, [scene])[scene])
    with fs2_b200.Context(device=0) as ctx:
        ctx.upload(hs.desc())
        cam = hs.camera()
        ctx.resize(640, 360)
        ctx.trace(cam, 4, seed=1, flags=abi.FS_TRACE_NO_RESOLVE)  # warm-up
        for (w, h) in RES:
            ctx.resize(w, h)
            for spp in SPP:
                ctx.reset()
                ctx.reset_stats()
                ctx.trace(cam, spp, seed=7)
                st = ctx.stats()
                ms = st["last_trace_ms"]
                print(json.dumps({"scene": scene, "width": w, "height": h, "spp": spp, "n_gpus": 1,
                                  "time_to_image_ms": round(ms, 3),
                                  "mpaths_per_s": round(w * h * spp / ms / 1e3, 1),
                                  "mrays_per_s": round(st["n_segments"] / ms / 1e3, 1)}), flush=True)
