#!/usr/bin/env python3
"""BASELINE.json config 5: resolution x spp sweep of configs 2 and 3 at 1 / 2 / 4 / 8 GPUs.
One process; N > 1 goes through one multi-device context (fs_create_multi: the path behind
base::SetTraceDevices / fs2_render --gpus).  Times are host wall clock around the synchronous
fs_trace call (render + film combine + resolve), after one untimed call per context.  Prints one
JSON line per point.

    python scripts/sweep.py [--gpus 1,2,4,8] [scene ...]        # default: 1 GPU, mesh100k glass3
"""
import json
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))

import fs2_b200  # noqa: E402
import gen_scenes  # noqa: E402
from fs2_b200 import host  # noqa: E402

RES = [(1280, 720), (1920, 1080), (2560, 1440), (3840, 2160)]
SPP = [16, 64, 256, 1024]
argv = sys.argv[1:]
gpus = [1]
if "--gpus" in argv:
    i = argv.index("--gpus")
    gpus = [int(x) for x in argv[i + 1].split(",")]
    del argv[i:i + 2]
scenes = argv or ["mesh100k", "glass3"]
cache = os.path.join(REPO, "scenes", "_cache")
for scene in scenes:
    hs = host.HostScene(gen_scenes.generate(cache, [scene])[scene])
    cam = hs.camera()
    for n in gpus:
        ctx = fs2_b200.Context(device=0) if n == 1 else fs2_b200.Context(devices=list(range(n)))
        with ctx:
            ctx.upload(hs.desc())
            ctx.resize(1920, 1080)
            ctx.trace(cam, 64, seed=1)  # warm-up: clocks, path pools
            for (w, h) in RES:
                ctx.resize(w, h)
                for spp in SPP:
                    ctx.reset()
                    t0 = time.perf_counter()
                    ctx.trace(cam, spp, seed=7)
                    ms = (time.perf_counter() - t0) * 1e3
                    print(json.dumps({"scene": scene, "width": w, "height": h, "spp": spp, "n_gpus": n,
                                      "time_to_image_ms": round(ms, 3),
                                      "mpaths_per_s": round(w * h * spp / ms / 1e3, 1)}), flush=True)
