#!/usr/bin/env python3
"""Generates the committed golden fixtures from the UNMODIFIED reference renderer
(oracle/_ref/fs2_ref, built by oracle/ref/Makefile from /root/reference/source).

    python tests/golden/make_golden.py            # needs /root/reference (this container)

For every small scene it stores, in tests/golden/NAME.npz:
  * the flattened scene exactly as the reference built it (object list, octrees,
    materials, textures, normal LUT) and the camera,
  * the reference's first hit per pixel (TraceRange ray construction),
  * the reference's float film after FRAMES TraceScene calls with BANDS worker threads and
    the scripted per-frame clock (so the xorshift64* streams can be replayed).
The GPU box has no /root/reference: tests there only read these files.
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "scenes"))

import fs2_b200  # noqa: E402,F401
from fs2_b200 import abi  # noqa: E402
from oracle import binding  # noqa: E402
import gen_scenes  # noqa: E402

W, H, FRAMES, BANDS = 96, 64, 2, 4
SCENES = ["api", "demo", "demo_dof", "many303", "mesh_tiny", "glass_small", "grid_tiny"]


def main():
    if binding.build_ref() is None:
        raise SystemExit("oracle/_ref/fs2_ref is not available (no /root/reference?)")
    with tempfile.TemporaryDirectory() as tmp:
        paths = gen_scenes.generate(tmp, [s for s in SCENES if s != "api"])
        gen_scenes.checker_texture(os.path.join(tmp, "checker.bmp"))
        paths["api"] = "@api:" + os.path.join(tmp, "checker.bmp")
        for name in SCENES:
            blob = os.path.join(tmp, name + ".fs2")
            fh = os.path.join(tmp, name + ".fh")
            film = os.path.join(tmp, name + ".film")
            binding.run_ref(paths[name], W, H, flatten=blob, firsthit=fh, render=film,
                            frames=FRAMES, threads=BANDS)
            scene = abi.FlatScene.load(blob)
            arrays = {"scene_" + n: getattr(scene, n) for n, _ in abi._ARRAYS}
            cam = np.fromfile(blob + ".camera", dtype=np.uint8)
            f = binding.read_ref_film(film)
            np.savez_compressed(
                os.path.join(HERE, name + ".npz"),
                scene_root=np.int64(scene.scene_root), sky_material=np.int64(scene.sky_material),
                camera=cam, first_hit=binding.read_ref_first_hit(fh),
                film_mean=f["mean"], film_count=f["count"], film_rgb8=f["rgb8"],
                film_normal=f["normal"], film_depth=f["depth"], film_mat_id=f["mat_id"],
                meta=np.array([W, H, FRAMES, BANDS], np.int64),
                seeds=binding.ref_seed_schedule(FRAMES), **arrays)
            print(name, os.path.getsize(os.path.join(HERE, name + ".npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
