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
In tests/golden/paths_NAME.npz: per-path TraceResults (radiance, ray count, normal, depth,
material id) of the unmodified reference fed the CUDA path's random streams
(oracle/_ref/fs2_ref_ctr, oracle/ref/ctr_rng.cpp) for PATH_SAMPLES sample indices, plus one
sample with Camera::fast_render_enabled (engine.cpp:68-70).
In tests/golden/exr_textures.npz: small .exr files (HALF / FLOAT x NONE / RLE / ZIPS / ZIP,
written by scenes/exr_write.py) and the texel arrays the reference's own loader — tinyexr's
LoadEXR behind DiffuseMaterial::LoadDiffuseTexture, source/material.cpp:70-92 — made of them.
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
SCENES = ["api", "demo", "demo_dof", "many303", "mesh_tiny", "glass_small", "grid_tiny",
          "needles", "far_mesh", "far_camera", "wide_range"]
PATH_SAMPLES, PATH_SEED = 2, 20260105
PATH_DT = np.dtype([("rgb", "<f4", 3), ("normal", "<f4", 3), ("depth", "<f4"), ("ray_count", "<u4"),
                    ("mat_id", "<u8"), ("pad", "<u8")])
CTR_BIN = os.path.join(REPO, "oracle", "_ref", "fs2_ref_ctr")


def dump_paths(scene_path, tmp, name, samples, fast):
    """Per-path records of the reference with the device's random streams."""
    import subprocess
    blob = os.path.join(tmp, name + ".ctr.fs2")
    out = os.path.join(tmp, name + ".paths")
    cmd = [CTR_BIN, scene_path, "--width", str(W), "--height", str(H), "--flatten", blob,
           "--paths", out, "--samples", str(samples), "--seed", str(PATH_SEED)]
    if fast:
        cmd.append("--fast")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise SystemExit("fs2_ref_ctr failed: %s %s" % (r.stdout[-1000:], r.stderr[-1000:]))
    raw = np.fromfile(out, dtype=np.uint8)
    hdr = raw[:16].view(np.uint32)
    assert hdr[0] == 0x54503246 and (hdr[1], hdr[2], hdr[3]) == (W, H, samples)
    return raw[16:].view(PATH_DT).reshape(samples, W * H), blob


def exr_goldens(tmp):
    """Eight small .exr files through the reference's own loader (one scene, one textured
    material per file; the flattened scene's texel array is what LoadEXR produced)."""
    from exr_write import write_exr  # scenes/exr_write.py
    rng = np.random.default_rng(11)
    out = {}
    text = ""
    names = []
    for dtype, dname in ((np.float16, "half"), (np.float32, "float")):
        for comp in ("none", "rle", "zips", "zip"):
            name = "%s_%s" % (dname, comp)
            w, h = 20 + len(names), 9 + len(names)  # odd sizes: scan-line blocks of 16 with a tail
            img = (rng.random((h, w, 3)) * 4.0).astype(dtype)
            img[0, 0] = (0.0, 1.0, 65504.0 if dtype == np.float16 else 1.0e30)
            img[h // 2, : w // 2] = 0.25
            path = os.path.join(tmp, name + ".exr")
            write_exr(path, {"R": img[..., 0], "G": img[..., 1], "B": img[..., 2]}, comp=comp)
            out["file_" + name] = np.fromfile(path, np.uint8)
            text += gen_scenes.material("m_" + name, color=(1, 1, 1), texture_scale=1.0, texture=path)
            text += gen_scenes.block("sphere", material="m_" + name,
                                     position=(3.0 * len(names), 0.0, 0.0), radius=1.0)
            names.append(name)
    scene_path = os.path.join(tmp, "exr.scene")
    with open(scene_path, "w") as f:
        f.write(text)
    blob = os.path.join(tmp, "exr.fs2")
    binding.run_ref(scene_path, 8, 8, flatten=blob)
    scene = abi.FlatScene.load(blob)
    mats = {int(m["texture"]): i for i, m in enumerate(scene.materials) if int(m["texture"]) >= 0}
    assert len(scene.textures) == len(names), "the reference's loader rejected a file"
    for k, name in enumerate(names):  # textures are created in material order
        t = scene.textures[k]
        n = int(t["width"]) * int(t["height"]) * 3
        out["texels_" + name] = np.array(scene.texels[int(t["offset"]):int(t["offset"]) + n], np.float32)
        out["dims_" + name] = np.array([int(t["width"]), int(t["height"])], np.int64)
    np.savez_compressed(os.path.join(HERE, "exr_textures.npz"), names=np.array(names), **out)
    print("exr_textures", os.path.getsize(os.path.join(HERE, "exr_textures.npz")) // 1024, "KiB")


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
            # per-path records with the device's streams, from the same (unmodified) reference
            rec, ctr_blob = dump_paths(paths[name], tmp, name, PATH_SAMPLES, fast=False)
            assert open(ctr_blob, "rb").read() == open(blob, "rb").read(), \
                "fs2_ref_ctr flattened a different scene than fs2_ref"
            fast, _ = dump_paths(paths[name], tmp, name, 1, fast=True)
            np.savez_compressed(
                os.path.join(HERE, "paths_" + name + ".npz"),
                meta=np.array([W, H, PATH_SAMPLES, PATH_SEED], np.int64),
                rgb=rec["rgb"], ray_count=rec["ray_count"], normal=rec["normal"], depth=rec["depth"],
                mat_id=rec["mat_id"], fast_rgb=fast["rgb"], fast_ray_count=fast["ray_count"])
            print("paths_" + name, os.path.getsize(os.path.join(HERE, "paths_" + name + ".npz")) // 1024, "KiB")
        exr_goldens(tmp)


if __name__ == "__main__":
    main()
