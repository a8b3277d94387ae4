"""fs2_render (headless replacement of the reference's main.cpp, source/main.cpp:47-168)
renders a scene file through the mirrored C++ API on the GPU and writes PPM / PFM."""
import os
import subprocess

import numpy as np
import pytest

import fs2_b200  # noqa: F401
import gen_scenes
from fs2_b200 import build

pytestmark = pytest.mark.gpu


def test_cli_renders_scene_file(tmp_path):
    exe = build.build_cli()
    scene = gen_scenes.generate(str(tmp_path), ["demo"])["demo"]
    ppm, pfm = str(tmp_path / "o.ppm"), str(tmp_path / "o.pfm")
    r = subprocess.run([exe, "-f", scene, "-w", "96", "-h", "64", "--spp", "32", "--seed", "3",
                        "--out", ppm, "--pfm", pfm, "--focus", "48,20"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "focal depth" in r.stdout
    raw = open(ppm, "rb").read()
    assert raw.startswith(b"P6\n96 64\n255\n") and len(raw) == 13 + 96 * 64 * 3
    img = np.frombuffer(raw[13:], np.uint8)
    assert img.mean() > 10 and img.std() > 5  # a lit, non-constant image
    mean = np.frombuffer(open(pfm, "rb").read()[len(b"PF\n96 64\n-1.0\n"):], np.float32)
    assert mean.size == 96 * 64 * 3 and np.isfinite(mean).all() and mean.mean() > 0.01


def test_cli_fails_loudly_on_missing_scene(tmp_path):
    exe = build.build_cli()
    r = subprocess.run([exe, "-f", str(tmp_path / "nope.scene")], capture_output=True, text=True)
    assert r.returncode != 0


def test_cli_checkpoint_resume_continues_the_sample_sequence(tmp_path):
    """3 spp + checkpoint, then resume + 5 spp in a new process == 8 spp in one go (same seed:
    sample k has the same random stream either way)."""
    exe = build.build_cli()
    scene = gen_scenes.generate(str(tmp_path), ["demo"])["demo"]
    common = [exe, "-f", scene, "-w", "80", "-h", "48", "--seed", "9"]
    ck = str(tmp_path / "film.ckpt")

    def run(extra, name):
        pfm = str(tmp_path / (name + ".pfm"))
        r = subprocess.run(common + extra + ["--out", str(tmp_path / (name + ".ppm")), "--pfm", pfm],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        return np.frombuffer(open(pfm, "rb").read()[len(b"PF\n80 48\n-1.0\n"):], np.float32)

    whole = run(["--spp", "8"], "whole")
    run(["--spp", "3", "--checkpoint", ck], "part")
    resumed = run(["--spp", "5", "--resume", ck], "resumed")
    assert np.allclose(whole, resumed, rtol=1e-5, atol=1e-6)
    bad = subprocess.run(common + ["-w", "64", "--spp", "1", "--resume", ck], capture_output=True, text=True)
    assert bad.returncode != 0  # a checkpoint of another frame size is refused
    # ... and so is one of another image: the header carries the seed and hashes of the flattened
    # scene and of the camera the sums were accumulated from
    other = gen_scenes.generate(str(tmp_path), ["demo_dof"])["demo_dof"]
    for argv in ([exe, "-f", other, "-w", "80", "-h", "48", "--seed", "9"],
                 [exe, "-f", scene, "-w", "80", "-h", "48", "--seed", "10"]):
        bad = subprocess.run(argv + ["--spp", "1", "--resume", ck, "--out", str(tmp_path / "bad.ppm")],
                             capture_output=True, text=True)
        assert bad.returncode != 0 and "another scene, camera or seed" in bad.stderr, bad.stderr


def test_cli_post_processing_output(tmp_path):
    """--post-out: the denoised image is smoother than the raw one at the same samples, bloom
    only adds light; the raw --out image is unchanged by the post chain."""
    exe = build.build_cli()
    scene = gen_scenes.generate(str(tmp_path), ["demo"])["demo"]
    common = [exe, "-f", scene, "-w", "96", "-h", "64", "--spp", "4", "--seed", "5"]

    def run(extra, name):
        out, post = str(tmp_path / (name + ".ppm")), str(tmp_path / (name + "_post.ppm"))
        r = subprocess.run(common + ["--out", out, "--post-out", post] + extra,
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        load = lambda f: np.frombuffer(open(f, "rb").read()[13:], np.uint8).reshape(64, 96, 3).astype(int)
        return load(out), load(post)

    raw, den = run(["--denoise", "4"], "den")
    rough = lambda a: np.abs(np.diff(a, axis=1)).mean()
    assert rough(den) < 0.8 * rough(raw)
    raw2, glow = run(["--bloom", "0.8,0.5,6", "--flare", "0.3,2"], "glow")
    assert (raw2 == raw).all()
    assert (glow >= raw2).all() and (glow > raw2).mean() > 0.05
    bad = subprocess.run(common + ["--post-out", str(tmp_path / "x.ppm"), "--denoise", "99"],
                         capture_output=True, text=True)
    assert bad.returncode != 0 and "post-processing failed" in bad.stderr


def test_cli_exr_output_equals_the_pfm(tmp_path):
    """--exr writes the same float mean image as --pfm (lossless, top row first in the file)."""
    import ctypes as C
    from fs2_b200 import host
    exe = build.build_cli()
    scene = gen_scenes.generate(str(tmp_path), ["demo"])["demo"]
    pfm, exr = str(tmp_path / "o.pfm"), str(tmp_path / "o.exr")
    r = subprocess.run([exe, "-f", scene, "-w", "72", "-h", "40", "--spp", "4", "--out", str(tmp_path / "o.ppm"),
                        "--pfm", pfm, "--exr", exr], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    mean = np.frombuffer(open(pfm, "rb").read()[len(b"PF\n72 40\n-1.0\n"):], np.float32).reshape(40, 72, 3)
    lib = host.load_host_library()
    lib.fs2h_load_image.restype = C.c_int
    lib.fs2h_load_image.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.c_void_p, C.c_size_t]
    w, h = C.c_uint32(), C.c_uint32()
    buf = np.zeros(72 * 40 * 3, np.float32)
    assert lib.fs2h_load_image(exr.encode(), C.byref(w), C.byref(h), buf.ctypes.data_as(C.c_void_p), buf.size) == 0
    assert (w.value, h.value) == (72, 40)
    assert np.array_equal(buf.reshape(40, 72, 3)[::-1], mean)


def test_cli_fast_preview_equals_the_abi_with_fast_render_enabled(tmp_path, ctx):
    """--fast sets Camera::fast_render_enabled (reference source/engine.cpp:68-70: paths stop after
    their second segment with radiance 1, what the reference renders while the mouse is down): the
    image equals fs_trace with that camera flag and the same seed, and differs from the full one.
    (The flag itself is pinned path by path against the reference in tests/test_gpu_paths.py.)"""
    from fs2_b200 import host
    exe = build.build_cli()
    scene = gen_scenes.generate(str(tmp_path), ["demo"])["demo"]
    common = [exe, "-f", scene, "-w", "96", "-h", "64", "--spp", "6", "--seed", "21"]

    def run(extra, name):
        pfm = str(tmp_path / (name + ".pfm"))
        r = subprocess.run(common + extra + ["--out", str(tmp_path / (name + ".ppm")), "--pfm", pfm],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        return np.frombuffer(open(pfm, "rb").read()[len(b"PF\n96 64\n-1.0\n"):], np.float32).reshape(64, 96, 3)

    fast, full = run(["--fast"], "fast"), run([], "full")
    hs = host.HostScene(scene)
    cam = hs.camera()
    cam.fast_render_enabled = 1
    ctx.upload(hs.desc())
    ctx.resize(96, 64)
    ctx.reset()
    ctx.trace(cam, 6, seed=21)
    want = ctx.read_film()["mean"].reshape(64, 96, 3)
    assert np.allclose(fast, want, rtol=2e-5, atol=1e-6)
    assert np.abs(fast - full).mean() > 1e-3
