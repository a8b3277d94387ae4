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
