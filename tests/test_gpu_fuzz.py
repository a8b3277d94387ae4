"""Seeded random scenes (scenes/gen_scenes.py scene_random): every primitive kind and material,
cameras inside and outside the object cloud, small meshes at random transforms, with and
without the scene octree.  GPU first hits must equal the CPU oracle's bit for bit and all
kernel variants must agree on the same seeded samples."""
import numpy as np
import pytest

import gen_scenes
from fs2_b200 import host
from oracle import binding
from test_gpu_image import _variants_agree

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", range(12))
def test_random_scene_first_hit_and_variants(ctx, seed, tmp_path_factory):
    d = str(tmp_path_factory.mktemp("fuzz"))
    name = "random%d" % seed
    hs = host.HostScene(gen_scenes.generate(d, [name])[name])
    cam = hs.camera()
    w, h = 112, 72
    ctx.upload(hs.desc())
    got = ctx.first_hit(cam, w, h)
    ref = binding.first_hit(hs, cam, w, h)
    assert (got["object"] == ref["object"]).all()
    assert (got["face"] == ref["face"]).all()
    assert (got["mat_id"] == ref["mat_id"]).all()
    assert (got["t"].view(np.uint32) == ref["t"].view(np.uint32)).all()
    assert (got["range"].view(np.uint32) == ref["range"].view(np.uint32)).all()
    assert (got["object"] != 0xFFFFFFFF).mean() > 0.1
    _variants_agree(ctx, hs.desc(), cam, 64, 40, 3)
