"""CUDA first-hit path (fs_first_hit through the C ABI) against the golden vectors of the
unmodified reference and against the CPU oracle: bit-exact parameters, ranges and IDs."""
import numpy as np
import pytest

from oracle import binding
from refdata import GOLDEN_SCENES, bits, golden, local_face

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", GOLDEN_SCENES)
def test_first_hit_matches_reference_golden(ctx, name):
    g = golden(name)
    ctx.upload(g.scene)
    o = ctx.first_hit(g.cam, g.w, g.h)
    fh = g.first_hit
    hit = fh["hit"] == 1
    # stated tolerance: none — every pixel must agree bit for bit (0 grazing pixels allowed)
    assert int((o["object"] != fh["object"]).sum()) == 0
    assert int((bits(o["t"]) != bits(fh["t"])).sum()) == 0
    assert int((bits(o["range"]) != bits(fh["range"])).sum()) == 0
    assert (o["mat_id"] == fh["mat_id"]).all()
    lf = local_face(g.scene, o["object"], o["face"])
    assert (lf[hit] == fh["face"][hit]).all()


@pytest.mark.parametrize("name,w,h", [("mesh_tiny", 333, 187), ("many303", 257, 129),
                                      ("demo", 512, 512), ("grid_tiny", 640, 360)])
def test_first_hit_matches_oracle_other_sizes(ctx, name, w, h):
    g = golden(name)
    ctx.upload(g.scene)
    ctx.reset_stats()
    o = ctx.first_hit(g.cam, w, h)
    ref = binding.first_hit(g.scene, g.cam, w, h)
    assert (o["object"] == ref["object"]).all()
    assert (bits(o["t"]) == bits(ref["t"])).all()
    assert (o["face"] == ref["face"]).all()
    assert (bits(o["range"]) == bits(ref["range"])).all()
    assert (o["mat_id"] == ref["mat_id"]).all()
    # work counters of the stats build equal the oracle's, test for test
    st = ctx.stats()
    c = ref["counters"]
    assert st["n_segments"] == c["n_segments"] == w * h
    assert st["n_node"] == c["n_node"]
    assert st["n_tri"] == c["n_tri"]
    assert st["n_prim"] == c["n_prim"]


def test_trace_range_single_ray(ctx):
    g = golden("demo")
    ctx.upload(g.scene)
    ref = binding.first_hit(g.scene, g.cam, g.w, g.h)["range"].reshape(g.h, g.w)
    for (x, y) in [(0, 0), (17, 40), (95, 63), (48, 5)]:
        r = ctx.trace_range(g.cam, g.w, g.h, float(x), float(y))
        assert np.float32(r).view(np.uint32) == ref[y, x].view(np.uint32)


def test_error_paths(ctx):
    import fs2_b200
    c2 = fs2_b200.Context(0)
    g = golden("demo")
    with pytest.raises(fs2_b200.Fs2Error) as e:
        c2.first_hit(g.cam, 8, 8)
    assert e.value.code == -3  # FS_E_NOSCENE
    c2.upload(g.scene)
    with pytest.raises(fs2_b200.Fs2Error) as e:
        c2.trace(g.cam, 1)
    assert e.value.code == -4  # FS_E_NOFILM
    bad = golden("demo").scene
    d = bad.desc()
    d.sky_material = 999
    with pytest.raises(fs2_b200.Fs2Error) as e:
        c2.upload(d)
    assert e.value.code == -1
    c2.close()
