"""The CPU restatement (oracle/) against the golden vectors produced by the unmodified
reference renderer: bit-exact first hits and bit-exact films (the reference's xorshift64*
streams are replayed), plus the survey's known-answer vectors."""
import ctypes as C

import numpy as np
import pytest

from oracle import binding
from refdata import GOLDEN_SCENES, bits, golden, local_face


@pytest.mark.parametrize("name", GOLDEN_SCENES)
def test_first_hit_bit_exact(name):
    g = golden(name)
    o = binding.first_hit(g.scene, g.cam, g.w, g.h)
    fh = g.first_hit
    hit = fh["hit"] == 1
    assert 0.05 < hit.mean() < 0.999  # the fixture exercises both hits and misses
    assert (bits(o["t"]) == bits(fh["t"])).all()
    assert (o["object"] == fh["object"]).all()
    assert (o["mat_id"] == fh["mat_id"]).all()
    assert (bits(o["range"]) == bits(fh["range"])).all()
    assert (bits(o["point"][hit]) == bits(fh["point"][hit])).all()
    assert (bits(o["normal"][hit]) == bits(fh["normal"][hit])).all()
    assert (bits(o["uv"][hit]) == bits(fh["uv"][hit])).all()
    lf = local_face(g.scene, o["object"], o["face"])
    assert (lf[hit] == fh["face"][hit]).all()


@pytest.mark.parametrize("name", GOLDEN_SCENES)
def test_film_bit_exact(name):
    g = golden(name)
    f = binding.trace_scene(g.scene, g.cam, g.w, g.h, g.frames, g.bands, seeds=g.seeds)
    assert (bits(f["mean"]) == bits(g.film["mean"])).all()
    assert (f["count"] == g.film["count"]).all()
    assert (f["rgb8"] == g.film["rgb8"]).all()
    assert (bits(f["normal"]) == bits(g.film["normal"])).all()
    assert (bits(f["depth"]) == bits(g.film["depth"])).all()
    assert (f["mat_id"] == g.film["mat_id"]).all()


def test_golden_covers_every_material_and_object_type():
    mats, objs = set(), set()
    for name in GOLDEN_SCENES:
        g = golden(name)
        mats |= set(int(t) for t in g.scene.materials["type"])
        objs |= set(int(t) for t in g.scene.objects["type"])
    assert mats == set(range(9))
    assert objs == set(range(6))
    assert golden("many303").scene.scene_root >= 0  # scene octree active
    assert golden("demo_dof").cam.aperture_size > 0


# ---- SURVEY.md App. C known answers (unmodified reference, g++ -O2 -ffp-contract=off)
def test_kat_rng():
    ints = np.zeros(3, np.uint64)
    floats = np.zeros(3, np.float32)
    binding.lib().fs2o_rng(C.c_uint64(521288629), 3, ints.ctypes.data_as(C.c_void_p),
                           floats.ctypes.data_as(C.c_void_p))
    assert [int(v) for v in ints] == [17856780258404322015, 14185171587269498017,
                                      264782984789228142]
    assert np.allclose(floats, [0.172053456, 0.0850002095, 0.611553192], rtol=0, atol=1e-9)


def test_kat_fast_sqrt():
    lib = binding.lib()
    assert np.float32(lib.fs2o_inv_sqrt(2.0)).view(np.uint32) == 0x3F34F95E
    assert np.float32(lib.fs2o_sqrt(2.0)).view(np.uint32) == 0x3FB51089


def test_kat_vectors():
    out = np.zeros(18, np.float32)
    v = np.array([0.3, -0.8, 0.52], np.float32)
    n = np.array([0, 1, 0], np.float32)
    binding.lib().fs2o_vec_kat(v.ctypes.data_as(C.c_void_p), n.ctypes.data_as(C.c_void_p),
                               C.c_float(0.75), C.c_float(0.7), out.ctypes.data_as(C.c_void_p))
    assert [hex(x) for x in out[0:3].view(np.uint32)] == ["0x3e994f8c", "0xbf4c6a10", "0x3f04de8a"]
    assert np.allclose(out[3:6], [0.299435019, 0.798493385, 0.519020677], atol=1e-8)
    assert [hex(x) for x in out[6:9].view(np.uint32)] == ["0x3e65bde1", "0xbf6443af", "0x3ec71c06"]
    assert np.allclose(out[9:12], [0.563382864, -0.798493385, 0.204067603], atol=2e-7)
    assert [hex(x) for x in out[12:14].view(np.uint32)] == ["0x3f1551fe", "0x3f663508"]
    # |(3,4,0)| with the approximate sqrt
    v2 = np.array([3, 4, 0], np.float32)
    binding.lib().fs2o_vec_kat(v2.ctypes.data_as(C.c_void_p), n.ctypes.data_as(C.c_void_p),
                               C.c_float(1.5), C.c_float(0.0), out.ctypes.data_as(C.c_void_p))
    assert abs(float(out[14]) - 5.00776815) < 1e-6
    # total internal reflection: normalize(0.9,-0.1,0.42).refract(+y, 1.5) = 0
    v3 = np.array([0.9, -0.1, 0.42], np.float32)
    binding.lib().fs2o_vec_kat(v3.ctypes.data_as(C.c_void_p), n.ctypes.data_as(C.c_void_p),
                               C.c_float(1.5), C.c_float(0.0), out.ctypes.data_as(C.c_void_p))
    assert (out[6:9] == 0).all()


def test_kat_film_quantisation():
    mean = np.zeros(3, np.float32)
    count = np.zeros(1, np.uint32)
    rgb8 = np.zeros(3, np.uint8)
    for c in ([0.2, 0.5, 1.7], [0.4, 0.1, 0.0]):
        rgb = np.array(c, np.float32)
        rc = binding.lib().fs2o_write_pixel(1, 1, mean.ctypes.data_as(C.c_void_p),
                                            count.ctypes.data_as(C.c_void_p),
                                            rgb8.ctypes.data_as(C.c_void_p), 0, 0,
                                            rgb.ctypes.data_as(C.c_void_p))
        assert rc == 0
    assert list(rgb8) == [148, 148, 237] and int(count[0]) == 2
