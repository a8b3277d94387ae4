"""Scene::Trace(const ray&, ObjectCollision*) and Scene::SampleSky (reference source/scene.h:128-131,
source/scene.cpp:157-162,227-249) on the device — fs_trace_rays / fs_sample_sky through the C ABI,
and Scene::Trace / SampleSky / DisplayFrame::WritePixel through the C++ mirror — against the CPU
oracle: every field bit for bit."""
import ctypes as C

import numpy as np
import pytest

import gen_scenes
from fs2_b200 import abi, host
from oracle import binding
from refdata import bits, golden

pytestmark = pytest.mark.gpu


def _segments(g, n, seed):
    """Random segments through the scene's neighbourhood: camera-like long rays, short rays that
    start near the geometry (some inside objects), and a few zero-length ones."""
    rng = np.random.default_rng(seed)
    o = np.array(list(g.cam.origin), np.float32)
    t = np.array(list(g.cam.target), np.float32)
    starts = np.empty((n, 3), np.float32)
    stops = np.empty((n, 3), np.float32)
    for i in range(n):
        k = i % 4
        if k == 0:      # from the camera into the scene
            d = (t - o) + rng.normal(size=3) * 14.0
            starts[i] = o
            stops[i] = o + d / np.linalg.norm(d) * 400.0
        elif k == 1:    # between two points near the target
            starts[i] = t + rng.normal(size=3) * 12.0
            stops[i] = t + rng.normal(size=3) * 12.0
        elif k == 2:    # upwards / outwards: mostly sky
            starts[i] = t + rng.normal(size=3) * 6.0
            stops[i] = starts[i] + np.abs(rng.normal(size=3)) * 900.0
        else:
            starts[i] = t + rng.normal(size=3) * 5.0
            stops[i] = starts[i] if i % 16 == 3 else starts[i] + rng.normal(size=3) * 0.5
    return starts.astype(np.float32), stops.astype(np.float32)


@pytest.mark.parametrize("name", ["demo", "mesh_tiny", "many303", "api", "glass_small", "grid_tiny"])
def test_trace_rays_equals_oracle(ctx, name):
    g = golden(name)
    ctx.upload(g.scene)
    starts, stops = _segments(g, 6000, 7)
    got = ctx.trace_rays(starts, stops)
    ref = binding.trace_rays(g.scene, starts, stops)
    hit = ref["object"] != 0xFFFFFFFF
    assert 0.15 < hit.mean() < 0.98  # the segments exercise both outcomes
    assert (got["object"] == ref["object"]).all()
    assert (bits(got["param"]) == bits(ref["param"])).all()
    assert (got["face"][hit] == ref["face"][hit]).all()
    assert (got["material"] == ref["material"]).all()
    assert (got["is_internal"] == ref["is_internal"]).all() and ref["is_internal"].any()
    for f in ("point", "normal"):
        assert (bits(got[f][hit]) == bits(ref[f][hit])).all(), f
    # texture coordinates of spheres go through atan2f (intersect.cpp:780-784): libdevice and
    # glibc round it differently in the last place, everything else is exact
    sphere = np.zeros(len(hit), bool)
    sphere[hit] = g.scene.objects["type"][ref["object"][hit]] == abi.FS_OBJ_SPHERE
    # a mesh without texture coordinates does not write the record's uv: the reference (and the
    # oracle) leave whatever an earlier, farther candidate stored there; the device reports (0, 0)
    # (stated deviation: such a mesh has nothing meaningful to sample a texture with)
    no_vt = np.zeros(len(hit), bool)
    ob = g.scene.objects[ref["object"][hit]]
    no_vt[hit] = (ob["type"] == abi.FS_OBJ_MESH) & ((ob["u"][:, 2] & 2) == 0)
    exact = hit & ~sphere & ~no_vt
    assert (bits(got["uv"][exact]) == bits(ref["uv"][exact])).all()
    assert np.allclose(got["uv"][sphere], ref["uv"][sphere], rtol=0, atol=2e-6)


@pytest.mark.parametrize("name", ["demo", "glass_small"])
def test_sample_sky_equals_oracle(ctx, name):
    g = golden(name)
    ctx.upload(g.scene)
    rng = np.random.default_rng(3)
    v = rng.normal(size=(4096, 3)).astype(np.float32)
    v /= np.linalg.norm(v, axis=1, keepdims=True).astype(np.float32)
    got = ctx.sample_sky(v)
    ref = binding.sample_sky(g.scene, v)
    assert (bits(got) == bits(ref)).all()
    assert got.max() > 0.0


def test_mirror_scene_trace_sample_sky_write_pixel(tmp_path):
    """The same calls through the C++ mirror (base::Scene::Trace / SampleSky,
    base::DisplayFrame::WritePixel followed by TraceSamples)."""
    lib = host.load_host_library()
    path = gen_scenes.generate(str(tmp_path), ["demo"])["demo"]
    hs = host.HostScene(path)
    lib.fs2h_scene_trace.restype = C.c_int
    lib.fs2h_scene_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.fs2h_scene_sample_sky.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.fs2h_frame_write_pixel.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    cam = hs.camera()
    start = np.array(list(cam.origin), np.float32)
    d = np.array(list(cam.target), np.float32) - start
    stop = (start + d / np.linalg.norm(d) * 500.0).astype(np.float32)
    out = np.zeros(11, np.float32)
    hit = lib.fs2h_scene_trace(hs.handle, start.ctypes.data, stop.ctypes.data, out.ctypes.data)
    ref = binding.trace_rays(hs, start[None], stop[None])
    assert hit == int(ref["object"][0] != 0xFFFFFFFF) == 1
    assert bits(out[0:1])[0] == bits(ref["param"])[0]
    assert (bits(out[1:4]) == bits(ref["point"][0])).all() and (bits(out[4:7]) == bits(ref["normal"][0])).all()
    mats = np.ctypeslib.as_array(C.cast(hs.desc().materials, C.POINTER(C.c_uint32)),
                                 shape=(hs.desc().n_materials, 16))
    assert int(out[9]) == int(np.float32(mats[ref["material"][0], 1]))  # Material::GetID of the struck surface
    up = np.array([0.1, 0.9, 0.2], np.float32)
    rgb = np.zeros(3, np.float32)
    lib.fs2h_scene_sample_sky(hs.handle, up.ctypes.data, rgb.ctypes.data)
    assert (bits(rgb) == bits(binding.sample_sky(hs, up[None])[0])).all()
    # WritePixel, then a traced sample on top: pixel (3, 2) holds (written + traced) / 2
    frame = host.HostFrame(32, 16)
    only = host.HostFrame(32, 16)
    px = np.array([0.25, 0.5, 0.75], np.float32)
    for y in range(16):
        for x in range(32):
            lib.fs2h_frame_write_pixel(frame.handle, px.ctypes.data, x, y)
    assert frame.count().min() == 1 and np.allclose(frame.mean().reshape(-1, 3), px)
    lib.fs2h_set_seed(11)
    frame.trace(hs, cam, 1)
    only.trace(hs, cam, 1)
    assert (frame.count() == 2).all()
    assert np.allclose(frame.mean(), (only.mean().reshape(-1, 3) + px) / 2.0, rtol=2e-6, atol=1e-7)
