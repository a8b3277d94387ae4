"""fs_trace_rays_static — the allocation-free variant of fs_trace_rays (include/fs2_cabi.h; no
counterpart in the reference): its staged segments and
hit records live in a fixed array inside the library's device image instead of a cudaMalloc'd
buffer.  Through the C ABI it must return what fs_trace_rays returns, bit for bit (same kernel,
same arithmetic: the tolerance is zero), on two calls in a row that share the array, and refuse
batches beyond FS_TRACE_RAYS_STATIC_MAX with a message that names the limit.

This file is named to be collected last: it owns its context and
passes alone in a fresh process as well as after the other tests."""
import ctypes as C

import numpy as np
import pytest

import fs2_b200
from fs2_b200 import abi
from refdata import golden

pytestmark = pytest.mark.gpu


def _segments(g, n, seed):
    rng = np.random.default_rng(seed)
    o = np.array(list(g.cam.origin), np.float32)
    t = np.array(list(g.cam.target), np.float32)
    starts = np.empty((n, 3), np.float32)
    stops = np.empty((n, 3), np.float32)
    half = n // 2
    d = (t - o)[None, :] + rng.normal(size=(half, 3)) * 12.0   # from the camera into the scene
    starts[:half] = o
    stops[:half] = o + d / np.linalg.norm(d, axis=1, keepdims=True) * 400.0
    starts[half:] = t + rng.normal(size=(n - half, 3)) * 10.0  # short segments near the geometry
    stops[half:] = t + rng.normal(size=(n - half, 3)) * 10.0
    return starts.astype(np.float32), stops.astype(np.float32)


def _same_records(a, b):
    assert a.dtype == b.dtype == abi.RAY_HIT_DT and len(a) == len(b)
    assert a.tobytes() == b.tobytes()  # every field of every record, bit for bit


def test_static_scratch_variant_equals_fs_trace_rays_twice_in_a_row():
    g = golden("mesh_tiny")
    with fs2_b200.Context(device=0) as ctx:
        ctx.upload(g.scene)
        # first call: a large batch fills most of the shared array
        s1, e1 = _segments(g, 20000, 11)
        want1 = ctx.trace_rays(s1, e1)
        got1 = ctx.trace_rays_static(s1, e1)
        _same_records(got1, want1)
        hit = want1["object"] != 0xFFFFFFFF
        assert 0.1 < hit.mean() < 0.99  # both outcomes occur
        # second call, right after: other, fewer rays — the array still holds the first call's
        # segments and records beyond them, none of which may show up
        s2, e2 = _segments(g, 777, 12)
        got2 = ctx.trace_rays_static(s2, e2)
        _same_records(got2, ctx.trace_rays(s2, e2))
        # and the first batch again gives the first answer again
        _same_records(ctx.trace_rays_static(s1, e1), want1)
        # the largest batch the array takes
        s3, e3 = _segments(g, abi.FS_TRACE_RAYS_STATIC_MAX, 13)
        _same_records(ctx.trace_rays_static(s3, e3), ctx.trace_rays(s3, e3))


def test_static_scratch_variant_refuses_what_does_not_fit():
    g = golden("mesh_tiny")
    with fs2_b200.Context(device=0) as ctx:
        n = abi.FS_TRACE_RAYS_STATIC_MAX + 1
        seg = np.zeros((n, 6), np.float32)
        out = np.zeros(n, abi.RAY_HIT_DT)
        # before a scene is resident: the size check comes first, nothing is touched
        rc = ctx.lib.fs_trace_rays_static(ctx.handle, seg.ctypes.data_as(C.c_void_p), n,
                                          out.ctypes.data_as(C.c_void_p))
        assert rc == -1  # FS_E_INVALID
        msg = ctx.lib.fs_last_error(ctx.handle).decode()
        assert str(abi.FS_TRACE_RAYS_STATIC_MAX) in msg and "FS_TRACE_RAYS_STATIC_MAX" in msg
        assert not out.tobytes().strip(b"\0")
        rc = ctx.lib.fs_trace_rays_static(ctx.handle, seg.ctypes.data_as(C.c_void_p), 4,
                                          out.ctypes.data_as(C.c_void_p))
        assert rc == -3  # FS_E_NOSCENE
        ctx.upload(g.scene)
        with pytest.raises(fs2_b200.Fs2Error) as e:
            ctx.trace_rays_static(seg[:, :3], seg[:, 3:])
        assert e.value.code == -1
        assert len(ctx.trace_rays_static(seg[:0, :3], seg[:0, 3:])) == 0  # empty batch: a no-op
