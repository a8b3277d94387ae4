"""ctypes binding of oracle/libfs2oracle.so (the CPU restatement) and a subprocess
wrapper around oracle/_ref/fs2_ref (the reference compiled unmodified).
TEST INFRASTRUCTURE — see oracle/__init__.py."""
import ctypes as C
import json
import os
import subprocess

import numpy as np

import fs2_b200
from fs2_b200 import abi

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "libfs2oracle.so")
REF_BIN = os.path.join(HERE, "_ref", "fs2_ref")

FIRST_HIT_DT = np.dtype([("t", "<f4"), ("point", "<f4", 3), ("normal", "<f4", 3), ("uv", "<f4", 2),
                         ("mat_id", "<u4"), ("object", "<u4"), ("face", "<u4"), ("range", "<f4"),
                         ("hit", "<u4")])


def build_oracle(force=False):
    """g++ build of the restatement with the pinned numerics flags."""
    srcs = [os.path.join(HERE, n) for n in
            ("fs2_oracle.cpp", "fs2_oracle_trace.inc", "fs2_oracle_shade.inc")]
    srcs.append(os.path.join(REPO, "include", "fs2_cabi.h"))
    if (not force and os.path.exists(LIB_PATH) and
            all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return LIB_PATH
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-shared", "-fPIC",
           "-I" + os.path.join(REPO, "include"), "-o", LIB_PATH, srcs[0]]
    subprocess.run(cmd, check=True)
    return LIB_PATH


def build_ref():
    """Builds oracle/_ref/fs2_ref when the reference sources are present (this container);
    on the GPU box the prebuilt binary travels with the snapshot."""
    if os.path.isdir("/root/reference/source"):
        subprocess.run(["make", "-s", "-C", os.path.join(HERE, "ref"), "-j8"], check=True)
    return REF_BIN if os.path.exists(REF_BIN) else None


_lib = None


def lib():
    global _lib
    if _lib is None:
        build_oracle()
        _lib = C.CDLL(LIB_PATH)
        _lib.fs2o_inv_sqrt.restype = C.c_float
        _lib.fs2o_inv_sqrt.argtypes = [C.c_float]
        _lib.fs2o_sqrt.restype = C.c_float
        _lib.fs2o_sqrt.argtypes = [C.c_float]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def first_hit(scene, cam, w, h):
    """Oracle first hit for every pixel → dict of numpy arrays (row 0 = bottom)."""
    n = w * h
    out = {"t": np.zeros(n, np.float32), "object": np.zeros(n, np.uint32),
           "face": np.zeros(n, np.uint32), "mat_id": np.zeros(n, np.uint32),
           "range": np.zeros(n, np.float32), "point": np.zeros((n, 3), np.float32),
           "normal": np.zeros((n, 3), np.float32), "uv": np.zeros((n, 2), np.float32)}
    counters = np.zeros(6, np.uint64)
    d = scene.desc()
    rc = lib().fs2o_first_hit(C.byref(d), C.byref(cam), w, h, _p(out["t"]), _p(out["object"]),
                              _p(out["face"]), _p(out["mat_id"]), _p(out["range"]),
                              _p(out["point"]), _p(out["normal"]), _p(out["uv"]), _p(counters))
    if rc != 0:
        raise RuntimeError("fs2o_first_hit failed: %d" % rc)
    out["counters"] = dict(zip(["n_segments", "n_node", "n_tri", "n_bary", "n_prim", "n_shade"],
                               [int(c) for c in counters]))
    return out


def trace_rays(scene, starts, stops):
    """Oracle Scene::Trace for explicit segments -> dict of arrays (fs_trace_rays' fields)."""
    seg = np.ascontiguousarray(np.concatenate([np.asarray(starts, np.float32).reshape(-1, 3),
                                               np.asarray(stops, np.float32).reshape(-1, 3)], axis=1))
    n = len(seg)
    out = {"param": np.zeros(n, np.float32), "point": np.zeros((n, 3), np.float32),
           "normal": np.zeros((n, 3), np.float32), "uv": np.zeros((n, 2), np.float32),
           "object": np.zeros(n, np.uint32), "face": np.zeros(n, np.uint32),
           "material": np.zeros(n, np.uint32), "is_internal": np.zeros(n, np.uint32)}
    d = scene.desc()
    rc = lib().fs2o_trace_rays(C.byref(d), _p(seg), n, _p(out["param"]), _p(out["point"]),
                               _p(out["normal"]), _p(out["uv"]), _p(out["object"]), _p(out["face"]),
                               _p(out["material"]), _p(out["is_internal"]))
    if rc != 0:
        raise RuntimeError("fs2o_trace_rays failed: %d" % rc)
    return out


def sample_sky(scene, views):
    v = np.ascontiguousarray(np.asarray(views, np.float32).reshape(-1, 3))
    out = np.zeros((len(v), 3), np.float32)
    d = scene.desc()
    rc = lib().fs2o_sample_sky(C.byref(d), _p(v), len(v), _p(out))
    if rc != 0:
        raise RuntimeError("fs2o_sample_sky failed: %d" % rc)
    return out


def ref_seed_schedule(frames, clock0=1700000000000):
    """The scripted per-frame clock of oracle/ref/harness.cpp (render())."""
    return np.array([clock0 + 1000 * k + 7 * k * k for k in range(frames)], np.uint64)


def trace_scene(scene, cam, w, h, frames, bands, seeds=None, aux=True):
    """Replays `frames` reference TraceScene calls on the CPU restatement."""
    n = w * h
    if seeds is None:
        seeds = ref_seed_schedule(frames)
    seeds = np.ascontiguousarray(seeds, np.uint64)
    film = {"mean": np.zeros((n, 3), np.float32), "count": np.zeros(n, np.uint32),
            "rgb8": np.zeros((n, 3), np.uint8)}
    if aux:
        film.update(normal=np.zeros((n, 3), np.float32), depth=np.zeros(n, np.float32),
                    mat_id=np.zeros(n, np.uint64))
    rays = C.c_uint64(0)
    counters = np.zeros(6, np.uint64)
    d = scene.desc()
    rc = lib().fs2o_trace_scene(C.byref(d), C.byref(cam), w, h, frames, bands, _p(seeds),
                                _p(film["mean"]), _p(film["count"]), _p(film["rgb8"]),
                                _p(film.get("normal")), _p(film.get("depth")),
                                _p(film.get("mat_id")), C.byref(rays), _p(counters))
    if rc != 0:
        raise RuntimeError("fs2o_trace_scene failed: %d" % rc)
    film["rays"] = int(rays.value)
    film["counters"] = dict(zip(["n_segments", "n_node", "n_tri", "n_bary", "n_prim", "n_shade"],
                                [int(c) for c in counters]))
    return film


FOG_SHARED_STREAM = 1 << 30


def trace_paths(scene, cam, w, h, n_samples, first_sample=0, stride=1, seed=1, flags=0):
    """Per-path records with the device's random streams (oracle RngMode kDevice): arrays of
    shape [n_samples, h*w(, 3)] — rgb, ray_count, normal, depth, mat_id — plus counters."""
    n = w * h
    out = {"rgb": np.zeros((n_samples, n, 3), np.float32),
           "ray_count": np.zeros((n_samples, n), np.uint32),
           "normal": np.zeros((n_samples, n, 3), np.float32),
           "depth": np.zeros((n_samples, n), np.float32),
           "mat_id": np.zeros((n_samples, n), np.uint64)}
    counters = np.zeros(6, np.uint64)
    d = scene.desc()
    rc = lib().fs2o_trace_paths(C.byref(d), C.byref(cam), w, h, C.c_uint32(first_sample),
                                C.c_uint32(n_samples), C.c_uint32(stride), C.c_uint64(seed),
                                C.c_int(flags), _p(out["rgb"]), _p(out["ray_count"]),
                                _p(out["normal"]), _p(out["depth"]), _p(out["mat_id"]),
                                _p(counters))
    if rc != 0:
        raise RuntimeError("fs2o_trace_paths failed: %d" % rc)
    out["counters"] = dict(zip(["n_segments", "n_node", "n_tri", "n_bary", "n_prim", "n_shade"],
                               [int(c) for c in counters]))
    return out


# ------------------------------------------------------------------ compiled reference
def have_ref():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def run_ref(scene_path, width, height, flatten=None, firsthit=None, render=None, frames=1,
            threads=1, realclock=False, aperture=None, focal=None, time_only=False, cores=None,
            timeout=3600, fog=None):
    """Runs oracle/_ref/fs2_ref; returns the parsed FS2REF json summary (if rendering)."""
    cmd = [REF_BIN, scene_path, "--width", str(width), "--height", str(height)]
    if flatten:
        cmd += ["--flatten", flatten]
    if firsthit:
        cmd += ["--firsthit", firsthit]
    if render:
        cmd += ["--render", render]
    elif time_only:
        cmd += ["--time"]
    cmd += ["--frames", str(frames), "--threads", str(threads)]
    if realclock:
        cmd += ["--realclock"]
    if aperture is not None:
        cmd += ["--aperture", str(aperture)]
    if focal is not None:
        cmd += ["--focal", str(focal)]
    if fog is not None:  # 14 floats: a FogMaterial cuboid added through the reference's C++ API
        cmd += ["--fog", ",".join("%.9g" % v for v in fog)]
    if cores:
        cmd = ["taskset", "-c", "0-%d" % (cores - 1)] + cmd
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    summary = None
    for line in r.stdout.splitlines():
        if line.startswith("FS2REF {"):
            summary = json.loads(line[len("FS2REF "):])
    if r.returncode != 0:
        raise RuntimeError("fs2_ref failed (%d): %s\n%s" % (r.returncode, r.stdout[-2000:],
                                                            r.stderr[-2000:]))
    return summary


def read_ref_first_hit(path):
    raw = np.fromfile(path, dtype=np.uint8)
    hdr = raw[:16].view(np.uint32)
    assert hdr[0] == 0x48463246 and hdr[3] == FIRST_HIT_DT.itemsize
    w, h = int(hdr[1]), int(hdr[2])
    return raw[16:].view(FIRST_HIT_DT).reshape(h, w)


def read_ref_film(path):
    raw = np.fromfile(path, dtype=np.uint8)
    hdr = raw[:16].view(np.uint32)
    assert hdr[0] == 0x4D463246
    w, h, frames = int(hdr[1]), int(hdr[2]), int(hdr[3])
    n = w * h
    off = 16
    out = {"w": w, "h": h, "frames": frames}
    for name, dt, k in (("mean", np.float32, 3), ("count", np.uint32, 1), ("rgb8", np.uint8, 3),
                        ("normal", np.float32, 3), ("depth", np.float32, 1),
                        ("mat_id", np.uint64, 1)):
        nb = n * k * np.dtype(dt).itemsize
        a = raw[off:off + nb].view(dt)
        out[name] = a.reshape(n, k) if k > 1 else a
        off += nb
    return out
