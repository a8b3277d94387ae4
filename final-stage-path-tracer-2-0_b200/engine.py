"""ctypes plumbing over libfs2cuda.so (include/fs2_cabi.h).  No tracing arithmetic here and
no fallback: if the CUDA library is missing or no device is present, calls raise."""
import ctypes as C
import os

import numpy as np

from . import abi, build

_SYMBOLS = {
    "fs_abi_version": (C.c_int, []),
    "fs_film_save": (C.c_int, [C.c_void_p, C.c_void_p]),
    "fs_film_load": (C.c_int, [C.c_void_p, C.c_void_p]),
    "fs_film_post": (C.c_int, [C.c_void_p, C.POINTER(abi.fs_post_params), C.c_void_p, C.c_void_p]),
    "fs_device_selftest": (C.c_int, [C.c_int, C.c_void_p, C.c_uint32]),
    "fs_host_register": (C.c_int, [C.c_void_p, C.c_size_t]),
    "fs_host_unregister": (C.c_int, [C.c_void_p]),
    "fs_mesh_tree_build": (C.c_int, [C.c_int, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p,
                                     C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                     C.POINTER(abi.fs_mesh_tree)]),
    "fs_mesh_tree_free": (None, [C.POINTER(abi.fs_mesh_tree)]),
    "fs_last_error": (C.c_char_p, [C.c_void_p]),
    "fs_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int]),
    "fs_create_multi": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int, C.c_int]),
    "fs_device_count": (C.c_int, [C.c_void_p]),
    "fs_set_tile_shard": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "fs_film_device_aux": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                     C.POINTER(C.c_size_t)]),
    "fs_destroy": (C.c_int, [C.c_void_p]),
    "fs_scene_upload": (C.c_int, [C.c_void_p, C.POINTER(abi.fs_scene_desc)]),
    "fs_film_resize": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "fs_film_reset": (C.c_int, [C.c_void_p]),
    "fs_first_hit": (C.c_int, [C.c_void_p, C.POINTER(abi.fs_camera), C.c_int, C.c_int,
                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fs_trace_range": (C.c_int, [C.c_void_p, C.POINTER(abi.fs_camera), C.c_int, C.c_int,
                                 C.c_float, C.c_float, C.POINTER(C.c_float)]),
    "fs_trace_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "fs_trace_rays_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "fs_sample_sky": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "fs_trace": (C.c_int, [C.c_void_p, C.POINTER(abi.fs_camera), C.c_uint32, C.c_uint32,
                           C.c_uint32, C.c_uint64, C.c_int]),
    "fs_film_read": (C.c_int, [C.c_void_p] + [C.c_void_p] * 6),
    "fs_film_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "fs_film_resolve": (C.c_int, [C.c_void_p]),
    "fs_get_stats": (C.c_int, [C.c_void_p, C.POINTER(abi.fs_stats)]),
    "fs_reset_stats": (C.c_int, [C.c_void_p]),
    "fs_get_stream": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "fs_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
}

_lib = None


class Fs2Error(RuntimeError):
    def __init__(self, code, message):
        super().__init__("fs2 error %d: %s" % (code, message))
        self.code = code


def load_library(path=None):
    """Loads libfs2cuda.so (building it in-tree if the sources are newer) and declares
    every entry point of include/fs2_cabi.h.  Raises if the library cannot be loaded."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    if path is None:
        path = os.environ.get("FS2_CUDA_LIB")  # experiments: an alternative build of the same ABI
    if path is None:
        path = build.CUDA_LIB
        if not os.path.exists(path):
            build.build_cuda()
    lib = C.CDLL(path)
    for name, (res, args) in _SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def exported_symbols():
    return sorted(_SYMBOLS)


def pin_host(array):
    """Page-locks a numpy array in place (fs_host_register) so film read-backs into it run at
    full PCIe speed; returns the array.  Call unpin_host before the array is freed."""
    lib = load_library()
    rc = lib.fs_host_register(array.ctypes.data_as(C.c_void_p), array.nbytes)
    if rc != 0:
        raise RuntimeError("fs_host_register: %s" % lib.fs_last_error(None).decode())
    return array


def unpin_host(array):
    load_library().fs_host_unregister(array.ctypes.data_as(C.c_void_p))


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Context:
    """One fs_ctx: a scene resident on one GPU plus its film — or, with `devices=[...]`, one
    multi-device context (fs_create_multi): the scene replicated on every listed GPU, fs_trace
    sharded across them, the film combined on devices[0]."""

    def __init__(self, device=0, stats=False, devices=None):
        self.lib = load_library()
        self.handle = C.c_void_p()
        if devices is not None:
            ids = (C.c_int * len(devices))(*devices)
            rc = self.lib.fs_create_multi(C.byref(self.handle), ids, len(devices), 1 if stats else 0)
            device = devices[0] if len(devices) else 0
        else:
            rc = self.lib.fs_create(C.byref(self.handle), device, 1 if stats else 0)
        if rc != 0:
            raise Fs2Error(rc, (self.lib.fs_last_error(None) or b"").decode())
        self.device = device
        self.devices = list(devices) if devices is not None else [device]
        self.width = self.height = 0

    def _check(self, rc):
        if rc != 0:
            raise Fs2Error(rc, (self.lib.fs_last_error(self.handle) or b"").decode())

    def close(self):
        if self.handle:
            self.lib.fs_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- scene / film ----------------------------------------------------------
    def upload(self, scene):
        d = scene.desc() if isinstance(scene, abi.FlatScene) else scene
        self._check(self.lib.fs_scene_upload(self.handle, C.byref(d)))

    def resize(self, width, height):
        self._check(self.lib.fs_film_resize(self.handle, width, height))
        self.width, self.height = width, height

    def reset(self):
        self._check(self.lib.fs_film_reset(self.handle))

    # ---- queries ---------------------------------------------------------------
    def first_hit(self, cam, width, height):
        n = width * height
        out = {"t": np.empty(n, np.float32), "object": np.empty(n, np.uint32),
               "face": np.empty(n, np.uint32), "mat_id": np.empty(n, np.uint32),
               "range": np.empty(n, np.float32)}
        self._check(self.lib.fs_first_hit(self.handle, C.byref(cam), width, height,
                                          _ptr(out["t"]), _ptr(out["object"]), _ptr(out["face"]),
                                          _ptr(out["mat_id"]), _ptr(out["range"])))
        return out

    def trace_range(self, cam, width, height, x, y):
        r = C.c_float()
        self._check(self.lib.fs_trace_range(self.handle, C.byref(cam), width, height, x, y,
                                            C.byref(r)))
        return r.value

    def trace_rays(self, starts, stops):
        """Scene::Trace for a batch of segments (fs_trace_rays): a structured array with
        param / point / normal / uv / object / face / material / is_internal per ray."""
        seg = np.ascontiguousarray(np.concatenate([np.asarray(starts, np.float32).reshape(-1, 3),
                                                   np.asarray(stops, np.float32).reshape(-1, 3)], axis=1))
        out = np.zeros(len(seg), abi.RAY_HIT_DT)
        self._check(self.lib.fs_trace_rays(self.handle, _ptr(seg), len(seg), _ptr(out)))
        return out

    def trace_rays_static(self, starts, stops):
        """fs_trace_rays_static: trace_rays without any device allocation (fixed scratch in the
        library image, at most abi.FS_TRACE_RAYS_STATIC_MAX rays per call)."""
        seg = np.ascontiguousarray(np.concatenate([np.asarray(starts, np.float32).reshape(-1, 3),
                                                   np.asarray(stops, np.float32).reshape(-1, 3)], axis=1))
        out = np.zeros(len(seg), abi.RAY_HIT_DT)
        self._check(self.lib.fs_trace_rays_static(self.handle, _ptr(seg), len(seg), _ptr(out)))
        return out

    def sample_sky(self, views):
        """Scene::SampleSky for a batch of view directions (fs_sample_sky): [n, 3] radiance."""
        v = np.ascontiguousarray(np.asarray(views, np.float32).reshape(-1, 3))
        out = np.zeros((len(v), 3), np.float32)
        self._check(self.lib.fs_sample_sky(self.handle, _ptr(v), len(v), _ptr(out)))
        return out

    def trace(self, cam, n_samples, first_sample=0, stride=1, seed=1, flags=0):
        self._check(self.lib.fs_trace(self.handle, C.byref(cam), first_sample, n_samples, stride,
                                      seed, flags))

    def resolve(self):
        self._check(self.lib.fs_film_resolve(self.handle))

    def read_film(self, mean=True, count=True, rgb8=True, aux=False, out=None):
        n = self.width * self.height
        film = out if out is not None else {}
        if mean and "mean" not in film:
            film["mean"] = np.empty((n, 3), np.float32)
        if count and "count" not in film:
            film["count"] = np.empty(n, np.uint32)
        if rgb8 and "rgb8" not in film:
            film["rgb8"] = np.empty((n, 3), np.uint8)
        if aux:
            film.setdefault("normal", np.empty((n, 3), np.float32))
            film.setdefault("depth", np.empty(n, np.float32))
            film.setdefault("mat_id", np.empty(n, np.uint64))
        self._check(self.lib.fs_film_read(self.handle, _ptr(film.get("mean")),
                                          _ptr(film.get("count")), _ptr(film.get("rgb8")),
                                          _ptr(film.get("normal")), _ptr(film.get("depth")),
                                          _ptr(film.get("mat_id"))))
        return film

    def post_film(self, params, rgb=True, rgb8=True):
        """Denoise / bloom / flare of the resolved film on the device (fs_film_post);
        params = abi.post_params(...).  The film itself is left as it is."""
        n = self.width * self.height
        out = {}
        if rgb:
            out["rgb"] = np.empty((n, 3), np.float32)
        if rgb8:
            out["rgb8"] = np.empty((n, 3), np.uint8)
        self._check(self.lib.fs_film_post(self.handle, C.byref(params), _ptr(out.get("rgb")),
                                          _ptr(out.get("rgb8"))))
        return out

    def save_film(self):
        """Accumulation film (rgb sum, count per pixel) as a numpy array: a checkpoint."""
        a = np.empty((self.width * self.height, 4), np.float32)
        self._check(self.lib.fs_film_save(self.handle, _ptr(a)))
        return a

    def load_film(self, sums):
        a = np.ascontiguousarray(sums, np.float32)
        assert a.size == self.width * self.height * 4
        self._check(self.lib.fs_film_load(self.handle, _ptr(a)))

    def film_device(self):
        """(device pointer, float count) of the rgb-sum/count film for an external reduce."""
        p = C.c_void_p()
        n = C.c_size_t()
        self._check(self.lib.fs_film_device(self.handle, C.byref(p), C.byref(n)))
        return p.value, n.value

    def film_device_aux(self):
        """(aux pointer, material-id pointer, pixel count): float4 normal+depth and u32 id planes."""
        a, m = C.c_void_p(), C.c_void_p()
        n = C.c_size_t()
        self._check(self.lib.fs_film_device_aux(self.handle, C.byref(a), C.byref(m), C.byref(n)))
        return a.value, m.value, n.value

    def set_tile_shard(self, rank, world):
        """Multi-process tile sharding: render only the tiles (tx + ty) % world == rank."""
        self._check(self.lib.fs_set_tile_shard(self.handle, rank, world))

    def stream(self):
        p = C.c_void_p()
        self._check(self.lib.fs_get_stream(self.handle, C.byref(p)))
        return p.value or 0

    def set_stream(self, stream):
        """Enqueue on a caller-owned CUDA stream (integer handle); None restores the own one."""
        self._check(self.lib.fs_set_stream(self.handle, C.c_void_p(stream) if stream else None))

    def stats(self):
        s = abi.fs_stats()
        self._check(self.lib.fs_get_stats(self.handle, C.byref(s)))
        return {k: getattr(s, k) for k, _ in abi.fs_stats._fields_}

    def reset_stats(self):
        self._check(self.lib.fs_reset_stats(self.handle))
