"""ctypes view of libfs2host.so: the C++ host mirror of the reference's Scene / Camera /
DisplayFrame / TraceScene API (host/fs2_host.h).  Scene loading, octree construction and
flattening happen in that library; this module only moves pointers around."""
import ctypes as C
import os

import numpy as np

from . import abi, build

_lib = None


def load_host_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(build.HOST_LIB):
        build.build_host()
    lib = C.CDLL(build.HOST_LIB)
    lib.fs2h_scene_new.restype = C.c_void_p
    lib.fs2h_scene_free.argtypes = [C.c_void_p]
    lib.fs2h_scene_load.argtypes = [C.c_void_p, C.c_char_p]
    lib.fs2h_scene_flatten.restype = C.POINTER(abi.fs_scene_desc)
    lib.fs2h_scene_flatten.argtypes = [C.c_void_p]
    lib.fs2h_scene_save_blob.argtypes = [C.c_void_p, C.c_char_p]
    lib.fs2h_scene_camera.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.fs_camera)]
    lib.fs2h_frame_new.restype = C.c_void_p
    lib.fs2h_frame_new.argtypes = [C.c_uint32, C.c_uint32]
    lib.fs2h_frame_free.argtypes = [C.c_void_p]
    lib.fs2h_frame_reset.argtypes = [C.c_void_p]
    lib.fs2h_frame_display.restype = C.POINTER(C.c_uint8)
    lib.fs2h_frame_display.argtypes = [C.c_void_p]
    lib.fs2h_frame_mean.restype = C.POINTER(C.c_float)
    lib.fs2h_frame_mean.argtypes = [C.c_void_p]
    lib.fs2h_frame_count.restype = C.POINTER(C.c_uint32)
    lib.fs2h_frame_count.argtypes = [C.c_void_p]
    lib.fs2h_trace_samples.argtypes = [C.c_void_p, C.POINTER(abi.fs_camera), C.c_void_p,
                                       C.c_uint32]
    lib.fs2h_trace_range.restype = C.c_float
    lib.fs2h_trace_range.argtypes = [C.c_void_p, C.POINTER(abi.fs_camera), C.c_void_p,
                                     C.c_float, C.c_float]
    lib.fs2h_set_device.argtypes = [C.c_int]
    lib.fs2h_set_devices.argtypes = [C.POINTER(C.c_int), C.c_int]
    lib.fs2h_set_sharding.argtypes = [C.c_int]
    lib.fs2h_device_count.restype = C.c_int
    lib.fs2h_scene_add_fog_cuboid.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    lib.fs2h_set_seed.argtypes = [C.c_uint64]
    lib.fs2h_set_gpu_build.argtypes = [C.c_int]
    lib.fs2h_gpu_build_ms.restype = C.c_float
    lib.fs2h_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def set_trace_devices(devices, sharding=0):
    """base::SetTraceDevices: the GPUs behind TraceScene / TraceSamples; sharding 0 automatic,
    1 sample indices, 2 interleaved tiles."""
    lib = load_host_library()
    ids = (C.c_int * len(devices))(*devices)
    lib.fs2h_set_devices(ids, len(devices))
    lib.fs2h_set_sharding(sharding)


def _np_view(ptr, count, dtype):
    if not count:
        return np.zeros(0, dtype)
    buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count)


class HostScene:
    """A scene loaded and flattened by the C++ host library, in a fresh random stream (the
    ids and the normal table equal what a fresh reference process produces)."""

    def __init__(self, path=None, initialize_materials=True):
        self.lib = load_host_library()
        self.lib.fs2h_reset_rng()
        self.handle = C.c_void_p(self.lib.fs2h_scene_new())
        self._desc = None
        if path is not None:
            if self.lib.fs2h_scene_load(self.handle, os.fsencode(path)) != 0:
                raise IOError("cannot load scene %s" % path)
            if initialize_materials:
                self.lib.fs2h_initialize_materials()

    def close(self):
        if self.handle:
            self.lib.fs2h_scene_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def desc(self):
        """fs_scene_desc (ctypes) pointing into the host library's arrays."""
        if self._desc is None:
            self._desc = self.lib.fs2h_scene_flatten(self.handle).contents
        return self._desc

    def camera(self, index=0):
        cam = abi.fs_camera()
        self.lib.fs2h_scene_camera(self.handle, index, C.byref(cam))
        return cam

    def add_fog_cuboid(self, position, size, rotation, colour, density):
        """FogMaterial cuboid through the mirrored C++ API (not expressible in a .scene file)."""
        v = (C.c_float * 14)(*position, *size, *rotation, *colour, density)
        if self.lib.fs2h_scene_add_fog_cuboid(self.handle, v) != 0:
            raise RuntimeError("AddCuboidObject failed")
        self._desc = None

    def save_blob(self, path):
        if self.lib.fs2h_scene_save_blob(self.handle, os.fsencode(path)) != 0:
            raise IOError("cannot write %s" % path)

    def to_flat(self):
        """Copies the flattened arrays into a numpy-backed abi.FlatScene."""
        d = self.desc()
        counts = {"nodes": d.n_nodes, "scene_refs": d.n_scene_refs, "tri_plane": 4 * d.n_tris,
                  "tri_aux": d.n_tris, "faces": d.n_faces, "objects": d.n_objects,
                  "materials": d.n_materials, "textures": d.n_textures, "texels": d.n_texels,
                  "lut": 3 * d.n_lut}
        arrays = {}
        for name, dt in abi._ARRAYS:
            ptr = getattr(d, name)
            arrays[name] = _np_view(ptr, counts[name], dt).copy() if ptr else np.zeros(0, dt)
        return abi.FlatScene(scene_root=d.scene_root, sky_material=d.sky_material, **arrays)

    def nbytes(self):
        d = self.desc()
        return (d.n_nodes * 32 + d.n_scene_refs * 4 + d.n_tris * 64 + d.n_faces * 64 +
                d.n_objects * 144 + d.n_materials * 64 + d.n_textures * 16 + d.n_texels * 4 +
                d.n_lut * 12)


class HostFrame:
    """DisplayFrame of the host mirror; trace() is the TraceScene drop-in path."""

    def __init__(self, width, height):
        self.lib = load_host_library()
        self.width, self.height = width, height
        self.handle = C.c_void_p(self.lib.fs2h_frame_new(width, height))

    def close(self):
        if self.handle:
            self.lib.fs2h_frame_free(self.handle)
            self.handle = C.c_void_p()

    def reset(self):
        self.lib.fs2h_frame_reset(self.handle)

    def trace(self, scene, cam, n_samples=1):
        rc = self.lib.fs2h_trace_samples(scene.handle, C.byref(cam), self.handle, n_samples)
        if rc != 0:
            raise RuntimeError("TraceScene failed: %s" % self.lib.fs2h_last_error().decode())

    def trace_range(self, scene, cam, x, y):
        return self.lib.fs2h_trace_range(scene.handle, C.byref(cam), self.handle, x, y)

    def mean(self):
        n = self.width * self.height
        return _np_view(C.cast(self.lib.fs2h_frame_mean(self.handle), C.c_void_p).value, n * 3,
                        np.float32).reshape(n, 3)

    def count(self):
        n = self.width * self.height
        return _np_view(C.cast(self.lib.fs2h_frame_count(self.handle), C.c_void_p).value, n,
                        np.uint32)

    def display(self):
        n = self.width * self.height
        return _np_view(C.cast(self.lib.fs2h_frame_display(self.handle), C.c_void_p).value,
                        n * 3, np.uint8).reshape(n, 3)
