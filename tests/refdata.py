"""Golden fixtures (tests/golden/*.npz, written by tests/golden/make_golden.py from the
unmodified reference renderer) and small comparison helpers."""
import ctypes as C
import os

import numpy as np

import fs2_b200  # noqa: F401
from fs2_b200 import abi

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_SCENES = ["api", "demo", "demo_dof", "many303", "mesh_tiny", "glass_small", "grid_tiny",
                 # adversarial meshes for the leaf BVH (scenes/gen_scenes.py): sliver triangles, a mesh
                 # at (9000, 9000, 9000), a camera 500 units from a unit-sized scene, coordinates
                 # spanning 1e-3 ... 5e3 in one mesh
                 "needles", "far_mesh", "far_camera", "wide_range"]


class Golden:
    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.name = name
        arrays = {n: z["scene_" + n] for n, _ in abi._ARRAYS}
        self.scene = abi.FlatScene(scene_root=int(z["scene_root"]),
                                   sky_material=int(z["sky_material"]), **arrays)
        self.cam = abi.fs_camera()
        raw = z["camera"].tobytes()
        C.memmove(C.addressof(self.cam), raw, len(raw))
        self.w, self.h, self.frames, self.bands = (int(v) for v in z["meta"])
        self.seeds = z["seeds"]
        self.first_hit = z["first_hit"].reshape(-1)
        self.film = {k[5:]: z[k] for k in z.files if k.startswith("film_")}


class GoldenPaths:
    """tests/golden/paths_NAME.npz: per-path TraceResults of the unmodified reference fed the
    CUDA path's random streams (oracle/_ref/fs2_ref_ctr), [samples, h*w] each."""

    def __init__(self, name):
        z = np.load(os.path.join(GOLDEN_DIR, "paths_" + name + ".npz"))
        self.w, self.h, self.samples, self.seed = (int(v) for v in z["meta"])
        for k in ("rgb", "ray_count", "normal", "depth", "mat_id", "fast_rgb", "fast_ray_count"):
            setattr(self, k, z[k])


_cache = {}


def golden(name):
    if name not in _cache:
        _cache[name] = Golden(name)
    return _cache[name]


def golden_paths(name):
    key = "paths_" + name
    if key not in _cache:
        _cache[key] = GoldenPaths(name)
    return _cache[key]


def ulp_distance(a, b):
    """Distance in units in the last place between two float32 arrays (same shape), treating
    +0 / -0 as equal; nan vs anything = huge."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    d = np.abs(ia - ib)
    d[np.isnan(a) | np.isnan(b)] = 1 << 40
    return d


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def local_face(scene, objects, faces):
    """Global face index (our convention) -> mesh-local index (the reference harness's)."""
    out = faces.copy()
    hit = objects != 0xFFFFFFFF
    is_mesh = np.zeros(len(objects), bool)
    is_mesh[hit] = scene.objects["type"][objects[hit]] == abi.FS_OBJ_MESH
    base = np.zeros(len(objects), np.uint32)
    base[is_mesh] = scene.objects["u"][objects[is_mesh], 1]
    out[is_mesh] = faces[is_mesh] - base[is_mesh]
    out[~is_mesh] = 0xFFFFFFFF
    return out


def rel_rmse(a, b):
    """Per-pixel relative RMSE of SURVEY.md §8(d): sqrt(mean((a-b)^2 / (b^2 + 0.01)))."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2 / (b ** 2 + 0.01))))
