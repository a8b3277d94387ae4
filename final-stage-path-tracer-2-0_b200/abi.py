"""ctypes mirror of include/fs2_cabi.h and a numpy reader/writer for .fs2 blobs
(include/fs2_blob.h).  Plumbing only: no tracing arithmetic lives here."""
import ctypes as C
import os

import numpy as np

FS2_ABI_VERSION = 1
FS_NODE_INTERNAL = 0xFFFFFFFF
FS_NO_HIT_PARAM = 2.0
FS_LUT_SIZE = 32768

FS_OBJ_SPHERE, FS_OBJ_PLANE, FS_OBJ_DISC, FS_OBJ_CUBOID, FS_OBJ_QUAD, FS_OBJ_MESH = range(6)
(FS_MAT_DIFFUSE, FS_MAT_LIGHT, FS_MAT_METAL, FS_MAT_MIRROR, FS_MAT_GLASS, FS_MAT_LIQUID,
 FS_MAT_CERAMIC, FS_MAT_GLOW, FS_MAT_FOG) = range(9)

FS_TRACE_DEFAULT = 0
FS_TRACE_NO_JITTER = 1
FS_TRACE_NO_DOF = 2
FS_TRACE_WAVEFRONT = 4
FS_TRACE_NO_RESOLVE = 8
FS_TRACE_ASYNC = 16
FS_TRACE_PER_THREAD = 32
FS_TRACE_COOP = 64
FS_TRACE_LBVH_LANE = 128
FS_TRACE_MEGAKERNEL = 256
FS_TRACE_SHARD_TILES = 512
FS_TRACE_SHARD_SAMPLES = 1024

# numpy dtypes of the flattened arrays -------------------------------------------------
NODE_DT = np.dtype([("bmin", "<f4", 3), ("first", "<u4"), ("bmax", "<f4", 3), ("count", "<u4")])
TRI_AUX_DT = np.dtype([("p1", "<f4", 3), ("face", "<u4"), ("v0a", "<f4"), ("v0b", "<f4"),
                       ("v1a", "<f4"), ("v1b", "<f4"), ("dot00", "<f4"), ("dot01", "<f4"),
                       ("dot11", "<f4"), ("inv_denom", "<f4")])
FACE_DT = np.dtype([("n", "<f4", (3, 3)), ("uv", "<f4", (3, 2)), ("pad", "<f4")])
OBJECT_DT = np.dtype([("type", "<u4"), ("material", "<u4"), ("u", "<u4", 4), ("f", "<f4", 24),
                      ("bmin", "<f4", 3), ("bmax", "<f4", 3)])
MATERIAL_DT = np.dtype([("type", "<u4"), ("id", "<u4"), ("diffuse", "<f4", 3),
                        ("emissive", "<f4", 3), ("roughness", "<f4"), ("index", "<f4"),
                        ("reflectivity", "<f4"), ("frost", "<f4"), ("density", "<f4"),
                        ("tex_scale", "<f4"), ("texture", "<i4"), ("pad", "<u4")])
TEXTURE_DT = np.dtype([("width", "<u4"), ("height", "<u4"), ("offset", "<u8")])
assert NODE_DT.itemsize == 32 and TRI_AUX_DT.itemsize == 48 and FACE_DT.itemsize == 64
assert OBJECT_DT.itemsize == 144 and MATERIAL_DT.itemsize == 64 and TEXTURE_DT.itemsize == 16

BLOB_MAGIC = 0x31325346
BLOB_HEADER_DT = np.dtype([("magic", "<u4"), ("abi_version", "<u4"), ("n_nodes", "<u4"),
                           ("n_scene_refs", "<u4"), ("n_tris", "<u4"), ("n_faces", "<u4"),
                           ("n_objects", "<u4"), ("n_materials", "<u4"), ("n_textures", "<u4"),
                           ("n_lut", "<u4"), ("n_texels", "<u8"), ("scene_root", "<i4"),
                           ("sky_material", "<u4"), ("offset", "<u8", 10), ("total_bytes", "<u8")])


class fs_scene_desc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("n_nodes", C.c_uint32),
                ("n_scene_refs", C.c_uint32), ("n_tris", C.c_uint32), ("n_faces", C.c_uint32),
                ("n_objects", C.c_uint32), ("n_materials", C.c_uint32),
                ("n_textures", C.c_uint32), ("n_texels", C.c_uint64), ("scene_root", C.c_int32),
                ("sky_material", C.c_uint32), ("n_lut", C.c_uint32), ("pad", C.c_uint32),
                ("nodes", C.c_void_p), ("scene_refs", C.c_void_p), ("tri_plane", C.c_void_p),
                ("tri_aux", C.c_void_p), ("faces", C.c_void_p), ("objects", C.c_void_p),
                ("materials", C.c_void_p), ("textures", C.c_void_p), ("texels", C.c_void_p),
                ("lut", C.c_void_p)]


FS_TRACE_RAYS_STATIC_MAX = 32768  # most rays per fs_trace_rays_static call (include/fs2_cabi.h)
# fs_ray_hit (fs_trace_rays)
RAY_HIT_DT = np.dtype([("param", "<f4"), ("point", "<f4", 3), ("normal", "<f4", 3), ("uv", "<f4", 2),
                       ("object", "<u4"), ("face", "<u4"), ("material", "<u4"), ("is_internal", "<u4")])


class fs_camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("target", C.c_float * 3), ("fov_y", C.c_float),
                ("aperture_size", C.c_float), ("focal_depth", C.c_float), ("z_near", C.c_float),
                ("z_far", C.c_float), ("fast_render_enabled", C.c_int32)]


class fs_mesh_tree(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("tri_plane", C.c_void_p), ("tri_aux", C.c_void_p),
                ("n_nodes", C.c_uint32), ("n_tris", C.c_uint32), ("build_ms", C.c_float)]


class fs_post_params(C.Structure):
    _fields_ = [("denoise_passes", C.c_int32), ("sigma_normal", C.c_float),
                ("sigma_depth", C.c_float), ("sigma_colour", C.c_float),
                ("bloom_threshold", C.c_float), ("bloom_strength", C.c_float),
                ("bloom_radius", C.c_int32), ("flare_strength", C.c_float),
                ("flare_ghosts", C.c_int32)]


def post_params(denoise_passes=4, sigma_normal=64.0, sigma_depth=0.02, sigma_colour=4.0,
                bloom_threshold=1.0, bloom_strength=0.0, bloom_radius=12, flare_strength=0.0,
                flare_ghosts=0):
    return fs_post_params(denoise_passes, sigma_normal, sigma_depth, sigma_colour,
                          bloom_threshold, bloom_strength, bloom_radius, flare_strength,
                          flare_ghosts)


class fs_stats(C.Structure):
    _fields_ = [("n_paths", C.c_uint64), ("n_segments", C.c_uint64), ("n_node", C.c_uint64),
                ("n_tri", C.c_uint64), ("n_bary", C.c_uint64), ("n_prim", C.c_uint64),
                ("n_shade", C.c_uint64), ("launches", C.c_uint64), ("last_trace_ms", C.c_float),
                ("last_kernel_ms", C.c_float), ("n_lbnode", C.c_uint64),
                ("n_tritest", C.c_uint64)]


def make_camera(origin=(-5.80, 7.05, -47.06), target=(0.0, 8.94, 0.0), fov_y=45.0,
                aperture_size=1.5, focal_depth=80.0, z_near=1.0, z_far=10000.0,
                fast_render_enabled=False):
    """Camera POD with the defaults of reference source/camera.cpp:6-24 and the default
    viewpoint of reference source/main.cpp:87-88."""
    c = fs_camera()
    c.origin[:] = [np.float32(v) for v in origin]
    c.target[:] = [np.float32(v) for v in target]
    c.fov_y = fov_y
    c.aperture_size = aperture_size
    c.focal_depth = focal_depth
    c.z_near = z_near
    c.z_far = z_far
    c.fast_render_enabled = 1 if fast_render_enabled else 0
    return c


def camera_from_file(path):
    c = fs_camera()
    with open(path, "rb") as f:
        data = f.read()
    assert len(data) == C.sizeof(fs_camera)
    C.memmove(C.addressof(c), data, len(data))
    return c


def camera_to_dict(c):
    return {"origin": list(c.origin), "target": list(c.target), "fov_y": c.fov_y,
            "aperture_size": c.aperture_size, "focal_depth": c.focal_depth,
            "z_near": c.z_near, "z_far": c.z_far,
            "fast_render_enabled": bool(c.fast_render_enabled)}


_ARRAYS = [("nodes", NODE_DT), ("scene_refs", np.dtype("<u4")), ("tri_plane", np.dtype("<f4")),
           ("tri_aux", TRI_AUX_DT), ("faces", FACE_DT), ("objects", OBJECT_DT),
           ("materials", MATERIAL_DT), ("textures", TEXTURE_DT), ("texels", np.dtype("<f4")),
           ("lut", np.dtype("<f4"))]


class FlatScene:
    """A flattened scene held as numpy arrays, convertible to an fs_scene_desc whose
    pointers stay valid as long as this object lives."""

    def __init__(self, scene_root=-1, sky_material=0, **arrays):
        self.scene_root = int(scene_root)
        self.sky_material = int(sky_material)
        for name, dt in _ARRAYS:
            a = arrays.get(name)
            if a is None:
                a = np.zeros(0, dt)
            a = np.ascontiguousarray(a, dtype=dt)
            setattr(self, name, a)
        self.tri_plane = self.tri_plane.reshape(-1)
        self.lut = self.lut.reshape(-1)

    # ---- blob io --------------------------------------------------------------
    @classmethod
    def load(cls, path):
        raw = np.fromfile(path, dtype=np.uint8)
        h = raw[:BLOB_HEADER_DT.itemsize].view(BLOB_HEADER_DT)[0]
        if h["magic"] != BLOB_MAGIC or h["abi_version"] != FS2_ABI_VERSION:
            raise ValueError("%s: not an fs2 blob of ABI %d" % (path, FS2_ABI_VERSION))
        if int(h["total_bytes"]) != raw.size:
            raise ValueError("%s: truncated blob" % path)
        counts = [int(h["n_nodes"]), int(h["n_scene_refs"]), 4 * int(h["n_tris"]),
                  int(h["n_tris"]), int(h["n_faces"]), int(h["n_objects"]),
                  int(h["n_materials"]), int(h["n_textures"]), int(h["n_texels"]),
                  3 * int(h["n_lut"])]
        arrays = {}
        for k, (name, dt) in enumerate(_ARRAYS):
            off = int(h["offset"][k])
            arrays[name] = raw[off:off + counts[k] * dt.itemsize].view(dt).copy()
        return cls(scene_root=int(h["scene_root"]), sky_material=int(h["sky_material"]), **arrays)

    def save(self, path):
        h = np.zeros(1, BLOB_HEADER_DT)
        h["magic"] = BLOB_MAGIC
        h["abi_version"] = FS2_ABI_VERSION
        h["n_nodes"] = len(self.nodes)
        h["n_scene_refs"] = len(self.scene_refs)
        h["n_tris"] = len(self.tri_aux)
        h["n_faces"] = len(self.faces)
        h["n_objects"] = len(self.objects)
        h["n_materials"] = len(self.materials)
        h["n_textures"] = len(self.textures)
        h["n_lut"] = len(self.lut) // 3
        h["n_texels"] = len(self.texels)
        h["scene_root"] = self.scene_root
        h["sky_material"] = self.sky_material
        pad16 = lambda n: (n + 15) & ~15
        off = pad16(BLOB_HEADER_DT.itemsize)
        chunks = []
        for k, (name, _) in enumerate(_ARRAYS):
            b = getattr(self, name).tobytes()
            h["offset"][0][k] = off
            chunks.append(b + b"\0" * (pad16(len(b)) - len(b)))
            off += pad16(len(b))
        h["total_bytes"] = off
        hb = h.tobytes()
        with open(path, "wb") as f:
            f.write(hb + b"\0" * (pad16(len(hb)) - len(hb)))
            for c in chunks:
                f.write(c)

    # ---- ctypes view -----------------------------------------------------------
    def desc(self):
        d = fs_scene_desc()
        d.abi_version = FS2_ABI_VERSION
        d.n_nodes = len(self.nodes)
        d.n_scene_refs = len(self.scene_refs)
        d.n_tris = len(self.tri_aux)
        d.n_faces = len(self.faces)
        d.n_objects = len(self.objects)
        d.n_materials = len(self.materials)
        d.n_textures = len(self.textures)
        d.n_texels = len(self.texels)
        d.scene_root = self.scene_root
        d.sky_material = self.sky_material
        d.n_lut = len(self.lut) // 3
        for name, _ in _ARRAYS:
            a = getattr(self, name)
            setattr(d, name, a.ctypes.data if a.size else None)
        return d

    def same_as(self, other):
        """Byte equality of every array (used to compare our octree build with the
        reference's)."""
        if self.scene_root != other.scene_root or self.sky_material != other.sky_material:
            return False
        return all(getattr(self, n).tobytes() == getattr(other, n).tobytes() for n, _ in _ARRAYS)

    def diff(self, other):
        out = []
        if self.scene_root != other.scene_root:
            out.append("scene_root %d != %d" % (self.scene_root, other.scene_root))
        if self.sky_material != other.sky_material:
            out.append("sky_material %d != %d" % (self.sky_material, other.sky_material))
        for n, _ in _ARRAYS:
            a, b = getattr(self, n), getattr(other, n)
            if a.shape != b.shape:
                out.append("%s: shape %s != %s" % (n, a.shape, b.shape))
            elif a.tobytes() != b.tobytes():
                av = a.view(np.uint8).reshape(len(a), -1) if len(a) else a
                bv = b.view(np.uint8).reshape(len(b), -1) if len(b) else b
                bad = np.nonzero((av != bv).any(axis=1))[0]
                out.append("%s: %d of %d records differ (first %d)" % (n, len(bad), len(a), bad[0]))
        return out

    def nbytes(self):
        return sum(getattr(self, n).nbytes for n, _ in _ARRAYS)


PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_DIR = os.path.dirname(PKG_DIR)
