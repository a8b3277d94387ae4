#!/usr/bin/env python3
"""Deterministic generators for every scene, mesh and texture used by the
tests and the bench.  The reference ships no assets, so the
configs of BASELINE.json are authored here with the reference's own `.scene`
grammar (reference source/scene.cpp:493-535).

    python scenes/gen_scenes.py OUTDIR [name ...]

Each scene NAME produces OUTDIR/NAME.scene (+ .obj / .bmp it refers to, by
absolute path, because the reference opens mesh/texture paths relative to the
process cwd).  Generation is seeded and byte-reproducible.
"""
import math
import os
import struct
import sys

import numpy as np

DEFAULT_CAMERA = ((-5.80, 7.05, -47.06), (0.00, 8.94, 0.00))  # reference main.cpp:87-88


def _f(x):
    return "%.6f" % x


def material(name, **kw):
    lines = ["material %s" % name, "{"]
    order = ["color", "emission", "metallic", "roughness", "index", "brdf", "reflectivity",
             "frostiness", "texture_scale", "texture"]  # texture_scale before texture (App. D)
    for k in order:
        if k not in kw:
            continue
        v = kw[k]
        if isinstance(v, (tuple, list)):
            lines.append("    %s %s" % (k, " ".join(_f(c) for c in v)))
        elif isinstance(v, str):
            lines.append("    %s %s" % (k, v))
        elif isinstance(v, int):
            lines.append("    %s %d" % (k, v))
        else:
            lines.append("    %s %s" % (k, _f(v)))
    lines.append("}")
    return "\n".join(lines) + "\n"


def block(kind, **kw):
    lines = [kind, "{"]
    for k, v in kw.items():
        if isinstance(v, (tuple, list)):
            lines.append("    %s %s" % (k, " ".join(_f(c) for c in v)))
        elif isinstance(v, str):
            lines.append("    %s %s" % (k, v))
        else:
            lines.append("    %s %s" % (k, _f(v)))
    lines.append("}")
    return "\n".join(lines) + "\n"


def camera(position=DEFAULT_CAMERA[0], target=DEFAULT_CAMERA[1], fov=45.0, aperture=0.0,
           focal_depth=47.0):
    return block("camera", position=position, target=target, fov=fov, aperture=aperture,
                 focal_depth=focal_depth)


# ----------------------------------------------------------------- assets
def write_bmp(path, rgb):
    """24-bit BMP, bottom-up rows as stored (reference source/bitmap.cpp:84-102)."""
    h, w, _ = rgb.shape
    row = (w * 3 + 3) & ~3
    data = bytearray()
    for y in range(h):
        line = rgb[y, :, ::-1].astype(np.uint8).tobytes()
        data += line + b"\0" * (row - w * 3)
    with open(path, "wb") as f:
        f.write(b"BM" + struct.pack("<IHHI", 54 + len(data), 0, 0, 54))
        f.write(struct.pack("<IiiHHIIiiII", 40, w, h, 1, 24, 0, len(data), 2835, 2835, 0, 0))
        f.write(data)


def checker_texture(path, n=64, cells=8, seed=3):
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:n, 0:n]
    c = ((xx // (n // cells)) + (yy // (n // cells))) & 1
    img = np.zeros((n, n, 3), np.float64)
    img[c == 0] = (225, 215, 190)
    img[c == 1] = (60, 90, 140)
    img += rng.randint(-10, 10, size=img.shape)
    write_bmp(path, np.clip(img, 0, 255))


def sky_texture(path, w=512, h=256):
    """Analytic sun + gradient environment (sphere-mapped, reference scene.cpp:157-162)."""
    v, u = np.mgrid[0:h, 0:w]
    el = v / (h - 1.0)
    img = np.zeros((h, w, 3), np.float64)
    img[..., 0] = 70 + 110 * el
    img[..., 1] = 110 + 100 * el
    img[..., 2] = 200 + 40 * el
    du = (u - 0.62 * w) / w
    dv = (v - 0.30 * h) / h
    sun = np.exp(-(du * du * 4 + dv * dv) * 900.0)
    img += sun[..., None] * 255.0
    write_bmp(path, np.clip(img, 0, 255))


def uv_sphere_obj(path, nu, nv, radius=1.0, bump=0.08, seed=7, with_attrs=True):
    """Displaced UV sphere: 2*nu*(nv-1) triangles with normals and uvs.
    (nu, nv) = (320, 161) gives 102 400 triangles (BASELINE config 2)."""
    rng = np.random.RandomState(seed)
    ph = rng.uniform(0, 2 * math.pi, size=6)
    lines = ["# fs2-b200 procedural displaced sphere nu=%d nv=%d seed=%d" % (nu, nv, seed)]
    verts = []
    norms = []
    uvs = []
    for j in range(nv + 1):
        th = math.pi * j / nv
        for i in range(nu):
            p = 2 * math.pi * i / nu
            d = 1.0 + bump * (math.sin(5 * p + ph[0]) * math.sin(4 * th + ph[1]) +
                              0.5 * math.sin(11 * p + ph[2]) * math.sin(9 * th + ph[3]) +
                              0.25 * math.sin(23 * p + ph[4]) * math.sin(17 * th + ph[5]))
            x = math.sin(th) * math.cos(p)
            y = math.cos(th)
            z = math.sin(th) * math.sin(p)
            verts.append((radius * d * x, radius * d * y, radius * d * z))
            norms.append((x, y, z))
            uvs.append((i / nu, j / nv))
    for v in verts:
        lines.append("v %.6f %.6f %.6f" % v)
    if with_attrs:
        for n in norms:
            lines.append("vn %.6f %.6f %.6f" % n)
        for t in uvs:
            lines.append("vt %.6f %.6f" % t)

    def idx(i, j):
        return j * nu + (i % nu) + 1

    def face(a, b, c):
        if with_attrs:
            return "f %d/%d/%d %d/%d/%d %d/%d/%d" % (a, a, a, b, b, b, c, c, c)
        return "f %d %d %d" % (a, b, c)

    for j in range(nv):
        for i in range(nu):
            a, b, c, d = idx(i, j), idx(i + 1, j), idx(i + 1, j + 1), idx(i, j + 1)
            if j != 0:
                lines.append(face(a, b, c))
            if j != nv - 1:
                lines.append(face(a, c, d))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return 2 * nu * (nv - 1)


def torus_knot_obj(path, nu, nv, seed=11):
    """(2,3) torus knot tube: 2*nu*nv triangles, concave, self-occluding."""
    lines = ["# fs2-b200 procedural torus knot nu=%d nv=%d" % (nu, nv)]
    P = np.zeros((nu, 3))
    for i in range(nu):
        t = 2 * math.pi * i / nu
        r = 2 + math.cos(3 * t)
        P[i] = (r * math.cos(2 * t), r * math.sin(2 * t), -math.sin(3 * t))
    T = np.roll(P, -1, 0) - np.roll(P, 1, 0)
    T /= np.linalg.norm(T, axis=1)[:, None]
    up = np.array([0.0, 0.0, 1.0])
    N = np.cross(T, up)
    N /= np.linalg.norm(N, axis=1)[:, None]
    B = np.cross(T, N)
    for i in range(nu):
        for j in range(nv):
            a = 2 * math.pi * j / nv
            n = math.cos(a) * N[i] + math.sin(a) * B[i]
            p = P[i] + 0.45 * n
            lines.append("v %.6f %.6f %.6f" % tuple(p))
    for i in range(nu):
        for j in range(nv):
            a = 2 * math.pi * j / nv
            n = math.cos(a) * N[i] + math.sin(a) * B[i]
            lines.append("vn %.6f %.6f %.6f" % tuple(n))
    for i in range(nu):
        for j in range(nv):
            lines.append("vt %.6f %.6f" % (i / nu, j / nv))

    def idx(i, j):
        return (i % nu) * nv + (j % nv) + 1

    for i in range(nu):
        for j in range(nv):
            a, b, c, d = idx(i, j), idx(i + 1, j), idx(i + 1, j + 1), idx(i, j + 1)
            lines.append("f %d/%d/%d %d/%d/%d %d/%d/%d" % (a, a, a, b, b, b, c, c, c))
            lines.append("f %d/%d/%d %d/%d/%d %d/%d/%d" % (a, a, a, c, c, c, d, d, d))
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return 2 * nu * nv


def grid_obj(path, nx, nz, size=60.0, amp=3.0, seed=5):
    """Displaced height-field grid: 2*nx*nz triangles (2048x1024 -> 4.19 M, config 4)."""
    rng = np.random.RandomState(seed)
    ph = rng.uniform(0, 2 * math.pi, size=8)
    xs = np.linspace(-size, size, nx + 1)
    zs = np.linspace(-size * nz / nx, size * nz / nx, nz + 1)
    X, Z = np.meshgrid(xs, zs)
    Y = amp * (np.sin(X * 0.21 + ph[0]) * np.cos(Z * 0.17 + ph[1]) +
               0.4 * np.sin(X * 0.83 + ph[2]) * np.sin(Z * 0.71 + ph[3]) +
               0.15 * np.sin(X * 2.9 + ph[4]) * np.cos(Z * 3.3 + ph[5]))
    with open(path, "w") as f:
        f.write("# fs2-b200 procedural grid %dx%d\n" % (nx, nz))
        V = np.stack([X.ravel(), Y.ravel(), Z.ravel()], 1)
        f.write("".join("v %.5f %.5f %.5f\n" % tuple(v) for v in V))
        j, i = np.mgrid[0:nz, 0:nx]
        a = (j * (nx + 1) + i + 1).ravel()
        b = a + 1
        d = a + nx + 1
        c = d + 1
        F = np.stack([a, c, b, a, d, c], 1).reshape(-1, 3)
        f.write("".join("f %d %d %d\n" % tuple(t) for t in F))
    return 2 * nx * nz


# ----------------------------------------------------------------- scenes
def common_materials(outdir, textured_floor=True):
    s = ""
    s += material("sky_mat", emission=(0.55, 0.65, 0.85))
    if textured_floor:
        tex = os.path.join(outdir, "checker.bmp")
        if not os.path.exists(tex):
            checker_texture(tex)
        s += material("floor_mat", color=(0.8, 0.8, 0.8), texture_scale=0.02, texture=tex)
    else:
        s += material("floor_mat", color=(0.7, 0.7, 0.65))
    s += material("red_diffuse", color=(0.85, 0.2, 0.15))
    s += material("green_ceramic", color=(0.2, 0.7, 0.3), roughness=0.6)
    s += material("gold_metal", color=(0.95, 0.75, 0.3), metallic=0.3)
    s += material("mirror_mat", color=(0.95, 0.95, 0.95), metallic=1.0)
    s += material("glass_mat", color=(0.95, 0.98, 1.0), brdf=2, index=0.75, reflectivity=0.1)
    s += material("frost_mat", color=(0.9, 0.95, 1.0), brdf=2, index=0.75, reflectivity=0.1,
                  frostiness=0.2)
    s += material("liquid_mat", color=(0.7, 0.85, 1.0), brdf=1, index=0.75, reflectivity=0.4)
    s += material("lamp_mat", emission=(12.0, 11.0, 9.0))
    s += material("blue_cuboid", color=(0.25, 0.35, 0.9), roughness=0.3)
    return s


def scene_demo(outdir, aperture=0.0):
    """Config 1: native primitives only, brute-force object loop (N < 182)."""
    s = "# fs2-b200 demo scene (BASELINE config 1)\n"
    s += common_materials(outdir)
    s += block("sky", material="sky_mat")
    s += camera(aperture=aperture)
    # floor: normal tilted off the y axis so the quad has non-zero tangents
    s += block("quad", material="floor_mat", position=(0, 0, 0), normal=(0.02, 1.0, 0.01),
               width=200.0, height=200.0)
    s += block("cuboid", material="blue_cuboid", position=(9.0, 4.0, 6.0), width=8.0, height=8.0,
               depth=8.0, rotation=(0.2, 1.0, 0.1, 0.6))
    s += block("sphere", material="green_ceramic", position=(-12.0, 5.0, 4.0), radius=5.0)
    s += block("sphere", material="gold_metal", position=(-1.5, 4.0, -6.0), radius=4.0)
    s += block("sphere", material="mirror_mat", position=(2.0, 12.5, 14.0), radius=6.0)
    s += block("sphere", material="glass_mat", position=(5.5, 3.0, -14.0), radius=3.0)
    s += block("sphere", material="liquid_mat", position=(-7.5, 2.5, -16.0), radius=2.5)
    s += block("sphere", material="red_diffuse", position=(14.0, 3.0, -8.0), radius=3.0)
    s += block("sphere", material="lamp_mat", position=(-4.0, 30.0, -10.0), radius=5.0)
    return s


def scene_many(outdir, n=300, seed=1):
    """>= 182 objects: the scene octree is built and used (scene.cpp:216-225)."""
    rng = np.random.RandomState(seed)
    s = "# fs2-b200 many-object scene: scene octree active\n"
    s += common_materials(outdir, textured_floor=False)
    s += block("sky", material="sky_mat")
    s += camera(position=(-10.0, 22.0, -75.0), target=(0.0, 6.0, 0.0))
    mats = ["red_diffuse", "green_ceramic", "gold_metal", "mirror_mat", "glass_mat", "liquid_mat",
            "blue_cuboid"]
    s += block("quad", material="floor_mat", position=(0, 0, 0), normal=(0.02, 1.0, 0.01),
               width=160.0, height=160.0)
    s += block("quad", material="floor_mat", position=(0, 20, 45), normal=(0.05, 0.1, -1.0),
               width=120.0, height=60.0)
    s += block("sphere", material="lamp_mat", position=(0.0, 60.0, -20.0), radius=8.0)
    for k in range(n):
        p = (rng.uniform(-40, 40), rng.uniform(1.0, 25.0), rng.uniform(-35, 40))
        if k % 10 == 9:
            s += block("cuboid", material=mats[k % len(mats)], position=p,
                       width=rng.uniform(1, 4), height=rng.uniform(1, 4), depth=rng.uniform(1, 4),
                       rotation=(rng.uniform(-1, 1), 1.0, rng.uniform(-1, 1), rng.uniform(0, 3)))
        else:
            s += block("sphere", material=mats[k % len(mats)], position=p,
                       radius=rng.uniform(0.6, 2.4))
    return s


def scene_mesh(outdir, nu, nv, name, aperture=0.0, knot=False):
    """Config 2 family: displaced-sphere meshes x3 (metal / ceramic / mirror) + floor + lamp."""
    obj = os.path.join(outdir, "%s_sphere_%dx%d.obj" % (name, nu, nv))
    if not os.path.exists(obj):
        uv_sphere_obj(obj, nu, nv)
    s = "# fs2-b200 mesh scene %s (BASELINE config 2 family)\n" % name
    s += common_materials(outdir)
    s += block("sky", material="sky_mat")
    s += camera(position=(-4.0, 11.0, -46.0), target=(0.0, 7.0, 0.0), aperture=aperture,
                focal_depth=46.0)
    s += block("quad", material="floor_mat", position=(0, 0, 0), normal=(0.02, 1.0, 0.01),
               width=200.0, height=200.0)
    s += block("sphere", material="lamp_mat", position=(-10.0, 40.0, -20.0), radius=6.0)
    s += block("mesh", file=obj, material="gold_metal", translation=(-13.0, 7.0, 2.0),
               scale=(6.0, 6.0, 6.0), rotation=(0.0, 1.0, 0.0, 0.4))
    s += block("mesh", file=obj, material="green_ceramic", translation=(0.0, 7.5, -4.0),
               scale=(6.5, 6.5, 6.5), rotation=(1.0, 0.0, 0.0, 0.3))
    s += block("mesh", file=obj, material="mirror_mat", translation=(13.0, 7.0, 4.0),
               scale=(6.0, 6.0, 6.0), rotation=(0.0, 0.0, 1.0, 0.7))
    if knot:
        kobj = os.path.join(outdir, "%s_knot.obj" % name)
        if not os.path.exists(kobj):
            torus_knot_obj(kobj, 96, 12)
        s += block("mesh", file=kobj, material="red_diffuse", translation=(2.0, 19.0, 10.0),
                   scale=(2.2, 2.2, 2.2), rotation=(1.0, 0.3, 0.0, 1.1))
    return s


def sky_texture_exr(path, w=1024, h=512):
    """The same analytic sky as a float .exr (HALF, ZIP): a sun of radiance ~40 that an 8-bit
    .bmp would clip ('float env texture written as .exr')."""
    from exr_write import write_exr
    v, u = np.mgrid[0:h, 0:w]
    el = v / (h - 1.0)
    img = np.zeros((h, w, 3), np.float64)
    img[..., 0] = (70 + 110 * el) / 255.0
    img[..., 1] = (110 + 100 * el) / 255.0
    img[..., 2] = (200 + 40 * el) / 255.0
    du = (u - 0.62 * w) / w
    dv = (v - 0.30 * h) / h
    sun = np.exp(-(du * du * 4 + dv * dv) * 900.0)
    img += sun[..., None] * 40.0
    img = img.astype(np.float16)
    write_exr(path, {"R": img[..., 0], "G": img[..., 1], "B": img[..., 2]}, comp="zip")


def scene_glass(outdir, aperture=0.0, sky_size=(512, 256), exr_sky=False):
    """Config 3: glass / frosted glass / liquid primitives, ceramic, env-map sky."""
    sky = os.path.join(outdir, "sky_%dx%d.%s" % (sky_size + ("exr" if exr_sky else "bmp",)))
    if not os.path.exists(sky):
        (sky_texture_exr if exr_sky else sky_texture)(sky, *sky_size)
    s = "# fs2-b200 glass + liquid scene under an environment texture (BASELINE config 3)\n"
    s += common_materials(outdir)
    s += material("env_mat", emission=(1.0, 1.0, 1.0), texture_scale=1.0, texture=sky)
    s += block("sky", material="env_mat")
    s += camera(aperture=aperture)
    s += block("quad", material="floor_mat", position=(0, 0, 0), normal=(0.02, 1.0, 0.01),
               width=200.0, height=200.0)
    s += block("sphere", material="glass_mat", position=(-8.0, 6.0, 0.0), radius=6.0)
    s += block("sphere", material="liquid_mat", position=(-8.0, 6.0, 0.0), radius=3.0)
    s += block("sphere", material="frost_mat", position=(6.0, 5.0, -6.0), radius=5.0)
    s += block("cuboid", material="glass_mat", position=(16.0, 5.0, 8.0), width=8.0, height=10.0,
               depth=8.0, rotation=(0.0, 1.0, 0.0, 0.5))
    s += block("sphere", material="green_ceramic", position=(16.0, 5.0, 8.0), radius=2.5)
    s += block("cuboid", material="liquid_mat", position=(-1.0, 2.0, -18.0), width=10.0,
               height=4.0, depth=6.0, rotation=(0.0, 1.0, 0.0, -0.3))
    s += block("sphere", material="mirror_mat", position=(0.0, 16.0, 22.0), radius=9.0)
    s += block("sphere", material="lamp_mat", position=(-20.0, 35.0, -5.0), radius=4.0)
    return s


def scene_grid(outdir, nx, nz, name, aperture=1.5):
    """Config 4 family: one multi-million-triangle height field + DOF."""
    obj = os.path.join(outdir, "%s_grid_%dx%d.obj" % (name, nx, nz))
    if not os.path.exists(obj):
        grid_obj(obj, nx, nz)
    s = "# fs2-b200 large-mesh scene %s (BASELINE config 4 family)\n" % name
    s += common_materials(outdir, textured_floor=False)
    s += block("sky", material="sky_mat")
    s += camera(position=(-8.0, 26.0, -78.0), target=(0.0, 2.0, 0.0), aperture=aperture,
                focal_depth=80.0)
    s += block("sphere", material="lamp_mat", position=(-30.0, 70.0, -20.0), radius=10.0)
    s += block("sphere", material="mirror_mat", position=(0.0, 12.0, 0.0), radius=7.0)
    s += block("sphere", material="glass_mat", position=(-18.0, 9.0, -14.0), radius=5.0)
    s += block("mesh", file=obj, material="green_ceramic", translation=(0.0, 0.001, 0.0),
               scale=(1.0, 1.0, 1.0), rotation=(0.0, 0.0, 0.0, 0.0))
    return s


# ----------------------------------------------------------------- adversarial meshes
# Fixtures aimed at the leaf BVH's exactness argument (csrc/fs2_lbvh_build.h record_box): the
# reference's linear leaf loop and the BVH walk must agree where float geometry is at its worst.
def _write_obj(path, verts, faces, comment):
    with open(path, "w") as f:
        f.write("# fs2-b200 %s\n" % comment)
        f.write("".join("v %.9g %.9g %.9g\n" % tuple(v) for v in verts))
        f.write("".join("f %d %d %d\n" % tuple(t) for t in faces))
    return len(faces)


def needles_obj(path, n=2400, seed=21):
    """Sliver triangles with aspect ratios 10 ... 10^5 (their stored barycentric denominators
    lose most or all of their bits), mixed with ordinary ones, in a +-12 cube."""
    rng = np.random.RandomState(seed)
    verts, faces = [], []
    for k in range(n):
        base = rng.uniform(-12, 12, 3)
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        length = rng.uniform(4.0, 18.0)
        side = np.cross(d, rng.normal(size=3))
        side /= np.linalg.norm(side)
        aspect = [1.5, 10.0, 100.0, 1.0e3, 1.0e4, 1.0e5][k % 6]
        a = base
        b = base + d * length
        c = base + d * length * rng.uniform(0.2, 0.8) + side * (length / aspect)
        i = len(verts)
        verts += [a, b, c]
        faces.append((i + 1, i + 2, i + 3))
    return _write_obj(path, verts, faces, "needle triangles n=%d seed=%d" % (n, seed))


def _small_sphere(nu, nv):
    """Unit sphere as (verts [n,3], faces [m,3] 1-based); 2*nu*(nv-1) triangles."""
    verts = []
    for j in range(nv + 1):
        th = math.pi * j / nv
        for i in range(nu):
            ph = 2 * math.pi * i / nu
            verts.append((math.sin(th) * math.cos(ph), math.cos(th), math.sin(th) * math.sin(ph)))
    faces = []
    for j in range(nv):
        for i in range(nu):
            a = j * nu + i
            b = j * nu + (i + 1) % nu
            c = a + nu
            d = b + nu
            if j > 0:
                faces.append((a + 1, b + 1, c + 1))
            if j < nv - 1:
                faces.append((b + 1, d + 1, c + 1))
    return np.array(verts), np.array(faces)


def wide_range_obj(path):
    """One mesh whose coordinates span 10^-3 ... 5*10^3: seven spheres of radius 10^-3 ... 10^3
    fanned out around the origin, each four radii away (so all are about equally large on
    screen), plus two ground triangles 8 000 units across."""
    sv, sf = _small_sphere(16, 9)
    verts, faces = [], []
    for k in range(7):
        r = 10.0 ** (k - 3)
        ang = math.radians(-36 + 12 * k)
        c = np.array([4.0 * r * math.sin(ang), 0.15 * r * (k - 3), 4.0 * r * math.cos(ang)])
        base = len(verts)
        verts += list(sv * r + c)
        faces += [tuple(int(i) + base for i in t) for t in sf]
    base = len(verts)
    verts += [np.array(v) for v in ((-4000.0, -1200.0, -100.0), (4000.0, -1200.0, -100.0),
                                    (4000.0, -1100.0, 4900.0), (-4000.0, -1100.0, 4900.0))]
    faces += [(base + 1, base + 3, base + 2), (base + 1, base + 4, base + 3)]
    return _write_obj(path, verts, faces, "wide coordinate range")


def scene_needles(outdir):
    obj = os.path.join(outdir, "needles.obj")
    if not os.path.exists(obj):
        needles_obj(obj)
    s = "# fs2-b200 adversarial: needle triangles\n"
    s += common_materials(outdir, textured_floor=False)
    s += block("sky", material="sky_mat")
    s += camera(position=(-6.0, 9.0, -44.0), target=(0.0, 1.0, 0.0))
    s += block("sphere", material="lamp_mat", position=(-10.0, 40.0, -20.0), radius=6.0)
    s += block("mesh", file=obj, material="red_diffuse", translation=(0.0, 2.0, 0.0),
               scale=(1.0, 1.0, 1.0), rotation=(0.0, 1.0, 0.0, 0.3))
    return s


def scene_far_mesh(outdir):
    """The config-2 layout moved to (9000, 9000, 9000): one ulp is ~10^-3 there."""
    obj = os.path.join(outdir, "far_sphere_48x25.obj")
    if not os.path.exists(obj):
        uv_sphere_obj(obj, 48, 25)
    o = 9000.0
    s = "# fs2-b200 adversarial: mesh far from the origin\n"
    s += common_materials(outdir, textured_floor=False)
    s += block("sky", material="sky_mat")
    s += camera(position=(o - 4.0, o + 11.0, o - 46.0), target=(o, o + 7.0, o))
    s += block("quad", material="floor_mat", position=(o, o, o), normal=(0.02, 1.0, 0.01),
               width=200.0, height=200.0)
    s += block("sphere", material="lamp_mat", position=(o - 10.0, o + 40.0, o - 20.0), radius=6.0)
    s += block("mesh", file=obj, material="gold_metal", translation=(o - 9.0, o + 7.0, o + 2.0),
               scale=(6.0, 6.0, 6.0), rotation=(0.0, 1.0, 0.0, 0.4))
    s += block("mesh", file=obj, material="mirror_mat", translation=(o + 9.0, o + 7.0, o + 4.0),
               scale=(6.0, 6.0, 6.0), rotation=(0.0, 0.0, 1.0, 0.7))
    return s


def scene_far_camera(outdir):
    """A unit-sized mesh at the origin seen from 500 units away: the camera lies far outside the
    scene's own coordinate scale (the leaf-BVH pads are re-sized for it, fs2_cabi.cu)."""
    obj = os.path.join(outdir, "far_sphere_48x25.obj")
    if not os.path.exists(obj):
        uv_sphere_obj(obj, 48, 25)
    s = "# fs2-b200 adversarial: camera far outside the scene scale\n"
    s += common_materials(outdir, textured_floor=False)
    s += block("sky", material="sky_mat")
    s += camera(position=(-300.0, 60.0, -400.0), target=(0.0, 0.0, 0.0), fov=0.6)
    s += block("sphere", material="lamp_mat", position=(-1.0, 4.0, -2.0), radius=0.6)
    s += block("mesh", file=obj, material="green_ceramic", translation=(0.0, 0.0, 0.0),
               scale=(1.0, 1.0, 1.0), rotation=(0.0, 1.0, 0.0, 0.0))
    s += block("mesh", file=obj, material="mirror_mat", translation=(2.2, 0.2, 0.5),
               scale=(0.8, 0.8, 0.8), rotation=(1.0, 0.0, 0.0, 0.5))
    return s


def scene_wide_range(outdir):
    obj = os.path.join(outdir, "wide_range.obj")
    if not os.path.exists(obj):
        wide_range_obj(obj)
    s = "# fs2-b200 adversarial: coordinates spanning 1e-3 ... 5e3 in one mesh\n"
    s += common_materials(outdir, textured_floor=False)
    s += block("sky", material="sky_mat")
    s += camera(position=(0.0, 0.0, 0.0), target=(0.0, -0.05, 1.0), fov=80.0)
    s += block("mesh", file=obj, material="green_ceramic", translation=(0.0, 0.0, 0.0),
               scale=(1.0, 1.0, 1.0), rotation=(0.0, 1.0, 0.0, 0.0))
    return s


def scene_random(outdir, seed):
    """Seeded random scene for the fuzz parity test: random camera (inside the cloud for odd
    seeds), spheres / rotated cuboids / tilted quads of every material, one or two small
    meshes at random transforms, and for every fourth seed enough objects to switch the scene
    octree on."""
    rng = np.random.RandomState(1000 + seed)
    s = "# fs2-b200 random scene %d\n" % seed
    s += common_materials(outdir, textured_floor=bool(seed & 1))
    s += block("sky", material="sky_mat")
    far = seed % 2 == 0
    pos = (rng.uniform(-30, 30), rng.uniform(4, 30), rng.uniform(-70, -40)) if far else \
          (rng.uniform(-10, 10), rng.uniform(3, 15), rng.uniform(-10, 10))
    s += camera(position=pos, target=(rng.uniform(-5, 5), rng.uniform(2, 10), rng.uniform(-5, 5)),
                fov=rng.uniform(30, 70), aperture=0.0)
    mats = ["red_diffuse", "green_ceramic", "gold_metal", "mirror_mat", "glass_mat", "frost_mat",
            "liquid_mat", "blue_cuboid", "floor_mat", "lamp_mat"]
    s += block("quad", material="floor_mat", position=(0, 0, 0),
               normal=(rng.uniform(-0.1, 0.1), 1.0, rng.uniform(-0.1, 0.1)), width=150.0, height=150.0)
    n = int(rng.randint(6, 30)) if seed % 4 else int(rng.randint(190, 260))
    for k in range(n):
        p3 = (rng.uniform(-35, 35), rng.uniform(0.5, 24.0), rng.uniform(-35, 35))
        m = mats[int(rng.randint(len(mats)))]
        kind = rng.randint(10)
        if kind < 6:
            s += block("sphere", material=m, position=p3, radius=rng.uniform(0.5, 5.0))
        elif kind < 9:
            s += block("cuboid", material=m, position=p3, width=rng.uniform(1, 8),
                       height=rng.uniform(1, 8), depth=rng.uniform(1, 8),
                       rotation=(rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(-1, 1),
                                 rng.uniform(0, 3)))
        else:
            s += block("quad", material=m, position=p3,
                       normal=(rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(-1, 1)),
                       width=rng.uniform(2, 20), height=rng.uniform(2, 20))
    for k in range(1 + seed % 2):
        if (seed + k) % 3 == 0:
            obj = os.path.join(outdir, "rnd_knot_%d.obj" % (seed % 3))
            if not os.path.exists(obj):
                torus_knot_obj(obj, 48 + 16 * (seed % 3), 8)
        else:
            nu = 16 + 8 * ((seed + k) % 4)
            obj = os.path.join(outdir, "rnd_sphere_%d.obj" % nu)
            if not os.path.exists(obj):
                uv_sphere_obj(obj, nu, nu // 2 + 1)
        sc = rng.uniform(2, 7)
        s += block("mesh", file=obj, material=mats[int(rng.randint(8))],
                   translation=(rng.uniform(-15, 15), rng.uniform(4, 14), rng.uniform(-15, 15)),
                   scale=(sc, sc * rng.uniform(0.6, 1.4), sc),
                   rotation=(rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(-1, 1),
                             rng.uniform(0, 3)))
    return s


SCENES = {
    "demo": lambda d: scene_demo(d),
    "demo_dof": lambda d: scene_demo(d, aperture=1.5),
    "many303": lambda d: scene_many(d, 300),
    "mesh_tiny": lambda d: scene_mesh(d, 24, 13, "tiny"),             # 576 tris/mesh
    "mesh_small": lambda d: scene_mesh(d, 64, 33, "small", knot=True),  # 4 096 tris/mesh + knot
    "mesh100k": lambda d: scene_mesh(d, 320, 161, "m100k"),           # 102 400 tris/mesh
    "glass": lambda d: scene_glass(d),
    "glass_small": lambda d: scene_glass(d, sky_size=(64, 32)),
    "glass_dof": lambda d: scene_glass(d, aperture=1.5),
    # config 3 as SURVEY 8(d) states it: float .exr environment; bench.py adds the fog cuboid
    # through the C++ API (FogMaterial is not expressible in the .scene grammar, App. A13)
    "glass3": lambda d: scene_glass(d, aperture=1.5, sky_size=(1024, 512), exr_sky=True),
    "grid_tiny": lambda d: scene_grid(d, 24, 12, "gtiny", aperture=0.0),
    "grid_small": lambda d: scene_grid(d, 64, 32, "gsmall"),
    "grid4m": lambda d: scene_grid(d, 2048, 1024, "g4m"),
    "needles": scene_needles,
    "far_mesh": scene_far_mesh,
    "far_camera": scene_far_camera,
    "wide_range": scene_wide_range,
}


def generate(outdir, names=None):
    os.makedirs(outdir, exist_ok=True)
    outdir = os.path.abspath(outdir)
    paths = {}
    for n in (names or SCENES.keys()):
        p = os.path.join(outdir, n + ".scene")
        if not os.path.exists(p):
            text = scene_random(outdir, int(n[6:])) if n.startswith("random") else SCENES[n](outdir)
            with open(p + ".tmp", "w") as f:
                f.write(text)
            os.replace(p + ".tmp", p)
        paths[n] = p
    return paths


if __name__ == "__main__":
    if len(sys.argv) < 2:
        print(__doc__)
        sys.exit(64)
    for k, v in generate(sys.argv[1], sys.argv[2:] or None).items():
        print(k, v)
