"""host/fs2_exr.cpp against the reference's own .exr loader (tinyexr's LoadEXR behind
DiffuseMaterial::LoadDiffuseTexture, reference source/material.cpp:70-92,
source/third_party/tiny_exr_loader.h): tests/golden/exr_textures.npz holds eight small files
(HALF / FLOAT x NONE / RLE / ZIPS / ZIP, odd sizes) and the texel arrays the unmodified reference
built from them (tests/golden/make_golden.py exr_goldens).  Dimensions and every texel bit must
agree."""
import ctypes as C
import os

import numpy as np
import pytest

from fs2_b200 import host

Z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "exr_textures.npz"))
NAMES = [str(n) for n in Z["names"]]


def _load(path):
    lib = host.load_host_library()
    lib.fs2h_load_image.restype = C.c_int
    lib.fs2h_load_image.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                    C.c_void_p, C.c_size_t]
    w, h = C.c_uint32(), C.c_uint32()
    buf = np.zeros(1 << 16, np.float32)
    rc = lib.fs2h_load_image(path.encode(), C.byref(w), C.byref(h), buf.ctypes.data_as(C.c_void_p), buf.size)
    assert rc == 0
    return w.value, h.value, buf[: w.value * h.value * 3].copy()


@pytest.mark.parametrize("name", NAMES)
def test_exr_reader_equals_reference_loader(name, tmp_path):
    path = str(tmp_path / (name + ".exr"))
    Z["file_" + name].tofile(path)
    w, h, tex = _load(path)
    assert (w, h) == tuple(int(v) for v in Z["dims_" + name])
    want = Z["texels_" + name]
    assert tex.size == want.size
    assert (tex.view(np.uint32) == want.view(np.uint32)).all()


def test_all_eight_variants_are_pinned():
    assert sorted(NAMES) == sorted("%s_%s" % (t, c) for t in ("half", "float")
                                   for c in ("none", "rle", "zips", "zip"))
