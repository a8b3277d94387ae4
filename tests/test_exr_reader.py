"""the .exr texture reader of the host side (host/fs2_exr.cpp) against
files written by scenes/exr_write.py from the published OpenEXR layout.  The reference reads .exr
through tinyexr's LoadEXR (source/material.cpp:70-92, source/third_party/tiny_exr_loader.h);
tests/test_exr_reference.py pins the reader against the texel arrays that loader produced
(tests/golden/exr_textures.npz).  This file covers the format itself: values must come back
exactly (HALF -> float conversion is exact), rows top to bottom, R / G / B picked by name."""
import ctypes as C
import os

import numpy as np
import pytest

from exr_write import write_exr
from fs2_b200 import host


def _load(path):
    lib = host.load_host_library()
    lib.fs2h_load_image.restype = C.c_int
    lib.fs2h_load_image.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                    C.c_void_p, C.c_size_t]
    w, h = C.c_uint32(), C.c_uint32()
    buf = np.zeros(1 << 20, np.float32)
    rc = lib.fs2h_load_image(path.encode(), C.byref(w), C.byref(h), buf.ctypes.data_as(C.c_void_p), buf.size)
    if rc != 0:
        return rc, None
    return 0, buf[: w.value * h.value * 3].reshape(h.value, w.value, 3).copy()


def _image(h, w, dtype, seed):
    rng = np.random.default_rng(seed)
    img = (rng.random((h, w, 3)) * 4.0).astype(dtype)
    img[0, 0] = (0.0, 1.0, 65504.0 if dtype == np.float16 else 1e30)
    img[h // 2, : w // 2] = 0.25  # runs, for RLE
    return img


@pytest.mark.parametrize("comp", ["none", "rle", "zips", "zip"])
@pytest.mark.parametrize("dtype", [np.float16, np.float32])
def test_exr_round_trip(tmp_path, comp, dtype):
    h, w = 37, 29  # not a multiple of ZIP's 16-line blocks
    img = _image(h, w, dtype, 5)
    path = str(tmp_path / "t.exr")
    write_exr(path, {"R": img[..., 0], "G": img[..., 1], "B": img[..., 2],
                     "A": np.ones((h, w), dtype)}, comp=comp)
    rc, got = _load(path)
    assert rc == 0
    assert got.shape == (h, w, 3)
    assert np.array_equal(got, img.astype(np.float32))


def test_exr_decreasing_line_order_and_window_origin(tmp_path):
    h, w = 20, 13
    img = _image(h, w, np.float16, 9)
    path = str(tmp_path / "d.exr")
    write_exr(path, {"B": img[..., 2], "G": img[..., 1], "R": img[..., 0]}, comp="zip",
              origin=(-7, 100), decreasing=True)
    rc, got = _load(path)
    assert rc == 0 and np.array_equal(got, img.astype(np.float32))


def test_exr_single_channel_becomes_grey_and_mixed_types(tmp_path):
    y = _image(9, 11, np.float32, 2)[..., 0]
    path = str(tmp_path / "y.exr")
    write_exr(path, {"Y": y}, comp="zips")
    rc, got = _load(path)
    assert rc == 0 and all(np.array_equal(got[..., c], y) for c in range(3))
    img = _image(9, 11, np.float32, 3)
    write_exr(path, {"R": img[..., 0].astype(np.float16), "G": img[..., 1], "B": img[..., 2],
                     "Z": img[..., 0]}, comp="rle")
    rc, got = _load(path)
    assert rc == 0
    assert np.array_equal(got[..., 0], img[..., 0].astype(np.float16).astype(np.float32))
    assert np.array_equal(got[..., 1:], img[..., 1:])


def test_exr_subnormal_and_special_halves(tmp_path):
    bits = np.array([[0x0001, 0x03FF, 0x0400, 0x8001, 0x7C00, 0xFC00, 0x3C00, 0x0000]], np.uint16)
    half = bits.view(np.float16)
    path = str(tmp_path / "s.exr")
    write_exr(path, {"R": half, "G": half, "B": half})
    rc, got = _load(path)
    assert rc == 0
    assert np.array_equal(got[..., 0].view(np.uint32), half.astype(np.float32).view(np.uint32))


def test_exr_rejects_damaged_files(tmp_path):
    img = _image(16, 16, np.float16, 1)
    path = str(tmp_path / "ok.exr")
    write_exr(path, {"R": img[..., 0], "G": img[..., 1], "B": img[..., 2]}, comp="zip")
    raw = open(path, "rb").read()
    for name, data in (("short", raw[: len(raw) // 2]), ("magic", b"\0" + raw[1:]),
                       ("tiled", raw[:5] + b"\x02" + raw[6:]), ("empty", b"")):
        p = str(tmp_path / (name + ".exr"))
        open(p, "wb").write(data)
        rc, _ = _load(p)
        assert rc == -1, name
    assert b"fs2h_load_image" in host.load_host_library().fs2h_last_error()
    rc, _ = _load(str(tmp_path / "missing.exr"))
    assert rc == -1


def test_scene_with_exr_texture_loads(tmp_path):
    """A material block with an .exr texture ends up in the flattened scene's texels."""
    img = _image(8, 8, np.float16, 4)
    write_exr(str(tmp_path / "tex.exr"), {"R": img[..., 0], "G": img[..., 1], "B": img[..., 2]}, comp="zip")
    scene = tmp_path / "s.scene"
    scene.write_text(
        "material m\n{\n color 1 1 1\n texture_scale 2\n texture %s\n}\n"
        "sphere\n{\n material m\n position 0 0 0\n radius 1\n}\n" % (tmp_path / "tex.exr"))
    hs = host.HostScene(str(scene))
    flat = hs.to_flat()
    tex = np.asarray(flat.texels, np.float32).reshape(-1)
    want = img.astype(np.float32).reshape(-1)
    assert tex.size >= want.size
    assert any(np.array_equal(tex[i:i + want.size], want) for i in range(0, tex.size - want.size + 1, 3))


def test_exr_writer_round_trips_through_the_reader(tmp_path):
    """save_exr (the CLI's --exr output: FLOAT B/G/R, ZIP, film rows bottom-up) read back by
    load_exr: exact values, top row first."""
    lib = host.load_host_library()
    lib.fs2h_save_exr.restype = C.c_int
    lib.fs2h_save_exr.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]
    for h, w, seed in ((37, 29, 1), (16, 8, 2), (1, 1, 3)):
        img = _image(h, w, np.float32, seed) if h > 1 else np.array([[[0.5, 2.0, -1.0]]], np.float32)
        if h > 1:
            img[3:30, :, :] = 0.125  # compressible blocks next to random (stored raw) ones
        path = str(tmp_path / ("w%d.exr" % seed))
        for bottom_up in (0, 1):
            assert lib.fs2h_save_exr(path.encode(), w, h, img.ctypes.data_as(C.c_void_p), bottom_up) == 0
            rc, got = _load(path)
            assert rc == 0
            assert np.array_equal(got, img[::-1] if bottom_up else img)
    assert lib.fs2h_save_exr(str(tmp_path / "no" / "dir.exr").encode(), 2, 2,
                             np.zeros(12, np.float32).ctypes.data_as(C.c_void_p), 0) != 0


def test_exr_against_an_independent_codec(tmp_path):
    """Cross-check with the OpenEXR library inside OpenCV (when the image has it): files of
    scenes/exr_write.py and of save_exr read by OpenCV, OpenCV's own file read by load_exr."""
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    cv2 = pytest.importorskip("cv2")
    img = _image(37, 29, np.float32, 11)
    img[0, 0] = (0.0, 1.0, 1e30)
    p = str(tmp_path / "x.exr")
    if not cv2.imwrite(p, img[..., ::-1].copy()):
        pytest.skip("this OpenCV build has no OpenEXR codec")
    rc, got = _load(p)  # OpenCV writer -> our reader
    assert rc == 0 and np.array_equal(got, img)
    for comp in ("none", "rle", "zips", "zip"):  # test writer -> OpenCV reader
        for dt in (np.float16, np.float32):
            im = img.clip(0, 60000).astype(dt)
            write_exr(p, {"R": im[..., 0], "G": im[..., 1], "B": im[..., 2]}, comp=comp)
            ref = cv2.imread(p, cv2.IMREAD_UNCHANGED)
            assert ref is not None and np.array_equal(ref[..., ::-1].astype(np.float32), im.astype(np.float32))
    lib = host.load_host_library()  # our writer -> OpenCV reader
    lib.fs2h_save_exr.restype = C.c_int
    lib.fs2h_save_exr.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int]
    assert lib.fs2h_save_exr(p.encode(), 29, 37, img.ctypes.data_as(C.c_void_p), 0) == 0
    ref = cv2.imread(p, cv2.IMREAD_UNCHANGED)
    assert ref is not None and np.array_equal(ref[..., ::-1], img)


def test_exr_reader_survives_corrupted_files(tmp_path):
    """Random byte damage anywhere in a valid file: the reader answers (image or error), it
    never crashes or over-allocates."""
    img = _image(24, 17, np.float16, 6)
    good = str(tmp_path / "g.exr")
    rng = np.random.default_rng(123)
    for comp in ("none", "rle", "zip"):
        write_exr(good, {"R": img[..., 0], "G": img[..., 1], "B": img[..., 2]}, comp=comp)
        raw = bytearray(open(good, "rb").read())
        for trial in range(150):
            bad = bytearray(raw)
            for _ in range(int(rng.integers(1, 6))):
                bad[int(rng.integers(0, len(bad)))] = int(rng.integers(0, 256))
            if trial % 10 == 0:
                bad = bad[: int(rng.integers(0, len(bad)))]
            p = str(tmp_path / "bad.exr")
            open(p, "wb").write(bytes(bad))
            rc, got = _load(p)
            assert rc in (0, -1, -2)
            if rc == 0:
                assert got.ndim == 3 and got.shape[2] == 3


def test_readers_refuse_a_directory(tmp_path):
    """fopen() succeeds on a directory and ftell() reports LONG_MAX: the readers must answer
    'unreadable' instead of sizing a buffer from it (found by fuzzing the .scene parser: a stray
    carriage return cut a mesh file name down to its directory)."""
    for ext in (".exr", ".bmp", ".hdr"):
        d = tmp_path / ("dir" + ext)
        d.mkdir()
        rc, _ = _load(str(d))
        assert rc == -1
    scene = tmp_path / "s.scene"
    scene.write_text("material m\n{\n color 1 1 1\n}\nmesh\n{\n file %s/\rgone.obj\n material m\n}\n"
                     "sphere\n{\n material m\n position 0 0 0\n radius 1\n}\n" % tmp_path)
    hs = host.HostScene(str(scene))  # the mesh fails to load, the scene still does
    assert hs.desc().n_objects >= 1


def test_hdr_header_cannot_size_an_allocation(tmp_path):
    """A 40-byte .hdr that claims 100 000 x 100 000 texels is refused before anything is
    allocated (it used to reach vector::assign and unwind through the C surface), and a real
    small file still loads."""
    big = tmp_path / "huge.hdr"
    big.write_bytes(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y 100000 +X 100000\n" + b"\x00" * 8)
    rc, _ = _load(str(big))
    assert rc == -1
    tall = tmp_path / "tall.hdr"  # plausible width, more lines than the file has bytes for
    tall.write_bytes(b"#?RADIANCE\n\n-Y 5000 +X 16\n" + b"\x80" * 64)
    rc, _ = _load(str(tall))
    assert rc == -1
    ok = tmp_path / "ok.hdr"
    ok.write_bytes(b"#?RADIANCE\n\n-Y 2 +X 3\n" + bytes([128, 64, 32, 129] * 6))
    rc, img = _load(str(ok))
    assert rc == 0 and img.shape == (2, 3, 3) and np.allclose(img[0, 0], [1.0, 0.5, 0.25])
