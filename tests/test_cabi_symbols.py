"""The C-ABI library loads and exports every entry point include/fs2_cabi.h declares.
No compute call is made here (there is no GPU in the CPU suite); without a device the
library must fail loudly instead of falling back to anything."""
import ctypes as C
import os
import re

import fs2_b200
from fs2_b200 import abi, engine

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(REPO, "include", "fs2_cabi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"^\s*(?:int|void|const char\s*\*)\s*(fs_[a-z_]+)\s*\(", text, re.M)))


def test_every_declared_symbol_is_exported():
    lib = engine.load_library()
    names = header_functions()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(engine.exported_symbols())
    assert lib.fs_abi_version() == abi.FS2_ABI_VERSION


def test_struct_sizes_match_header():
    assert C.sizeof(abi.fs_camera) == 48
    assert C.sizeof(abi.fs_stats) == 88
    assert C.sizeof(abi.fs_scene_desc) == 56 + 10 * 8
    assert abi.NODE_DT.itemsize == 32 and abi.OBJECT_DT.itemsize == 144
    assert abi.RAY_HIT_DT.itemsize == 52  # fs_ray_hit


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        return  # covered by the gpu suite
    lib = engine.load_library()
    h = C.c_void_p()
    rc = lib.fs_create(C.byref(h), 0, 0)
    assert rc == -2 and not h.value  # FS_E_CUDA, no context, no CPU fallback
    assert b"CUDA" in lib.fs_last_error(None) or b"device" in lib.fs_last_error(None)
    try:
        fs2_b200.Context(0)
    except fs2_b200.Fs2Error as e:
        assert e.code == -2
    else:
        raise AssertionError("Context() must raise without a GPU")


def test_product_does_not_import_oracle():
    pkg = os.path.join(REPO, "final-stage-path-tracer-2-0_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(root, f)).read()
                assert "oracle" not in text.replace("oracle/", "").lower() or f == "__init__.py" \
                    or "import oracle" not in text, f
                assert "from oracle" not in text and "import oracle" not in text, f
                assert "libfs2oracle" not in text and "fs2o_" not in text, f
