"""the device functions the kernels are built from, evaluated by one
device thread (fs_device_selftest, csrc/fs2_selftest.cuh), against the bits the unmodified
reference produces for the same probes (tests/golden/kat_words.txt).  Parameters, barycentrics,
vector operations, the random stream and the film quantiser must match bit for bit; hit points
(start + dir * t, printed by the reference from its own collision record) within 2 ulp."""
import ctypes as C

import numpy as np
import pytest

from fs2_b200 import engine
from test_kat_fixture import kat_words

pytestmark = pytest.mark.gpu

EXACT = [0, 1, 2, 3, 7, 8, 9, 10, 11, 12, 13, 20] + list(range(24, 48))
POINTS = [4, 5, 6, 14, 15, 16, 17, 18, 19, 21, 22, 23]


def test_device_functions_reproduce_reference_bits():
    lib = engine.load_library()
    got = np.zeros(48, np.uint32)
    rc = lib.fs_device_selftest(0, got.ctypes.data_as(C.c_void_p), 48)
    assert rc == 0, lib.fs_last_error(None)
    ref = kat_words()
    bad = [(i, hex(int(got[i])), hex(int(ref[i]))) for i in EXACT if got[i] != ref[i]]
    assert not bad, "word, device, reference: %r" % bad
    gf, rf = got.view(np.float32), ref.view(np.float32)
    for i in POINTS:
        assert abs(float(gf[i]) - float(rf[i])) <= 2 * np.spacing(np.abs(rf[i])) + 1e-6, (i, gf[i], rf[i])


def test_selftest_rejects_short_buffer():
    lib = engine.load_library()
    got = np.zeros(8, np.uint32)
    assert lib.fs_device_selftest(0, got.ctypes.data_as(C.c_void_p), 8) != 0
