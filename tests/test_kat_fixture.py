"""The known-answer fixture tests/golden/kat_words.txt (output of oracle/ref/kat_probe.cpp linked
against the unmodified reference) carries the values SURVEY.md Appendix C lists; when the probe
binary is present it still prints the fixture."""
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def kat_words():
    with open(os.path.join(ROOT, "tests", "golden", "kat_words.txt")) as f:
        return np.array([int(w, 16) for w in f.read().split()], np.uint32)


def test_fixture_holds_the_survey_values():
    w = kat_words()
    assert len(w) == 48
    f = w.view(np.float32)
    assert w[0] == 0x3F34F95E and w[1] == 0x3FB51089            # inv_sqrt(2), base::sqrtf(2)
    assert f[2] == np.float32(5.00776815)                        # |(3,4,0)|
    assert w[3] == 0x3E3322F1 and w[7] == 0x3BB91B4E             # plane y=0, triangle t
    assert w[8] == 0x3D6CB903 and w[9] == 0x3E319C6E             # barycentric u, v
    assert w[13] == 0x3B9001E8 and w[20] == 0x3B9D0599           # sphere, aabb
    assert list(w[24:27]) == [0x3E994F8C, 0xBF4C6A10, 0x3F04DE8A]  # normalize
    assert list(w[30:33]) == [0x3E65BDE1, 0xBF6443AF, 0x3EC71C06]  # refract 0.75
    assert list(w[36:38]) == [0x3F1551FE, 0x3F663508]            # sphere map
    assert (w[38:41] == 0).all()                                 # total internal reflection
    ints = [int(w[41 + 2 * i]) | int(w[42 + 2 * i]) << 32 for i in range(3)]
    assert ints == [17856780258404322015, 14185171587269498017, 264782984789228142]
    assert [int(w[47]) >> s & 255 for s in (0, 8, 16)] == [148, 148, 237]


def test_reference_probe_still_prints_the_fixture():
    exe = os.path.join(ROOT, "oracle", "_ref", "kat_probe")
    if not os.path.exists(exe):
        import pytest
        pytest.skip("oracle/_ref/kat_probe not built (needs /root/reference)")
    out = subprocess.check_output([exe], text=True).split()
    assert [int(x, 16) for x in out] == [int(x) for x in kat_words()]
