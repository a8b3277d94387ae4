"""Regenerates tests/golden/kat_words.txt from the UNMODIFIED reference's math: builds
oracle/_ref/kat_probe (oracle/ref/Makefile, needs /root/reference) and stores its output — the
48 words of SURVEY.md Appendix C's probes in the layout of csrc/fs2_selftest.cuh."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if __name__ == "__main__":
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle", "ref")])
    out = subprocess.check_output([os.path.join(ROOT, "oracle", "_ref", "kat_probe")], text=True)
    words = out.split()
    assert len(words) == 48
    with open(os.path.join(ROOT, "tests", "golden", "kat_words.txt"), "w") as f:
        f.write("\n".join(words) + "\n")
    print("wrote %d words" % len(words))
