"""In-tree builds of the native libraries (no JIT cache: the .so files travel with the tree).

    libfs2cuda.so  csrc/fs2_cabi.cu        the C ABI + sm_100a kernels (nvcc)
    libfs2host.so  host/*.cpp              C++ mirror of the reference's Scene / Camera /
                                           DisplayFrame / TraceScene API (g++)

Numerics flags are part of the contract: no FMA contraction on either
side, IEEE division and square root, no fast-math.
"""
import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_DIR = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
HOST = os.path.join(PKG_DIR, "host")
CUDA_LIB = os.path.join(PKG_DIR, "libfs2cuda.so")
HOST_LIB = os.path.join(PKG_DIR, "libfs2host.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off", "--expt-relaxed-constexpr",
              "--extended-lambda", "-shared"]
GXX_FLAGS = ["-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _sources(directory, exts):
    out = []
    for root, _, files in os.walk(directory):
        out += [os.path.join(root, f) for f in files if f.endswith(exts)]
    return sorted(out)


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found; libfs2cuda.so cannot be built")
    return p


def build_cuda(force=False, verbose=False, extra=()):
    deps = _sources(CSRC, (".cu", ".cuh")) + _sources(os.path.join(REPO_DIR, "include"), (".h",))
    if not force and not _newer(CUDA_LIB, deps):
        return CUDA_LIB
    cmd = [nvcc_path()] + NVCC_FLAGS + list(extra) + ["-I" + os.path.join(REPO_DIR, "include"),
                                                      "-o", CUDA_LIB,
                                                      os.path.join(CSRC, "fs2_cabi.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.run(cmd, check=True)
    return CUDA_LIB


def build_host(force=False):
    srcs = _sources(HOST, (".cpp",))
    if not srcs:
        return None
    deps = srcs + _sources(HOST, (".h",)) + _sources(os.path.join(REPO_DIR, "include"), (".h",))
    if not force and not _newer(HOST_LIB, deps):
        return HOST_LIB
    build_cuda()  # libfs2host.so links against the C ABI library next to it
    cmd = ["g++"] + GXX_FLAGS + ["-I" + os.path.join(REPO_DIR, "include"), "-I" + HOST,
                                 "-o", HOST_LIB] + srcs + ["-L" + PKG_DIR, "-lfs2cuda",
                                                           "-Wl,-rpath,$ORIGIN", "-lpthread", "-lz"]
    subprocess.run(cmd, check=True)
    return HOST_LIB


CLI_DIR = os.path.join(PKG_DIR, "cli")
CLI_EXE = os.path.join(PKG_DIR, "fs2_render")


def build_cli(force=False):
    """fs2_render: headless replacement of the reference's main.cpp (window/OpenGL loop)."""
    src = os.path.join(CLI_DIR, "fs2_render.cpp")
    deps = [src, os.path.join(HOST, "fs2_host.h")]
    build_host()
    if not force and not _newer(CLI_EXE, deps + [HOST_LIB]):
        return CLI_EXE
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(REPO_DIR, "include"), "-I" + HOST,
           src, "-o", CLI_EXE, "-L" + PKG_DIR, "-lfs2host", "-lfs2cuda", "-Wl,-rpath,$ORIGIN",
           "-lpthread"]
    subprocess.run(cmd, check=True)
    return CLI_EXE


def build_all(force=False):
    return build_cuda(force), build_host(force), build_cli(force)


if __name__ == "__main__":
    import sys
    build_cuda(force=True, verbose="-v" in sys.argv)
    build_host(force=True)
    build_cli(force=True)
    print(CUDA_LIB)
