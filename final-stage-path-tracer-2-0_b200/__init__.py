"""fs2_b200 — B200-native per-sample path-tracing hot path of Final Stage 2.0.

The directory name is the project's (`final-stage-path-tracer-2-0_b200`), which is not
a Python identifier; import it through the repo-root shim `fs2_b200.py`:

    import fs2_b200
    ctx = fs2_b200.Context(device=0)

Layout: csrc/ (hand-written sm_100a CUDA kernels + the C ABI of include/fs2_cabi.h),
host/ (C++ mirror of the reference's Scene/Camera/Material/DisplayFrame/TraceScene API
that flattens scenes for the C ABI), abi.py / engine.py (ctypes plumbing), build.py.
The product never imports anything under oracle/.
"""
from .abi import *  # noqa: F401,F403
from .abi import FlatScene, fs_camera, fs_scene_desc, fs_stats, make_camera  # noqa: F401
from . import build  # noqa: F401


def __getattr__(name):
    # engine/host need the compiled libraries; import lazily so that CPU-only tooling
    # (blob io, builders) works before build() has run.
    if name in ("Context", "Fs2Error", "load_library"):
        from . import engine
        return getattr(engine, name)
    if name in ("HostScene", "load_host_library"):
        from . import host
        return getattr(host, name)
    raise AttributeError(name)
