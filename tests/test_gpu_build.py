"""the per-mesh octree built on the GPU (csrc/fs2_build.cuh, C ABI
fs_mesh_tree_build) must equal the host builder's output byte for byte — same nodes, same
leaf membership and order, same packed triangle records.  The host builder itself is pinned
against the reference's flattening in tests/test_host_flatten.py."""
import time

import numpy as np
import pytest

import gen_scenes
from fs2_b200 import host

pytestmark = pytest.mark.gpu


def _flat(path, device):
    lib = host.load_host_library()
    lib.fs2h_set_gpu_build(device)
    try:
        t0 = time.time()
        hs = host.HostScene(path)
        flat = hs.to_flat()
        return flat, time.time() - t0
    finally:
        lib.fs2h_set_gpu_build(-1)


@pytest.mark.parametrize("name", ["mesh_tiny", "mesh_small", "grid_small", "mesh100k"])
def test_gpu_built_octree_is_byte_identical(name, tmp_path_factory):
    d = str(tmp_path_factory.mktemp("build"))
    path = gen_scenes.generate(d, [name])[name]
    lib = host.load_host_library()
    ms0 = lib.fs2h_gpu_build_ms()
    cpu, t_cpu = _flat(path, -1)
    gpu, t_gpu = _flat(path, 0)
    assert lib.fs2h_gpu_build_ms() > ms0, "the GPU builder did not run"
    assert gpu.diff(cpu) == []
    assert gpu.same_as(cpu)
    print("%s: %d nodes, %d leaf records; load+build host %.2f s, gpu %.2f s (device %.1f ms)" %
          (name, len(cpu.nodes), len(cpu.tri_aux), t_cpu, t_gpu, lib.fs2h_gpu_build_ms() - ms0))


def test_gpu_built_octree_config4(tmp_path_factory):
    """4.19 M triangles, 13.9 M leaf references."""
    d = str(tmp_path_factory.mktemp("build4m"))
    path = gen_scenes.generate(d, ["grid4m"])["grid4m"]
    lib = host.load_host_library()
    ms0 = lib.fs2h_gpu_build_ms()
    cpu, t_cpu = _flat(path, -1)
    gpu, t_gpu = _flat(path, 0)
    assert len(cpu.tri_aux) > 10_000_000
    assert np.array_equal(gpu.nodes, cpu.nodes)
    assert np.array_equal(gpu.tri_plane.view(np.uint32), cpu.tri_plane.view(np.uint32))
    assert gpu.tri_aux.tobytes() == cpu.tri_aux.tobytes()
    print("grid4m: load+build host %.2f s, gpu %.2f s (device %.1f ms)" %
          (t_cpu, t_gpu, lib.fs2h_gpu_build_ms() - ms0))
