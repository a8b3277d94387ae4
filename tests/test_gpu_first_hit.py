"""CUDA first-hit path (fs_first_hit through the C ABI) against the golden vectors of the
unmodified reference and against the CPU oracle: bit-exact parameters, ranges and IDs."""
import numpy as np
import pytest

from oracle import binding
from refdata import GOLDEN_SCENES, bits, golden, local_face

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", GOLDEN_SCENES)
def test_first_hit_matches_reference_golden(ctx, name):
    g = golden(name)
    ctx.upload(g.scene)
    o = ctx.first_hit(g.cam, g.w, g.h)
    fh = g.first_hit
    hit = fh["hit"] == 1
    # stated tolerance: none — every pixel must agree bit for bit (0 grazing pixels allowed)
    assert int((o["object"] != fh["object"]).sum()) == 0
    assert int((bits(o["t"]) != bits(fh["t"])).sum()) == 0
    assert int((bits(o["range"]) != bits(fh["range"])).sum()) == 0
    assert (o["mat_id"] == fh["mat_id"]).all()
    lf = local_face(g.scene, o["object"], o["face"])
    assert (lf[hit] == fh["face"][hit]).all()


@pytest.mark.parametrize("name,w,h", [("mesh_tiny", 333, 187), ("many303", 257, 129),
                                      ("demo", 512, 512), ("grid_tiny", 640, 360)])
def test_first_hit_matches_oracle_other_sizes(ctx, name, w, h):
    g = golden(name)
    ctx.upload(g.scene)
    ctx.reset_stats()
    o = ctx.first_hit(g.cam, w, h)
    ref = binding.first_hit(g.scene, g.cam, w, h)
    assert (o["object"] == ref["object"]).all()
    assert (bits(o["t"]) == bits(ref["t"])).all()
    assert (o["face"] == ref["face"]).all()
    assert (bits(o["range"]) == bits(ref["range"])).all()
    assert (o["mat_id"] == ref["mat_id"]).all()
    # work counters of the stats build equal the oracle's, test for test
    st = ctx.stats()
    c = ref["counters"]
    assert st["n_segments"] == c["n_segments"] == w * h
    assert st["n_node"] == c["n_node"]
    assert st["n_tri"] == c["n_tri"]
    assert st["n_prim"] == c["n_prim"]


def test_trace_range_single_ray(ctx):
    g = golden("demo")
    ctx.upload(g.scene)
    ref = binding.first_hit(g.scene, g.cam, g.w, g.h)["range"].reshape(g.h, g.w)
    for (x, y) in [(0, 0), (17, 40), (95, 63), (48, 5)]:
        r = ctx.trace_range(g.cam, g.w, g.h, float(x), float(y))
        assert np.float32(r).view(np.uint32) == ref[y, x].view(np.uint32)


def test_error_paths(ctx):
    import fs2_b200
    c2 = fs2_b200.Context(0)
    g = golden("demo")
    with pytest.raises(fs2_b200.Fs2Error) as e:
        c2.first_hit(g.cam, 8, 8)
    assert e.value.code == -3  # FS_E_NOSCENE
    c2.upload(g.scene)
    with pytest.raises(fs2_b200.Fs2Error) as e:
        c2.trace(g.cam, 1)
    assert e.value.code == -4  # FS_E_NOFILM
    bad = golden("demo").scene
    d = bad.desc()
    d.sky_material = 999
    with pytest.raises(fs2_b200.Fs2Error) as e:
        c2.upload(d)
    assert e.value.code == -1
    c2.close()


def test_malformed_trees_are_rejected_at_upload(ctx):
    """fs_scene_upload walks every octree once: leaf ranges outside the record arrays, child
    links that do not follow their parent (cycles), shared subtrees and trees deeper than the
    kernels' frame stacks return FS_E_INVALID instead of reading out of bounds or silently
    dropping geometry."""
    import fs2_b200
    from fs2_b200 import abi
    INTERNAL = 0xFFFFFFFF

    def upload_broken(name, mutate):
        g = golden(name)
        arrays = {n: getattr(g.scene, n) for n, _ in abi._ARRAYS}
        arrays["nodes"] = mutate(g.scene.nodes.copy(), g.scene)
        bad = abi.FlatScene(scene_root=g.scene.scene_root, sky_material=g.scene.sky_material, **arrays)
        with pytest.raises(fs2_b200.Fs2Error) as e:
            ctx.upload(bad)
        assert e.value.code == -1, name
        return str(e.value)

    def mesh_root(sc):
        return int(sc.objects["u"][sc.objects["type"] == abi.FS_OBJ_MESH][0, 0])

    def walk(nodes, root):
        """(node, depth, parent) in depth-first order."""
        todo = [(root, 0, -1)]
        while todo:
            i, d, p = todo.pop()
            yield i, d, p
            if nodes["count"][i] == INTERNAL:
                todo += [(int(nodes["first"][i]) + c, d + 1, i) for c in range(8)]

    def leaf_overrun(nodes, sc):
        i = next(i for i, _, _ in walk(nodes, mesh_root(sc)) if 0 < nodes["count"][i] != INTERNAL)
        nodes["count"][i] = len(sc.tri_aux) + 5
        return nodes
    assert "leaf range" in upload_broken("mesh_tiny", leaf_overrun)

    def cycle(nodes, sc):
        r = mesh_root(sc)
        assert nodes["count"][r] == INTERNAL
        nodes["first"][r] = max(0, r - 7)  # the root becomes its own descendant
        return nodes
    assert "follow their parent" in upload_broken("mesh_tiny", cycle)

    def shared_subtree(nodes, sc):
        for i, _, p in walk(nodes, mesh_root(sc)):
            if p >= 0 and nodes["count"][i] == INTERNAL:
                first = int(nodes["first"][p])
                sibling = first + ((i - first + 1) & 7)
                nodes[sibling] = nodes[i]
                return nodes
        raise AssertionError("no internal child")
    assert "reachable twice" in upload_broken("mesh_tiny", shared_subtree)

    def too_deep(nodes, sc):
        # hang fresh blocks of eight empty leaves below the deepest leaf until an internal node
        # sits at depth 4 (mesh trees: internal nodes at depth 0..3, reference mesh.cpp:15-16)
        i, depth = max(((i, d) for i, d, _ in walk(nodes, mesh_root(sc)) if nodes["count"][i] != INTERNAL),
                       key=lambda t: t[1])
        while depth <= 4:
            block = np.zeros(8, nodes.dtype)
            nodes["count"][i] = INTERNAL
            nodes["first"][i] = len(nodes)
            nodes = np.concatenate([nodes, block])
            i, depth = len(nodes) - 8, depth + 1
        return nodes
    assert "deeper" in upload_broken("mesh_tiny", too_deep)

    def scene_leaf_overrun(nodes, sc):
        i = next(i for i, _, _ in walk(nodes, int(sc.scene_root)) if 0 < nodes["count"][i] != INTERNAL)
        nodes["first"][i] = len(sc.scene_refs)
        return nodes
    assert "leaf range" in upload_broken("many303", scene_leaf_overrun)
    # and the unbroken scenes still upload
    ctx.upload(golden("mesh_tiny").scene)
    ctx.upload(golden("many303").scene)
