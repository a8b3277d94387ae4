// fs2_phase.cuh — warp-synchronous phase machine for the path-tracing megakernel.
//
// Why: ncu on the per-lane walkers (k_trace_mega) shows 2.2-2.8 active threads per executed
// warp instruction: nested data-dependent loops (octree walk -> leaf BVH walk -> triangle
// loop) let the 32 lanes of a warp drift into different loop nests and the hardware runs
// them one after the other.  Here every lane keeps an explicit state and the warp executes
// one *phase* at a time — all lanes that are in state X run step X together, then the warp
// re-converges and moves on to the next phase:
//
//     OBJ -> OCT_ENTER -> LB_NODE -> LB_TRIS -> OCT_RETURN -> SHADE (batched)
//
// A step is small and loop-free (one octree node, one leaf-BVH node, <= 4 triangles, one
// sibling step, one material evaluation), so divergence is bounded by the number of
// distinct states instead of by trip counts.  The arithmetic of every step is the same
// device function the per-lane walkers use (fs2_traverse.cuh, fs2_lbvh.cuh), so results
// are bit-identical (tests/test_gpu_image.py compares all variants).
//
// Per-lane stacks live in shared memory, one 4-byte column per thread (conflict-free):
//   mesh-octree frames   FS2_PH_MESH_FRAMES x 5 words (node | state bits, three split-plane
//                        parameters, the entry parameter the entry point is rebuilt from)
//   leaf-BVH stack       FS2_PH_LB_STACK x 1 word: (node - leaf root) << 16 | bf16(entry t),
//                        a leaf's tree has < 65 536 nodes (fs2_lbvh_build.h) and a truncated
//                        entry parameter only makes the pop-time cull conservative
//   scene-octree frames  FS2_PH_SCENE_FRAMES x 5 words, allocated only when the scene has a
//                        tree (>= 182 objects): 152 B instead of 252 B per lane otherwise, which
//                        is what bounds the CTAs per SM (and what is left for L1)
#pragma once
#include "fs2_shade.cuh"

namespace fs2 {

#define FS2_PH_THREADS 128
#define FS2_PH_MESH_FRAMES 4   // mesh trees: internal nodes at depth 0..3 (mesh.cpp:16)
#define FS2_PH_SCENE_FRAMES 5  // scene trees: internal nodes at depth 0..4 (<= ~2 M objects)
#define FS2_PH_LB_STACK 18     // 6 inner leaf-BVH levels x 3 pending siblings: the worst case of any tree the
                               // builder makes (fs2_lbvh_build.h caps the inner levels).  Measured and dropped: a
                               // 10-entry stack (L1 124 instead of 88 KB: config 2 +0.7 %) with an exact linear-scan
                               // fallback on overflow - the fallback's call site alone cost config 2 10 % (spills)
#define FS2_PH_FRAME_WORDS 5    // node | bits << 25, tx, ty, tz, entry parameter
#define FS2_PH_NODE_BITS 25     // octree node indices below 2^25 (checked by fs_scene_upload)
#define FS2_PH_SLOTS_NOTREE (FS2_PH_MESH_FRAMES * FS2_PH_FRAME_WORDS + FS2_PH_LB_STACK)
#define FS2_PH_SLOTS (FS2_PH_SLOTS_NOTREE + FS2_PH_SCENE_FRAMES * FS2_PH_FRAME_WORDS)
#define FS2_PH_SMEM_BYTES (FS2_PH_SLOTS * FS2_PH_THREADS * 4)
// dynamic shared memory of a phase-machine kernel for a scene with / without a scene tree
static inline unsigned ph_smem_bytes(bool scene_tree) {
  return (scene_tree ? FS2_PH_SLOTS : FS2_PH_SLOTS_NOTREE) * FS2_PH_THREADS * 4u;
}

#ifndef FS2_PH_SHADE_BATCH
#define FS2_PH_SHADE_BATCH 8   // shade when this many lanes wait (or nothing else can run)
#endif
#ifndef FS2_PH_REPEAT_MIN
#define FS2_PH_REPEAT_MIN 12   // repeat a phase only while this many lanes are still in it
#endif

#ifndef FS2_PH_FUSE_DONE
#define FS2_PH_FUSE_DONE 1  // a walk that ends (root miss, last frame popped) leaves the tree in the same step
#endif

#ifndef FS2_PH_EARLY_ROOT
#define FS2_PH_EARLY_ROOT 1  // measured: config 2 709 -> 723, grid4m 606 -> 614 Mpaths/s
#endif
#ifndef FS2_PH_SPHERE_BATCH
#define FS2_PH_SPHERE_BATCH 4  // consecutive sphere references of a scene-tree leaf tested per OBJ step.  Measured on the
                               // 303-object scene with 64 phase repeats (Mpaths/s): 1 -> 375, 4 -> 423, 6 -> 419, 8 -> 411
#endif
#ifndef FS2_PH_SPHERE_FAST
#define FS2_PH_SPHERE_FAST 1  // scene-tree leaves test spheres from DScene::sref_sphere
#endif
#ifndef FS2_PH_SKIP_EMPTY
#define FS2_PH_SKIP_EMPTY 1  // children that are empty leaves are answered from the parent's occupancy byte
#endif

enum PhState {
  PH_IDLE = 0,
  PH_SEG_START,   // begin Scene::Trace for the lane's ray
  PH_OBJ,         // next object of the brute-force list / of the current scene-tree leaf
  PH_ENTER,       // "call" octree node `cur`
  PH_LB_NODE,     // visit leaf-BVH node `lb_cur`
  PH_LB_TRIS,     // test one pending group of <= 4 triangles
  PH_RETURN,      // return `ret` into the enclosing octree frame
  PH_SHADE        // segment finished: material / sky / film, next ray
};

struct PhLane {
  uint32_t st;
  Hit rec;   // active record: scene-level best, or the leaf-local record of scene.cpp:84
  Hit glob;  // scene-level best while a scene-tree leaf is processed
  uint32_t obj_i;                              // brute-force cursor
  uint32_t cur_obj;                            // mesh object being walked
  uint32_t sleaf_first, sleaf_count, sleaf_i;  // scene-tree leaf iteration
  uint32_t cur;                                // octree node to enter
  int msp, ssp;                                // frame counts of the mesh / scene walk
  uint32_t flags;                              // bit0 walking a mesh tree, bit1 ret, bit2 scene leaf improved
  // leaf BVH walk
  uint32_t lb_cur, lb_root;
  int lb_sp;
  uint32_t qbase, qcount, qmask;  // pending triangle groups of the current node; bit31 of qcount: found
};

#define PH_F_MESH 1u
#define PH_F_RET 2u
#define PH_F_SIMPROVED 4u
#define PH_F_FOUND 8u
#define PH_F_FAR 16u  // segment starts beyond DScene::lb_origin_limit: linear mesh-leaf scans

// ---- shared-memory columns
#ifndef FS2_WF_STAGE_TOP
#define FS2_WF_STAGE_TOP 0  // experiment: root + first level of the first 8 objects' mesh trees in shared memory.
                            // Measured: config 2 657 vs 720 Mpaths/s, grid4m 565 vs 612 — off
#endif
#define FS2_STAGE_OBJECTS 8
#define FS2_STAGE_NODES 9   // a root and its eight children (numbered depth first: contiguous)

struct PhSmem {
  uint32_t* col;  // &smem[threadIdx.x]
  const float4* top = nullptr;         // [FS2_STAGE_OBJECTS][2 * FS2_STAGE_NODES] staged node words
  const uint32_t* top_root = nullptr;  // [FS2_STAGE_OBJECTS] root node index, 0xFFFFFFFF = not staged
  __device__ __forceinline__ uint32_t& w(int slot) const { return col[slot * FS2_PH_THREADS]; }
  __device__ __forceinline__ float& f(int slot) const {
    return reinterpret_cast<float*>(col)[slot * FS2_PH_THREADS];
  }
  // word i of frame sp: mesh frames first, then the leaf-BVH stack, scene frames last
  __device__ __forceinline__ int fslot(bool mesh, int sp, int i) const {
    return (mesh ? 0 : FS2_PH_SLOTS_NOTREE) + sp * FS2_PH_FRAME_WORDS + i;
  }
  __device__ __forceinline__ int lslot(int sp) const {
    return FS2_PH_MESH_FRAMES * FS2_PH_FRAME_WORDS + sp;
  }
};

// Node record of a mesh-tree walk: from the staged copy when the node is a staged root or one of
// its children, from global memory otherwise.
__device__ __forceinline__ NodeRec ph_load_node(const DScene& S, const PhSmem& sm, uint32_t obj, bool mesh,
                                                uint32_t i) {
#if FS2_WF_STAGE_TOP
  if (mesh && obj < FS2_STAGE_OBJECTS && sm.top) {  // (only k_wf_trace stages)
    const uint32_t rel = i - sm.top_root[obj];
    if (rel < FS2_STAGE_NODES) {
      const float4 a = sm.top[obj * (2 * FS2_STAGE_NODES) + 2 * rel], b = sm.top[obj * (2 * FS2_STAGE_NODES) + 2 * rel + 1];
      NodeRec n;
      n.mn = v3(a.x, a.y, a.z);
      n.mx = v3(b.x, b.y, b.z);
      n.count = __float_as_uint(b.w);
      const uint32_t f = __float_as_uint(a.w);
      const bool internal = n.count == FS_NODE_INTERNAL;
      n.first = internal ? (f & ((1u << FS2_NODE_INDEX_BITS) - 1u)) : f;
      n.cmask = internal ? (f >> FS2_NODE_INDEX_BITS) : 0u;
      return n;
    }
  }
#endif
  return load_node(S, i);
}

// frame words: 0 node | bits << 25 (bits: 0-2 child, 3-4 visits, 5 returned a hit, 6 origin
// inside), 1..3 tx ty tz, 4 entry parameter

// Begin Scene::Trace for the lane's next ray (scene.cpp:227-237).
// (TREE = false in the ph_* steps: the caller knows the scene has no scene octree, so the
// scene-leaf bookkeeping folds away at compile time.)
template <bool TREE = true>
__device__ __forceinline__ LbRay ph_segment_start(const DScene& S, PhLane& w, const Ray& r) {
  w.rec = no_hit();
  w.flags = lb_origin_far(S, r.start) ? PH_F_FAR : 0u;
  w.ssp = 0;
  if (!TREE || S.scene_root < 0) {
    w.obj_i = 0;
    w.st = PH_OBJ;
  } else {
    w.cur = (uint32_t)S.scene_root;
    w.st = PH_ENTER;
  }
  return make_lbray(r);
}

// ------------------------------------------------------------------ octree steps
// The active tree's walk ended with `ret` (no frame left): Scene::Trace is complete
// (scene.cpp:234-236) or MeshObject::Trace returns into the object loop (mesh.cpp:277-322).
template <bool TREE>
__device__ __forceinline__ void ph_tree_done(const DScene& S, PhLane& w, bool mesh, bool ret) {
  if (!mesh) {
    w.st = PH_SHADE;
    return;
  }
  w.flags &= ~PH_F_MESH;
  if (TREE && S.scene_root >= 0 && ret && w.rec.param < w.glob.param) {
    w.glob = w.rec;
    w.flags |= PH_F_SIMPROVED;
  }
  w.st = PH_OBJ;
}

// Node w.cur answered "no hit" before anything was pushed for it: return false to the enclosing
// frame, or — at the root — end the walk in the same step (a ray that misses a mesh's root box
// is the common case: one step instead of ENTER + RETURN).
template <bool TREE>
__device__ __forceinline__ void ph_enter_miss(const DScene& S, PhLane& w, bool mesh) {
  w.flags &= ~PH_F_RET;
  if (FS2_PH_FUSE_DONE && (mesh ? w.msp : w.ssp) == 0)
    ph_tree_done<TREE>(S, w, mesh, false);
  else
    w.st = PH_RETURN;
}

// "call" node w.cur of the active tree (reference mesh.cpp:78-84 / scene.cpp:76-83, then
// bvh.h:160-186).  Same arithmetic as walk_octree().
template <bool STATS, bool TREE = true>
__device__ __forceinline__ void ph_enter(const DScene& S, PhLane& w, const Ray& r,
                                         const PhSmem& sm, Counters& c) {
  const bool mesh = !TREE || (w.flags & PH_F_MESH) != 0;  // without a scene tree only meshes walk
  NodeRec n = ph_load_node(S, sm, w.cur_obj, mesh, w.cur);
#if FS2_PH_EARLY_ROOT
  // the leaf-BVH root of a mesh leaf, fetched beside the node instead of after its box test
  // (one memory round trip less per leaf visit; wasted for internal nodes and misses)
  const uint32_t early_root = mesh ? __ldg(&S.lb_leaf[w.cur].x) : 0u;
#endif
  float nparam;
  V3 entry = r.start;
  bool inside;
  if (STATS) c.n_node++;
  if (!ray_box(n.mn, n.mx, r, nparam, entry, inside) || nparam > w.rec.param) {
    ph_enter_miss<TREE>(S, w, mesh);
    return;
  }
  if (n.count != FS_NODE_INTERNAL) {
    if (!n.count) {
      ph_enter_miss<TREE>(S, w, mesh);
    } else if (!mesh) {
      // scene.cpp:84-99: one leaf-local record, seeded with "no hit"
      w.sleaf_first = n.first;
      w.sleaf_count = n.count;
      w.sleaf_i = 0;
      w.flags &= ~PH_F_SIMPROVED;
      w.glob = w.rec;
      w.rec = no_hit();
      w.st = PH_OBJ;
    } else {
      // mesh leaf (mesh.cpp:86-111)
#if FS2_PH_EARLY_ROOT
      uint32_t root = early_root;
#else
      uint32_t root = __ldg(&S.lb_leaf[w.cur].x);
#endif
      if (r.zero_length || (w.flags & PH_F_FAR) || (root == 0xFFFFFFFFu && n.count > 16u)) {
        // rare: point query / far-away origin / leaf beyond the BVH level cap -> linear scan in
        // reference order
        ColdResult cr = mesh_leaf_cold<STATS>(S, n.first, n.count, r, w.cur_obj, w.rec, c);
        w.rec = cr.h;
        c = cr.c;
        w.flags = cr.ret ? (w.flags | PH_F_RET) : (w.flags & ~PH_F_RET);
        w.st = PH_RETURN;
      } else {
        if (STATS) c.n_tri += n.count;
        // the leaf's running best lives in w.rec itself: a record of this leaf overwrites it
        // only when it is accepted; PH_F_FOUND says that it holds one (rec.prim = its leaf rank)
        w.flags &= ~PH_F_FOUND;
        w.lb_sp = 0;
        w.lb_root = root;
        if (root == 0xFFFFFFFFu) {
          w.qbase = n.first;
          w.qcount = n.count;
          w.qmask = (1u << ((n.count + 3u) >> 2)) - 1u;
          prefetch_l1(S.lb_plane + n.first);
          w.st = PH_LB_TRIS;
        } else {
          w.lb_cur = root;
          prefetch_l1(S.lb_nodes + 8 * (size_t)root);
          w.st = PH_LB_NODE;
        }
      }
    }
    return;
  }
  int& sp = mesh ? w.msp : w.ssp;
  const int maxd = mesh ? FS2_PH_MESH_FRAMES : FS2_PH_SCENE_FRAMES;
  if (sp >= maxd) {  // deeper than any tree the host builds
    ph_enter_miss<TREE>(S, w, mesh);
    return;
  }
  V3 cen = (n.mn + n.mx) * 0.5f;
  V3 st = inside ? r.start : entry;
  uint32_t bits = (st.x - cen.x >= 0.0f ? 1u : 0u) | (st.z - cen.z >= 0.0f ? 2u : 0u) |
                  (st.y - cen.y >= 0.0f ? 4u : 0u);
  if (inside) bits |= 64u;
  sm.w(sm.fslot(mesh, sp, 0)) = w.cur | (bits << FS2_PH_NODE_BITS);
  sm.f(sm.fslot(mesh, sp, 1)) = split_param(st.x, r.dir.x, cen.x);
  sm.f(sm.fslot(mesh, sp, 2)) = split_param(st.y, r.dir.y, cen.y);
  sm.f(sm.fslot(mesh, sp, 3)) = split_param(st.z, r.dir.z, cen.z);
  sm.f(sm.fslot(mesh, sp, 4)) = nparam;  // st is rebuilt from it (outside origins only)
  sp++;
  w.cur = n.first + (bits & 7u);
  if (FS2_PH_SKIP_EMPTY && !((n.cmask >> (bits & 7u)) & 1u)) {
    // the child is an empty leaf: whatever its box test says, it returns "no hit"
    // (mesh.cpp:78-111 / scene.cpp:76-99 with an empty list) — answered here, without the visit
    if (STATS) c.n_node++;
    w.flags &= ~PH_F_RET;
    w.st = PH_RETURN;
    return;
  }
  prefetch_l1(S.nodes + 2 * (size_t)w.cur);
  // stays in PH_ENTER
}

// Return (w.flags & PH_F_RET) into the enclosing frame of the active tree (bvh.h:189-228),
// or finish the tree when no frame is left.  One frame per call.
template <bool STATS, bool TREE = true>
__device__ __forceinline__ void ph_return(const DScene& S, PhLane& w, const Ray& r,
                                          const PhSmem& sm, Counters& c) {
  const bool mesh = !TREE || (w.flags & PH_F_MESH) != 0;
  int& sp = mesh ? w.msp : w.ssp;
  bool ret = (w.flags & PH_F_RET) != 0;
  if (sp == 0) {  // (a leaf at the root: the only way here without a frame)
    ph_tree_done<TREE>(S, w, mesh, ret);
    return;
  }
  const int f = sp - 1;
  const uint32_t packed = sm.w(sm.fslot(mesh, f, 0));
  const uint32_t node = packed & ((1u << FS2_PH_NODE_BITS) - 1u);
  uint32_t bits = packed >> FS2_PH_NODE_BITS;
  NodeRec pn = ph_load_node(S, sm, w.cur_obj, mesh, node);
  if (ret) {
    bits |= 32u;
    NodeRec ch = ph_load_node(S, sm, w.cur_obj, mesh, pn.first + (bits & 7u));
    if (point_in_box(ch.mn, ch.mx, r.start + r.dir * w.rec.param)) {
      // ret stays true; continue returning into the parent frame, or finish with the last one
      if (--sp == 0 && FS2_PH_FUSE_DONE) ph_tree_done<TREE>(S, w, mesh, true);
      return;
    }
  }
  float tx = sm.f(sm.fslot(mesh, f, 1)), ty = sm.f(sm.fslot(mesh, f, 2)),
        tz = sm.f(sm.fslot(mesh, f, 3));
  uint32_t visits = ((bits >> 3) & 3u) + 1u;
  bool x_hit = tx <= 1.0f, y_hit = ty <= 1.0f, z_hit = tz <= 1.0f;
  float t = 0.f;
  uint32_t flip = 0u;
  int consumed = 0;
  if (x_hit && tx < ty && tx < tz) {
    t = tx;
    consumed = 1;
    flip = 1u;
  } else if (y_hit && ty < tz && ty < tx) {
    t = ty;
    consumed = 2;
    flip = 4u;
  } else if (z_hit && tz < ty && tz < tx) {
    t = tz;
    consumed = 3;
    flip = 2u;
  }
  bool go = flip != 0u && visits < 4u;
  if (go) {
    // the point the node was entered at: the origin, or ray_box's entry point, rebuilt with
    // ray_box's own expression (start + (stop - start) * param)
    V3 st = r.start;
    if (!(bits & 64u)) st = r.start + (r.stop - r.start) * sm.f(sm.fslot(mesh, f, 4));
    go = point_in_box(pn.mn, pn.mx, st + r.dir * t);
  }
  if (!go) {
    w.flags = (bits & 32u) ? (w.flags | PH_F_RET) : (w.flags & ~PH_F_RET);
    if (--sp == 0 && FS2_PH_FUSE_DONE) ph_tree_done<TREE>(S, w, mesh, (bits & 32u) != 0u);
    return;  // keep returning
  }
  sm.f(sm.fslot(mesh, f, consumed)) = FS2_INFINITY;
  bits = ((bits ^ flip) & ~(3u << 3)) | (visits << 3);
  sm.w(sm.fslot(mesh, f, 0)) = node | (bits << FS2_PH_NODE_BITS);
  w.cur = pn.first + (bits & 7u);
  if (FS2_PH_SKIP_EMPTY && !((pn.cmask >> (bits & 7u)) & 1u)) {
    // empty sibling: it returns "no hit" into this frame; the next step looks for the one after
    if (STATS) c.n_node++;
    w.flags &= ~PH_F_RET;
    return;  // stays in PH_RETURN
  }
  prefetch_l1(S.nodes + 2 * (size_t)w.cur);
  w.st = PH_ENTER;
}

// ------------------------------------------------------------------ object iteration
// One object of the brute-force list (scene.cpp:230-233) or of the current scene-tree leaf
// (scene.cpp:84-99).  Primitives are tested on the spot, a mesh starts its octree walk.
// Brute-force list entries per OBJ step (template parameter, chosen per scene by the host).  With
// meshes in the list one entry per step is best (config 2: 1 -> 651, 4 -> 629 Mpaths/s: lanes
// leave for their octree walks at different entries).  Without any, every lane of the warp sits in
// PH_OBJ for the whole segment and the loop overhead per entry is what is left: 16 per step gives
// config 3's scene 3 388 -> 3 713 and config 1's 4 077 -> 4 497 Mpaths/s.  (A run-time count costs
// config 2 4 %: the loop's registers.)
#define FS2_PH_OBJ_BATCH_PRIMS 16
template <bool STATS, bool TREE = true, int BATCH = 1>
__device__ __forceinline__ void ph_object(const DScene& S, PhLane& w, const Ray& r, Counters& c) {
  const bool tree = TREE && S.scene_root >= 0;
  if (!tree) {
    // a few list entries per step: primitives are tested on the spot, the end of the list is
    // noticed in the step that tested the last one
#pragma unroll 1
    for (int k = 0; k < BATCH; k++) {
      if (w.obj_i >= S.n_objects) {
        w.st = PH_SHADE;
        return;
      }
      const uint32_t oi = w.obj_i++;
      float4 w0 = obj_word(S, oi, 0);
      if (__float_as_uint(w0.x) == FS_OBJ_MESH) {
        uint32_t root = __float_as_uint(w0.z);
        if (root == 0xFFFFFFFFu) continue;
        w.cur_obj = oi;
        w.cur = root;
        prefetch_l1(S.nodes + 2 * (size_t)root);
        w.msp = 0;
        w.flags |= PH_F_MESH;
        w.st = PH_ENTER;
        return;
      }
      primitive_trace<STATS>(S, oi, __float_as_uint(w0.x), r, w.rec, c);
    }
    return;
  }
  if (w.sleaf_i >= w.sleaf_count) {
    // leaf done: hand the leaf's verdict back to the scene walk
    w.rec = w.glob;
    w.flags = (w.flags & PH_F_SIMPROVED) ? (w.flags | PH_F_RET) : (w.flags & ~PH_F_RET);
    w.st = PH_RETURN;
    return;
  }
  uint32_t k = w.sleaf_first + w.sleaf_i;
  w.sleaf_i++;
#if FS2_PH_SPHERE_FAST
  {
    // spheres (most of what a many-object scene lists) are tested from the 16-byte copy of their
    // centre and radius that sits beside the reference: one load, no dependent chain.  A step takes
    // a run of up to FS2_PH_SPHERE_BATCH consecutive sphere references (same order, same tests as
    // one per step: the leaf-local record is updated in leaf order).
    float4 sp = __ldg(S.sref_sphere + k);
    if (sp.w == sp.w) {
#pragma unroll 1
      for (int b = 0;;) {
        if (STATS) c.n_prim++;
        float t;
        if (sphere_test(v3(sp.x, sp.y, sp.z), sp.w, r, t) && t < w.rec.param) {
          w.rec.param = t;
          w.rec.object = __ldg(S.scene_refs + k);
          w.rec.prim = 0xFFFFFFFFu;
          if (w.rec.param < w.glob.param) {
            w.glob = w.rec;
            w.flags |= PH_F_SIMPROVED;
          }
        }
        if (++b >= FS2_PH_SPHERE_BATCH || w.sleaf_i >= w.sleaf_count) return;
        sp = __ldg(S.sref_sphere + k + 1);
        if (!(sp.w == sp.w)) return;  // the next reference is not a sphere: its own step
        k++;
        w.sleaf_i++;
      }
    }
  }
#endif
  const uint32_t oi = __ldg(S.scene_refs + k);
  float4 w0 = obj_word(S, oi, 0);
  if (__float_as_uint(w0.x) == FS_OBJ_MESH) {
    uint32_t root = __float_as_uint(w0.z);
    if (root != 0xFFFFFFFFu) {
      w.cur_obj = oi;
      w.cur = root;
      prefetch_l1(S.nodes + 2 * (size_t)root);
      w.msp = 0;
      w.flags |= PH_F_MESH;
      w.st = PH_ENTER;
    }
    return;
  }
  bool improved = primitive_trace<STATS>(S, oi, __float_as_uint(w0.x), r, w.rec, c);
  if (improved && w.rec.param < w.glob.param) {
    w.glob = w.rec;
    w.flags |= PH_F_SIMPROVED;
  }
}

// Warp-cooperative form of the scene-tree leaf loop (scene.cpp:84-99), called by the whole
// (converged) warp; `owners` = lanes in PH_OBJ.  The reference's tree is shallow — a leaf lists
// tens of objects — and the per-lane form tests one of them per step at 5-8 active lanes (ncu on
// the 303-object scene: profiles/r2_*_many303_*).  Here the pending objects of every owner are
// spread over G = 4 ... 32 lanes (32 / owners, a power of two), one object per lane:
//   sequential rule: object i replaces the leaf-local record when it is hit at t_i < record.t,
//   so after a run of primitives the record is the lexicographic minimum (t_i, i) over the hits
//   below the record the run started with — every lane tests its object against that starting
//   record and a G-lane min-reduction picks the survivor.  (CuboidObject::Trace's own face loop
//   has the same form and runs unchanged inside primitive_trace.)
//   A mesh ends the run: it must see the record left by the objects before it (its octree walk
//   culls with it), so a group only takes the primitives in front of its first mesh, and an
//   owner whose next object is a mesh starts that walk itself.
// slot_owner: 8 bytes of shared memory per warp.
template <bool STATS>
__device__ __forceinline__ void ph_object_coop(const DScene& S, PhLane& w, const Ray& r, Counters& c,
                                               unsigned owners, volatile uint8_t* slot_owner) {
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const unsigned below = (1u << lane) - 1u;
  const bool own = (owners >> lane) & 1u;
  if (own && w.sleaf_i >= w.sleaf_count) {
    // leaf done: hand the leaf's verdict back to the scene walk
    w.rec = w.glob;
    w.flags = (w.flags & PH_F_SIMPROVED) ? (w.flags | PH_F_RET) : (w.flags & ~PH_F_RET);
    w.st = PH_RETURN;
  }
  unsigned todo = __ballot_sync(FULL, own && w.st == PH_OBJ);
  if (!todo) return;
  // lanes per owner: 32 / owners rounded down to a power of two, at least 4
  const uint32_t n_own = (uint32_t)__popc(todo);
  const uint32_t lg = n_own <= 1u ? 5u : n_own <= 2u ? 4u : n_own <= 4u ? 3u : 2u;
  const uint32_t G = 1u << lg, slots = 32u >> lg;
  bool start_mesh = false;
  while (todo) {
    const uint32_t rank = (uint32_t)__popc(todo & below);
    const bool served = ((todo >> lane) & 1u) && rank < slots;
    if (served) slot_owner[rank] = (uint8_t)lane;
    __syncwarp();
    const uint32_t n_served = min((uint32_t)__popc(todo), slots);
    const uint32_t slot = (uint32_t)lane >> lg, j = (uint32_t)lane & (G - 1u);
    const int src = slot < n_served ? (int)slot_owner[slot] : 0;
    __syncwarp();
    const uint32_t base = __shfl_sync(FULL, w.sleaf_first + w.sleaf_i, src);
    const uint32_t left = __shfl_sync(FULL, w.sleaf_count - w.sleaf_i, src);
    Ray q;
    q.start = v3(__shfl_sync(FULL, r.start.x, src), __shfl_sync(FULL, r.start.y, src),
                 __shfl_sync(FULL, r.start.z, src));
    q.dir = v3(__shfl_sync(FULL, r.dir.x, src), __shfl_sync(FULL, r.dir.y, src),
               __shfl_sync(FULL, r.dir.z, src));
    q.stop = q.start;  // (not read by the primitive tests)
    q.zero_length = __shfl_sync(FULL, r.zero_length ? 1u : 0u, src) != 0u;
    const float bt = __shfl_sync(FULL, w.rec.param, src);
    const bool mine = slot < n_served && j < left;
    uint32_t oi = 0, type = FS_OBJ_MESH;
    if (mine) {
      oi = __ldg(S.scene_refs + base + j);
      type = __float_as_uint(obj_word(S, oi, 0).x);
    }
    // primitives in front of the group's first mesh
    const unsigned mesh_b = __ballot_sync(FULL, mine && type == FS_OBJ_MESH);
    const unsigned gmask = G == 32u ? FULL : ((1u << G) - 1u) << (slot << lg);
    const unsigned gm = mesh_b & gmask;
    const uint32_t first_mesh = gm ? (uint32_t)__ffs((int)(gm >> (slot << lg))) - 1u : G;
    Hit h;
    h.param = FS2_INFINITY;
    h.object = 0xFFFFFFFFu;
    h.prim = 0xFFFFFFFFu;
    h.u = h.v = 0.f;
    uint32_t key = 0xFFFFFFFFu;  // position in the run: ties keep the first object
    if (mine && j < first_mesh) {
      Hit t = h;
      t.param = bt;
      if (primitive_trace<STATS>(S, oi, type, q, t, c)) {
        h = t;
        key = j;
      }
    }
#pragma unroll
    for (uint32_t off = 1; off < 32u; off <<= 1) {
      float ot = __shfl_xor_sync(FULL, h.param, off);
      uint32_t ok = __shfl_xor_sync(FULL, key, off);
      uint32_t oo = __shfl_xor_sync(FULL, h.object, off), op = __shfl_xor_sync(FULL, h.prim, off);
      if (off < G && (ot < h.param || (ot == h.param && ok < key))) {
        h.param = ot;
        key = ok;
        h.object = oo;
        h.prim = op;
      }
    }
    // hand the slot's verdict back to its owner
    const int from = served ? (int)(rank << lg) : 0;
    const float rt = __shfl_sync(FULL, h.param, from);
    const uint32_t rk = __shfl_sync(FULL, key, from);
    const uint32_t ro = __shfl_sync(FULL, h.object, from), rp = __shfl_sync(FULL, h.prim, from);
    const uint32_t rfm = __shfl_sync(FULL, first_mesh, from);
    if (served) {
      const uint32_t avail = min(w.sleaf_count - w.sleaf_i, G);
      const uint32_t take = min(avail, rfm);
      w.sleaf_i += take;
      if (rk != 0xFFFFFFFFu) {
        w.rec.param = rt;
        w.rec.object = ro;
        w.rec.prim = rp;
        if (w.rec.param < w.glob.param) {
          w.glob = w.rec;
          w.flags |= PH_F_SIMPROVED;
        }
      }
      if (take == 0u) start_mesh = true;  // the next object is a mesh
    }
    todo &= ~__ballot_sync(FULL, served);
  }
  if (start_mesh) {
    // (not ph_object: a second inlined copy of the primitive tests costs 8 KB of instruction cache)
    const uint32_t oi = __ldg(S.scene_refs + w.sleaf_first + w.sleaf_i);
    w.sleaf_i++;
    const uint32_t root = __float_as_uint(obj_word(S, oi, 0).z);
    if (root != 0xFFFFFFFFu) {
      w.cur_obj = oi;
      w.cur = root;
      prefetch_l1(S.nodes + 2 * (size_t)root);
      w.msp = 0;
      w.flags |= PH_F_MESH;
      w.st = PH_ENTER;
    }
  }
}

// ------------------------------------------------------------------ leaf BVH steps
// Leaf finished: publish the best of this leaf and return to the octree walk.
__device__ __forceinline__ void ph_leaf_done(PhLane& w) {
  if (w.flags & PH_F_FOUND) {
    w.flags |= PH_F_RET;  // w.rec already is the leaf's best
  } else {
    w.flags &= ~PH_F_RET;
  }
  w.st = PH_RETURN;
}

// Next pending leaf-BVH node that can still matter, or leaf done.
__device__ __forceinline__ void ph_lb_pop(PhLane& w, const DScene& S, const PhSmem& sm) {
  while (w.lb_sp > 0) {
    --w.lb_sp;
    uint32_t e = sm.w(sm.lslot(w.lb_sp));
    if (__uint_as_float(e << 16) <= fminf(w.rec.param, 1.0f)) {
      w.lb_cur = w.lb_root + (e >> 16);
      prefetch_l1(S.lb_nodes + 8 * (size_t)w.lb_cur);
      w.st = PH_LB_NODE;
      return;
    }
  }
  ph_leaf_done(w);
}

template <bool STATS>
__device__ __forceinline__ void ph_lb_node(const DScene& S, PhLane& w, const Ray& r,
                                           const LbRay& q, const PhSmem& sm, Counters& c) {
  const float4* nd = S.lb_nodes + 8 * (size_t)w.lb_cur;
  float4 lx = ldg_keep(nd), ly = ldg_keep(nd + 1), lz = ldg_keep(nd + 2);
  float4 hx = ldg_keep(nd + 3), hy = ldg_keep(nd + 4), hz = ldg_keep(nd + 5);
  uint4 meta = __ldg((const uint4*)(nd + 6));
  if (STATS) c.n_lbnode++;
  const V3 s = r.start;
  float lim = fminf(w.rec.param, 1.0f);
  float t0 = lb_child(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, s, q, lim);
  float t1 = lb_child(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, s, q, lim);
  float t2 = lb_child(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, s, q, lim);
  float t3 = lb_child(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, s, q, lim);
  if (meta.y) {
    uint32_t mask = (t0 < 3.0f ? 1u : 0u) | (t1 < 3.0f ? 2u : 0u) | (t2 < 3.0f ? 4u : 0u) |
                    (t3 < 3.0f ? 8u : 0u);
    if (mask) {
      w.qbase = meta.x;
      w.qcount = meta.z;
      w.qmask = mask;
      prefetch_l1(S.lb_plane + meta.x + 4u * ((uint32_t)__ffs((int)mask) - 1u));
      w.st = PH_LB_TRIS;
    } else {
      ph_lb_pop(w, S, sm);
    }
    return;
  }
  uint32_t c0 = meta.x, c1 = meta.x + 1, c2 = meta.x + 2, c3 = meta.x + 3;
#define FS2_CSWAP(ta, ca, tb, cb) \
  if (tb < ta) {                  \
    float tt = ta;                \
    ta = tb;                      \
    tb = tt;                      \
    uint32_t cc = ca;             \
    ca = cb;                      \
    cb = cc;                      \
  }
  FS2_CSWAP(t0, c0, t1, c1)
  FS2_CSWAP(t2, c2, t3, c3)
  FS2_CSWAP(t0, c0, t2, c2)
  FS2_CSWAP(t1, c1, t3, c3)
  FS2_CSWAP(t1, c1, t2, c2)
#undef FS2_CSWAP
  if (t3 < 3.0f && w.lb_sp < FS2_PH_LB_STACK) {
    sm.w(sm.lslot(w.lb_sp)) = ((c3 - w.lb_root) << 16) | (__float_as_uint(t3) >> 16);
    w.lb_sp++;
  }
  if (t2 < 3.0f && w.lb_sp < FS2_PH_LB_STACK) {
    sm.w(sm.lslot(w.lb_sp)) = ((c2 - w.lb_root) << 16) | (__float_as_uint(t2) >> 16);
    w.lb_sp++;
  }
  if (t1 < 3.0f && w.lb_sp < FS2_PH_LB_STACK) {
    sm.w(sm.lslot(w.lb_sp)) = ((c1 - w.lb_root) << 16) | (__float_as_uint(t1) >> 16);
    w.lb_sp++;
  }
  if (t0 < 3.0f) {
    w.lb_cur = c0;  // stays in PH_LB_NODE
    prefetch_l1(S.lb_nodes + 8 * (size_t)c0);
    return;
  }
  ph_lb_pop(w, S, sm);
}

// Warp-cooperative form of ph_lb_node, called by the whole (converged) warp; `owners` = lanes in
// PH_LB_NODE.  ncu on the per-lane form (profiles/r2_c_*): the four child box tests run at 9 of
// 32 lanes, the sort / push code at 5.5, together 18 % of the kernel's instructions, and the six
// float4 child planes set its register peak.  Here a node's four children go to four lanes (up to
// 8 owners per pass): each lane loads one child's box (6 scalars), runs one lb_child, ranks itself
// among its three neighbours by (entry parameter, child index) with three shuffles, and the lanes
// of the far children write the owner's stack column in shared memory directly; the owner only
// takes the nearest child (or the hit mask of a quad-level node) back.  The visiting order is
// nearest first like the per-lane form's sorting network; exactness does not depend on it (the
// walk returns the lexicographic minimum (t, leaf rank) of what it reaches, and culls with <=).
template <bool STATS>
__device__ __forceinline__ void ph_lb_node_coop(const DScene& S, PhLane& w, const Ray& r,
                                                const LbRay& q, const PhSmem& sm, Counters& c,
                                                unsigned owners, volatile uint8_t* slot_owner) {
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const unsigned below = (1u << lane) - 1u;
  const bool own = (owners >> lane) & 1u;
  if (STATS && own) c.n_lbnode++;
  const float my_lim = fminf(w.rec.param, 1.0f);
  bool none = false;  // this owner's node had no child to continue with: pop
  unsigned todo = owners;
  while (todo) {
    const uint32_t rank = (uint32_t)__popc(todo & below);
    const bool served = ((todo >> lane) & 1u) && rank < 8u;
    if (served) slot_owner[rank] = (uint8_t)lane;
    __syncwarp();
    const uint32_t n_served = min((uint32_t)__popc(todo), 8u);
    const uint32_t slot = (uint32_t)lane >> 2, j = (uint32_t)lane & 3u;
    const bool active = slot < n_served;
    const int src = active ? (int)slot_owner[slot] : 0;
    __syncwarp();
    const uint32_t node = __shfl_sync(FULL, w.lb_cur, src);
    const V3 s = v3(__shfl_sync(FULL, r.start.x, src), __shfl_sync(FULL, r.start.y, src),
                    __shfl_sync(FULL, r.start.z, src));
    LbRay qq;
    qq.idx = __shfl_sync(FULL, q.idx, src);
    qq.idy = __shfl_sync(FULL, q.idy, src);
    qq.idz = __shfl_sync(FULL, q.idz, src);
    const float lim = __shfl_sync(FULL, my_lim, src);
    const uint32_t sp = (uint32_t)__shfl_sync(FULL, w.lb_sp, src);
    const uint32_t root = __shfl_sync(FULL, w.lb_root, src);
    float t = 3.0f;
    uint4 meta = make_uint4(0u, 0u, 0u, 0u);
    if (active) {
      const float* nd = reinterpret_cast<const float*>(S.lb_nodes + 8 * (size_t)node);
      meta = __ldg(reinterpret_cast<const uint4*>(nd + 24));
      t = lb_child(__ldg(nd + j), __ldg(nd + 4 + j), __ldg(nd + 8 + j), __ldg(nd + 12 + j),
                   __ldg(nd + 16 + j), __ldg(nd + 20 + j), s, qq, lim);
    }
    const bool hit = t < 3.0f;
    const unsigned hit_b = __ballot_sync(FULL, hit);
    const uint32_t hits4 = (hit_b >> (slot << 2)) & 15u;
    // children in front of this one: smaller entry parameter, or the same and a lower index
    uint32_t before = 0;
#pragma unroll
    for (uint32_t d = 1; d < 4; d++) {
      const float ot = __shfl_xor_sync(FULL, t, d);
      before += (ot < t || (ot == t && (j ^ d) < j)) ? 1u : 0u;
    }
    const unsigned first_b = __ballot_sync(FULL, hit && before == 0u);
    const uint32_t nhit = (uint32_t)__popc(hits4);
    if (active && hit && !meta.y && before >= 1u) {
      // pending sibling: the farthest lowest on the stack, so the nearest pops first
      const uint32_t pos = sp + (nhit - 1u - before);
      if (pos < (uint32_t)FS2_PH_LB_STACK)
        (sm.col + (src - lane))[sm.lslot((int)pos) * FS2_PH_THREADS] =
            ((meta.x + j - root) << 16) | (__float_as_uint(t) >> 16);
    }
    // hand the node's verdict back to its owner (read from the slot's first lane)
    const int from = served ? (int)(rank << 2) : 0;
    const uint32_t jbest = (uint32_t)__ffs((int)((first_b >> (slot << 2)) & 15u));  // 1-based, 0 = none
    const uint32_t packed = hits4 | (jbest << 4) | (meta.y ? 256u : 0u);
    const uint32_t r_packed = __shfl_sync(FULL, packed, from);
    const uint32_t r_x = __shfl_sync(FULL, meta.x, from), r_z = __shfl_sync(FULL, meta.z, from);
    if (served) {
      const uint32_t h4 = r_packed & 15u, jb = (r_packed >> 4) & 7u;
      if (!h4) {
        none = true;
      } else if (r_packed & 256u) {
        w.qbase = r_x;
        w.qcount = r_z;
        w.qmask = h4;
        prefetch_l1(S.lb_plane + r_x + 4u * ((uint32_t)__ffs((int)h4) - 1u));
        w.st = PH_LB_TRIS;
      } else {
        w.lb_cur = r_x + jb - 1u;  // stays in PH_LB_NODE
        prefetch_l1(S.lb_nodes + 8 * (size_t)w.lb_cur);
        const uint32_t nsp = (uint32_t)w.lb_sp + (uint32_t)__popc(h4) - 1u;
        w.lb_sp = (int)(nsp < (uint32_t)FS2_PH_LB_STACK ? nsp : (uint32_t)FS2_PH_LB_STACK);
      }
    }
    todo &= ~__ballot_sync(FULL, served);
  }
  __syncwarp();  // the stack columns written by other lanes are read by ph_lb_pop
  if (none) ph_lb_pop(w, S, sm);
}

template <bool STATS>
__device__ __forceinline__ void ph_lb_tris(const DScene& S, PhLane& w, const Ray& r,
                                           const PhSmem& sm, Counters& c) {
  uint32_t g = (uint32_t)__ffs((int)w.qmask) - 1u;
  w.qmask &= w.qmask - 1u;
  uint32_t rem = w.qcount - 4u * g;
  bool found = (w.flags & PH_F_FOUND) != 0;
  if (w.qmask) prefetch_l1(S.lb_plane + w.qbase + 4u * ((uint32_t)__ffs((int)w.qmask) - 1u));
  lb_test_group<STATS>(S, w.qbase + 4u * g, rem < 4u ? rem : 4u, r.start, r.dir, w.rec.param,
                       w.rec.prim, found, w.rec.u, w.rec.v, c);
  if (found) {
    w.rec.object = w.cur_obj;
    w.flags |= PH_F_FOUND;
  }
  if (!w.qmask) ph_lb_pop(w, S, sm);
}

// Warp-cooperative form of ph_lb_tris, called by the whole (converged) warp.  `owners` = lanes
// in PH_LB_TRIS.  ncu (profiles/r1_h): the per-lane form ran at 6 of 32 lanes and was 26 % of the
// traversal kernel's instructions.  Here every owner's group of <= 4 records is spread over 4
// lanes: up to 8 owners per pass, one record per lane — plane stage, acceptance filter and
// barycentric stage all run once, in parallel, and a 4-lane (t, leaf rank) min-reduction picks
// the record the sequential loop would have kept:
//   sequential rule (lb_test_tris): accept when inside and (t < bt, or t == bt and a record of
//   this leaf is already held and rank < its rank); the survivor of a group is therefore the
//   lexicographic minimum (t, rank) over the records that pass that test against the state the
//   group started with.
// slot_owner: 8 bytes of shared memory per warp.
template <bool STATS>
__device__ __forceinline__ void ph_lb_tris_coop(const DScene& S, PhLane& w, const Ray& r,
                                                const PhSmem& sm, Counters& c, unsigned owners,
                                                volatile uint8_t* slot_owner) {
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const unsigned below = (1u << lane) - 1u;
  const bool own = (owners >> lane) & 1u;
  uint32_t k0 = 0, n = 0;
  if (own) {
    uint32_t g = (uint32_t)__ffs((int)w.qmask) - 1u;
    w.qmask &= w.qmask - 1u;
    uint32_t rem = w.qcount - 4u * g;
    n = rem < 4u ? rem : 4u;
    k0 = w.qbase + 4u * g;
    if (STATS) c.n_tritest += n;
    if (w.qmask) prefetch_l1(S.lb_plane + w.qbase + 4u * ((uint32_t)__ffs((int)w.qmask) - 1u));
  }
  unsigned todo = owners;
  while (todo) {
    // the first 8 owners take the 8 four-lane slots
    const uint32_t rank = (uint32_t)__popc(todo & below);
    const bool served = ((todo >> lane) & 1u) && rank < 8u;
    if (served) slot_owner[rank] = (uint8_t)lane;
    __syncwarp();
    const uint32_t n_served = min((uint32_t)__popc(todo), 8u);
    const uint32_t slot = (uint32_t)lane >> 2, j = (uint32_t)lane & 3u;
    const int src = slot < n_served ? (int)slot_owner[slot] : 0;
    __syncwarp();
    const uint32_t ko = __shfl_sync(FULL, k0, src), no = __shfl_sync(FULL, n, src);
    const V3 s = v3(__shfl_sync(FULL, r.start.x, src), __shfl_sync(FULL, r.start.y, src),
                    __shfl_sync(FULL, r.start.z, src));
    const V3 d = v3(__shfl_sync(FULL, r.dir.x, src), __shfl_sync(FULL, r.dir.y, src),
                    __shfl_sync(FULL, r.dir.z, src));
    const float bt = __shfl_sync(FULL, w.rec.param, src);
    const uint32_t border = __shfl_sync(FULL, w.rec.prim, src);
    const bool found = __shfl_sync(FULL, w.flags & PH_F_FOUND, src) != 0u;
    float kt = FS2_INFINITY, u = 0.f, v = 0.f;
    uint32_t kord = 0xFFFFFFFFu;
    if (slot < n_served && j < no) {
      const uint32_t k = ko + j;
      float4 pl = ldg_once(S.lb_plane + k);
      float ts = pl.x * d.x + pl.y * d.y + pl.z * d.z;
      if (!near_zero(ts)) {
        float ns = -(pl.x * s.x + pl.y * s.y + pl.z * s.z + pl.w);
        float t = ns / ts;
        if (t >= 0.0f && t <= 1.0f && (t < bt || (t == bt && found))) {
          uint32_t ord = ldg_once(S.lb_order + k);
          if (t < bt || ord < border) {
            if (STATS) c.n_bary++;
            float uu, vv;
            if (tri_inside(S.lb_aux + 3 * (size_t)k, s + d * t, uu, vv)) {
              kt = t;
              kord = ord;
              u = uu;
              v = vv;
            }
          }
        }
      }
    }
#pragma unroll
    for (int off = 1; off <= 2; off <<= 1) {
      float ot = __shfl_xor_sync(FULL, kt, off);
      uint32_t oo = __shfl_xor_sync(FULL, kord, off);
      float ou = __shfl_xor_sync(FULL, u, off), ov = __shfl_xor_sync(FULL, v, off);
      if (ot < kt || (ot == kt && oo < kord)) {
        kt = ot;
        kord = oo;
        u = ou;
        v = ov;
      }
    }
    // hand the slot's winner back to its owner
    const int from = served ? (int)(4u * rank) : 0;
    const float rt = __shfl_sync(FULL, kt, from);
    const uint32_t ro = __shfl_sync(FULL, kord, from);
    const float ru = __shfl_sync(FULL, u, from), rv = __shfl_sync(FULL, v, from);
    if (served && ro != 0xFFFFFFFFu) {
      w.rec.param = rt;
      w.rec.object = w.cur_obj;
      w.rec.prim = ro;
      w.rec.u = ru;
      w.rec.v = rv;
      w.flags |= PH_F_FOUND;
    }
    const unsigned served_mask = __ballot_sync(FULL, served);
    todo &= ~served_mask;
  }
  if (own && !w.qmask) ph_lb_pop(w, S, sm);
}

}  // namespace fs2
