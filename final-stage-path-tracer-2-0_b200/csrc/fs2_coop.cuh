// fs2_coop.cuh — resumable per-lane traversal + warp-cooperative mesh-leaf scans.
//
// Why: with one thread walking its own leaf (fs2_traverse.cuh: mesh_leaf) the octree's fat
// leaves (mean 105, up to 2 600 triangles at the depth cap of 4) make lanes of a warp loop
// for wildly different trip counts over scattered addresses; ncu on the 3x100k-triangle
// scene shows 2.2 active threads per executed warp instruction.  Here every lane runs its
// traversal as a small state machine that *yields* when it reaches a non-empty mesh leaf;
// the warp then serves the posted leaves one after the other with all 32 lanes testing 32
// consecutive triangle records per step (coalesced 16-byte plane loads), followed by a
// (param, index) min-reduction that reproduces the reference's sequential "first strictly
// smaller parameter wins" rule (reference source/mesh.cpp:89-110).
#pragma once
#include "fs2_traverse.cuh"

namespace fs2 {

// Leaves with at most this many triangles are scanned by the owning lane on the spot.
#ifndef FS2_COOP_MIN_LEAF
#define FS2_COOP_MIN_LEAF 4
#endif

enum WalkResult { WALK_LEAF = 0, WALK_DONE = 1 };

// Resumable form of walk_octree (fs2_traverse.cuh).  Start with resume = false and `cur`
// = root; the call returns WALK_LEAF with (first, count) at every non-empty leaf and must
// be re-entered with resume = true, ret = "the leaf improved the record"; WALK_DONE
// delivers the walk's result in `ret`.  best_param is the active record's parameter.
template <bool STATS>
__device__ __forceinline__ WalkResult walk_resume(const DScene& S, Frame* stack, int maxd, int& sp,
                                                  uint32_t& cur, const Ray& r, float best_param,
                                                  bool resume, bool& ret, uint32_t& first,
                                                  uint32_t& count, Counters& c) {
  bool returning = resume;
  for (;;) {
    if (!returning) {
      NodeRec n = load_node(S, cur);
      float nparam;
      V3 entry = r.start;
      bool inside;
      if (STATS) c.n_node++;
      if (!ray_box(n.mn, n.mx, r, nparam, entry, inside) || nparam > best_param) {
        ret = false;
        returning = true;
      } else if (n.count != FS_NODE_INTERNAL) {
        if (n.count) {
          first = n.first;
          count = n.count;
          return WALK_LEAF;
        }
        ret = false;
        returning = true;
      } else if (sp < maxd) {
        V3 cen = (n.mn + n.mx) * 0.5f;
        V3 st = inside ? r.start : entry;
        Frame f;
        f.node = cur;
        f.bits = (st.x - cen.x >= 0.0f ? 1u : 0u) | (st.z - cen.z >= 0.0f ? 2u : 0u) |
                 (st.y - cen.y >= 0.0f ? 4u : 0u);
        f.tx = split_param(st.x, r.dir.x, cen.x);
        f.ty = split_param(st.y, r.dir.y, cen.y);
        f.tz = split_param(st.z, r.dir.z, cen.z);
        f.sx = st.x;
        f.sy = st.y;
        f.sz = st.z;
        stack[sp++] = f;
        cur = n.first + (f.bits & 7u);
        continue;
      } else {
        ret = false;
        returning = true;
      }
    }
    // ---- return `ret` into the enclosing frame
    if (sp == 0) return WALK_DONE;
    Frame& f = stack[sp - 1];
    NodeRec pn = load_node(S, f.node);
    if (ret) {
      f.bits |= 32u;
      NodeRec ch = load_node(S, pn.first + (f.bits & 7u));
      if (point_in_box(ch.mn, ch.mx, r.start + r.dir * best_param)) {
        sp--;
        ret = true;
        continue;
      }
    }
    uint32_t visits = ((f.bits >> 3) & 3u) + 1u;
    bool x_hit = f.tx <= 1.0f, y_hit = f.ty <= 1.0f, z_hit = f.tz <= 1.0f;
    float t = 0.f;
    uint32_t flip = 0u;
    if (x_hit && f.tx < f.ty && f.tx < f.tz) {
      t = f.tx;
      f.tx = FS2_INFINITY;
      flip = 1u;
    } else if (y_hit && f.ty < f.tz && f.ty < f.tx) {
      t = f.ty;
      f.ty = FS2_INFINITY;
      flip = 4u;
    } else if (z_hit && f.tz < f.ty && f.tz < f.tx) {
      t = f.tz;
      f.tz = FS2_INFINITY;
      flip = 2u;
    }
    bool go = flip != 0u && visits < 4u;
    if (go) go = point_in_box(pn.mn, pn.mx, v3(f.sx, f.sy, f.sz) + r.dir * t);
    if (!go) {
      ret = (f.bits & 32u) != 0u;
      sp--;
      continue;
    }
    f.bits = ((f.bits ^ flip) & ~(3u << 3)) | (visits << 3);
    cur = pn.first + (f.bits & 7u);
    returning = false;
  }
}

// ------------------------------------------------------------------ per-lane state machine
enum Pc {
  PC_IDLE = 0,       // lane has no segment in flight
  PC_SEG_START,      // begin Scene::Trace for `ray`
  PC_NEXT_OBJ,       // brute force: next entry of objects[]
  PC_SCENE_RUN,      // (re)enter the scene-tree walk
  PC_LEAF_NEXT_OBJ,  // scene-tree leaf: next referenced object
  PC_MESH_RUN,       // (re)enter the mesh-tree walk of the current object
  PC_LEAF_WAIT,      // a mesh leaf is posted for the cooperative scan
  PC_SEG_DONE        // nearest hit of the segment is in `rec`
};

struct LaneWalk {
  uint32_t pc;
  Hit rec;   // active record (the scene-level best, or the leaf-local record of scene.cpp:84)
  Hit glob;  // scene-level best while a scene-tree leaf is being processed
  // object iteration
  uint32_t obj_i;      // brute force cursor
  uint32_t cur_obj;    // object being traced
  uint32_t sleaf_first, sleaf_count, sleaf_i;
  bool sleaf_improved;
  // walkers
  uint32_t scur, mcur;
  int ssp, msp;
  bool resume, ret;
  // posted mesh leaf
  uint32_t leaf_first, leaf_count;
};

// Runs the lane's traversal until it needs a cooperative leaf scan (PC_LEAF_WAIT) or the
// segment is finished (PC_SEG_DONE).
template <bool STATS>
__device__ __forceinline__ void lane_advance(const DScene& S, LaneWalk& w, const Ray& r,
                                             Frame* sstack, Frame* mstack, Counters& c) {
  while (w.pc != PC_LEAF_WAIT && w.pc != PC_SEG_DONE && w.pc != PC_IDLE) {
    switch (w.pc) {
      case PC_SEG_START: {
        w.rec = no_hit();
        if (S.scene_root < 0) {
          w.obj_i = 0;
          w.pc = PC_NEXT_OBJ;
        } else {
          w.scur = (uint32_t)S.scene_root;
          w.ssp = 0;
          w.resume = false;
          w.pc = PC_SCENE_RUN;
        }
        break;
      }
      case PC_NEXT_OBJ: {  // scene.cpp:230-233
        if (w.obj_i >= S.n_objects) {
          w.pc = PC_SEG_DONE;
          break;
        }
        uint32_t oi = w.obj_i++;
        float4 w0 = obj_word(S, oi, 0);
        if (__float_as_uint(w0.x) == FS_OBJ_MESH) {
          uint32_t root = __float_as_uint(w0.z);
          if (root != 0xFFFFFFFFu) {
            w.cur_obj = oi;
            w.mcur = root;
            w.msp = 0;
            w.resume = false;
            w.pc = PC_MESH_RUN;
          }
        } else {
          object_trace<STATS>(S, oi, r, w.rec, c);
        }
        break;
      }
      case PC_SCENE_RUN: {  // scene.cpp:76-103 over bvh.h:154-231
        uint32_t first, count;
        bool ret = w.ret;
        WalkResult wr = walk_resume<STATS>(S, sstack, FS2_SCENE_STACK, w.ssp, w.scur, r,
                                           w.rec.param, w.resume, ret, first, count, c);
        if (wr == WALK_DONE) {
          w.pc = PC_SEG_DONE;
        } else {
          // scene.cpp:84-99: one leaf-local record, seeded with "no hit"
          w.sleaf_first = first;
          w.sleaf_count = count;
          w.sleaf_i = 0;
          w.sleaf_improved = false;
          w.glob = w.rec;
          w.rec = no_hit();
          w.pc = PC_LEAF_NEXT_OBJ;
        }
        break;
      }
      case PC_LEAF_NEXT_OBJ: {
        if (w.sleaf_i >= w.sleaf_count) {
          w.rec = w.glob;
          w.resume = true;
          w.ret = w.sleaf_improved;
          w.pc = PC_SCENE_RUN;
          break;
        }
        uint32_t oi = __ldg(S.scene_refs + w.sleaf_first + w.sleaf_i);
        w.sleaf_i++;
        float4 w0 = obj_word(S, oi, 0);
        if (__float_as_uint(w0.x) == FS_OBJ_MESH) {
          uint32_t root = __float_as_uint(w0.z);
          if (root != 0xFFFFFFFFu) {
            w.cur_obj = oi;
            w.mcur = root;
            w.msp = 0;
            w.resume = false;
            w.pc = PC_MESH_RUN;
          }
        } else if (object_trace<STATS>(S, oi, r, w.rec, c) && w.rec.param < w.glob.param) {
          w.glob = w.rec;
          w.sleaf_improved = true;
        }
        break;
      }
      case PC_MESH_RUN: {  // mesh.cpp:78-115,277-322
        uint32_t first, count;
        bool ret = w.ret;
        WalkResult wr = walk_resume<STATS>(S, mstack, FS2_MESH_STACK, w.msp, w.mcur, r,
                                           w.rec.param, w.resume, ret, first, count, c);
        if (wr == WALK_LEAF) {
          if (r.zero_length || count <= FS2_COOP_MIN_LEAF) {
            w.ret = mesh_leaf<STATS>(S, first, count, r, w.cur_obj, w.rec, c);
            w.resume = true;
          } else {
            w.leaf_first = first;
            w.leaf_count = count;
            w.pc = PC_LEAF_WAIT;
          }
        } else {
          // MeshObject::Trace finished with `ret`
          if (S.scene_root < 0) {
            w.pc = PC_NEXT_OBJ;
          } else {
            if (ret && w.rec.param < w.glob.param) {
              w.glob = w.rec;
              w.sleaf_improved = true;
            }
            w.pc = PC_LEAF_NEXT_OBJ;
          }
        }
        break;
      }
      default:
        w.pc = PC_SEG_DONE;
        break;
    }
  }
}

// Serves every lane that posted a leaf.  Must be called by the full warp.
template <bool STATS>
__device__ __forceinline__ void warp_scan_leaves(const DScene& S, LaneWalk& w, const Ray& r,
                                                 Counters& c) {
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  unsigned want = __ballot_sync(FULL, w.pc == PC_LEAF_WAIT);
  while (want) {
    int src = __ffs(want) - 1;
    want &= want - 1;
    const float sx = __shfl_sync(FULL, r.start.x, src), sy = __shfl_sync(FULL, r.start.y, src),
                sz = __shfl_sync(FULL, r.start.z, src);
    const float dx = __shfl_sync(FULL, r.dir.x, src), dy = __shfl_sync(FULL, r.dir.y, src),
                dz = __shfl_sync(FULL, r.dir.z, src);
    const float limit = __shfl_sync(FULL, w.rec.param, src);
    const uint32_t first = __shfl_sync(FULL, w.leaf_first, src);
    const uint32_t count = __shfl_sync(FULL, w.leaf_count, src);
    float bt = limit, bu = 0.f, bv = 0.f;
    uint32_t bk = 0xFFFFFFFFu;
    uint32_t nbary = 0;
    for (uint32_t base = 0; base < count; base += 32) {
      uint32_t i = base + lane;
      if (i < count) {
        uint32_t k = first + i;
        float4 pl = __ldg(S.tri_plane + k);
        float ts = pl.x * dx + pl.y * dy + pl.z * dz;
        if (!near_zero(ts)) {
          float ns = -(pl.x * sx + pl.y * sy + pl.z * sz + pl.w);
          float t = ns / ts;
          if (t >= 0.0f && t <= 1.0f) {
            if (STATS) nbary++;
            if (t < bt) {
              float u, v;
              V3 p = v3(sx, sy, sz) + v3(dx, dy, dz) * t;
              if (tri_inside(S.tri_aux + 3 * (size_t)k, p, u, v)) {
                bt = t;
                bk = k;
                bu = u;
                bv = v;
              }
            }
          }
        }
      }
    }
    // nearest candidate of the warp; equal parameters resolve to the lowest record index,
    // i.e. the one the sequential loop meets first.  -0.0 orders like +0.0.
    uint32_t key = (bk != 0xFFFFFFFFu) ? __float_as_uint(bt + 0.0f) : 0xFFFFFFFFu;
    uint32_t kmin = __reduce_min_sync(FULL, key);
    bool improved = kmin != 0xFFFFFFFFu;
    uint32_t cand = (key == kmin) ? bk : 0xFFFFFFFFu;
    uint32_t wk = __reduce_min_sync(FULL, cand);
    int winner = __ffs(__ballot_sync(FULL, improved && cand == wk)) - 1;
    if (winner < 0) winner = 0;
    float wt = __shfl_sync(FULL, bt, winner), wu = __shfl_sync(FULL, bu, winner),
          wv = __shfl_sync(FULL, bv, winner);
    if (STATS) {
      nbary = __reduce_add_sync(FULL, nbary);
      if (lane == src) {
        c.n_tri += count;
        c.n_bary += nbary;
      }
    }
    if (lane == src) {
      if (improved) {
        w.rec.param = wt;
        w.rec.object = w.cur_obj;
        w.rec.prim = wk;
        w.rec.u = wu;
        w.rec.v = wv;
      }
      w.ret = improved;
      w.resume = true;
      w.pc = PC_MESH_RUN;
    }
  }
}

}  // namespace fs2
