// fs2_traverse.cuh — nearest-hit query of one ray segment against the flattened scene:
// the reference's object loop / scene octree / per-mesh octree with sibling stepping,
// restructured as an iterative walker with an explicit frame stack (the reference is
// recursive: source/scene.cpp:76-103,227-249, source/bvh.h:154-231,
// source/mesh.cpp:78-115,277-322, source/object.cpp:23-237).
//
// Only (param, object, prim, u, v) are tracked while walking; hit point, normal and
// texture coordinates are rebuilt for the winner afterwards with the same expressions
// the reference evaluates at test time, so the record is bit-identical.
#pragma once
#include "fs2_math.cuh"
#include "../../include/fs2_cabi.h"

namespace fs2 {

struct DScene {
  const float4* nodes;      // 2 x float4 per fs_node
  const uint32_t* scene_refs;
  const float4* sref_sphere;  // per scene ref: sphere centre xyz + radius, or a NaN radius (fs2_cabi.cu)
  const float4* tri_plane;  // 1 x float4 per packed triangle
  const float4* tri_aux;    // 3 x float4 per fs_tri_aux
  const float4* faces;      // 4 x float4 per fs_face_attr
  const float4* objects;    // 9 x float4 per fs_object
  const float4* materials;  // 4 x float4 per fs_material
  const fs_texture* textures;
  const float* texels;
  const float4* lut;        // xyz + pad
  // leaf BVH (fs2_lbvh_build.h): permuted triangle records + 128-byte 4-wide nodes
  const float4* lb_nodes;   // 8 x float4 per LbNode
  const uint2* lb_leaf;     // per octree node: (root LbNode or 0xFFFFFFFF, first record)
  const float4* lb_plane;
  const float4* lb_aux;
  const uint32_t* lb_order; // permuted index -> original packed index
  uint32_t n_objects;
  uint32_t n_lut;
  uint32_t sky_material;
  int32_t scene_root;
  // segments that start farther than this from the origin (max-norm) scan mesh leaves linearly:
  // the leaf BVH's box pads bound the rounding of hit points only up to that distance
  // (fs2_lbvh_build.h record_box (b))
  float lb_origin_limit;
};

__device__ __forceinline__ bool lb_origin_far(const DScene& S, V3 s) {
  return !(fmaxf(fmaxf(fabsf(s.x), fabsf(s.y)), fabsf(s.z)) <= S.lb_origin_limit);
}

struct Counters {
  uint32_t n_node, n_tri, n_bary, n_prim;  // reference-equivalent work
  uint32_t n_lbnode, n_tritest;            // leaf-BVH kernel's own work: nodes visited, planes tested
};

// Running nearest hit.  prim: packed-triangle index for meshes, face plane for cuboids.
struct Hit {
  float param;
  uint32_t object;
  uint32_t prim;
  float u, v;
};

__device__ __forceinline__ Hit no_hit() {
  Hit h;
  h.param = FS_NO_HIT_PARAM;
  h.object = 0xFFFFFFFFu;
  h.prim = 0xFFFFFFFFu;
  h.u = 0.f;
  h.v = 0.f;
  return h;
}

#ifndef FS2_CUBOID_REJECT
#define FS2_CUBOID_REJECT 1  // bounding-sphere early out in front of the cuboid's six-plane loop
#endif
#define FS2_MESH_STACK 5
#define FS2_SCENE_STACK 8

// ------------------------------------------------------------------ plane / box tests
// reference source/math/intersect.cpp:728-744.  Returns the parameter, or -1 for "no hit"
// (valid parameters are in [0,1]).
__device__ __forceinline__ float plane_param(float4 pl, V3 start, V3 dir) {
  float ts = pl.x * dir.x + pl.y * dir.y + pl.z * dir.z;
  if (near_zero(ts)) return -1.0f;
  float ns = -(pl.x * start.x + pl.y * start.y + pl.z * start.z + pl.w);
  float t = ns / ts;
  return (t >= 0.0f && t <= 1.0f) ? t : -1.0f;
}

// One slab of reference source/math/intersect.cpp:490-529.
__device__ __forceinline__ bool slab(float start, float vec, float smin, float smax, float& tnear,
                                     float& tfar) {
  if (near_zero(vec)) return !(start < smin || start > smax);
  float t1 = (smin - start) / vec;
  float t2 = (smax - start) / vec;
  if (t1 > t2) {
    float tmp = t1;
    t1 = t2;
    t2 = tmp;
  }
  if (t1 > tnear) tnear = t1;
  if (t2 < tfar) tfar = t2;
  return true;
}

// reference source/math/intersect.cpp:531-597
__device__ __forceinline__ bool box_worker(V3 mn, V3 mx, V3 start, V3 vec, float& param) {
  float tnear = -FS2_INFINITY, tfar = FS2_INFINITY;
  if (!slab(start.x, vec.x, mn.x, mx.x, tnear, tfar)) return false;
  if (!slab(start.y, vec.y, mn.y, mx.y, tnear, tfar)) return false;
  if (!slab(start.z, vec.z, mn.z, mx.z, tnear, tfar)) return false;
  if (tnear > tfar) return false;
  if (tfar < 0.0f || tnear < 0.0f || tnear > 1.0f) return false;
  param = tnear;
  return true;
}

// reference source/math/intersect.cpp:599-654.  `inside` reports point_in_bounds(start);
// entry point is only produced for an outside origin (the only case that reads it).
__device__ __forceinline__ bool ray_box(V3 mn, V3 mx, const Ray& r, float& param, V3& entry,
                                        bool& inside) {
  inside = point_in_box(mn, mx, r.start);
  if (r.zero_length) {
    param = 0.0f;
    return inside;
  }
  if (inside && point_in_box(mn, mx, r.stop)) {
    param = 0.0f;
    return true;
  }
  // One instance of the slab worker serves both cases (code size: the traversal loop has
  // to stay inside the 32 KB instruction cache).  Inside origin: the reversed segment
  // (intersect.cpp:621-649); outside: the worker recomputes stop - start, which for bounce
  // rays is not bit-equal to r.dir (engine.cpp:120-127 offsets start and dir separately).
  V3 o = inside ? r.stop : r.start;
  V3 vec = inside ? (r.start - r.stop) : (r.stop - r.start);
  float p;
  if (!box_worker(mn, mx, o, vec, p)) return false;
  if (inside) {
    param = 1.0f - p;
  } else {
    param = p;
    entry = r.start + vec * p;
  }
  return true;
}

// ------------------------------------------------------------------ triangles
// Barycentric stage on a packed record (reference source/math/intersect.cpp:96-129,
// 146-177; constants hoisted by include/fs2_pack.h with the same operations).
__device__ __forceinline__ bool tri_inside(const float4* __restrict__ aux, V3 point, float& u,
                                           float& v) {
  float4 a0 = ldg_once(aux), a1 = ldg_once(aux + 1), a2 = ldg_once(aux + 2);
  V3 vp = point - v3(a0.x, a0.y, a0.z);
  uint32_t drop = __float_as_uint(a0.w) >> 30;
  float vpa = (drop == 0u) ? vp.y : vp.x;
  float vpb = (drop == 2u) ? vp.y : vp.z;
  float dot02 = a1.x * vpa + a1.y * vpb;
  float dot12 = a1.z * vpa + a1.w * vpb;
  float uu = (a2.z * dot02 - a2.y * dot12) * a2.w;
  float vv = (a2.x * dot12 - a2.y * dot02) * a2.w;
  if (uu >= 0.0f && vv >= 0.0f && uu + vv <= 1.0f) {
    u = uu;
    v = vv;
    return true;
  }
  return false;
}

// Leaf loop of reference source/mesh.cpp:86-111 with ray_intersect_triangle
// (source/math/intersect.cpp:357-419) inlined; one thread walks the leaf.
template <bool STATS>
__device__ __forceinline__ bool mesh_leaf(const DScene& S, uint32_t first, uint32_t count,
                                          const Ray& r, uint32_t object, Hit& h, Counters& c) {
  bool improved = false;
  if (STATS) c.n_tri += count;
  if (r.zero_length) {
    // point_in_triangle path (intersect.cpp:364-373 → :318-345): the plane is a vector4 and
    // its dot with the promoted point adds w twice.
    for (uint32_t i = 0; i < count; i++) {
      uint32_t k = first + i;
      float4 pl = __ldg(S.tri_plane + k);
      float d = (pl.x * r.start.x + pl.y * r.start.y + pl.z * r.start.z + pl.w * 1.0f) + pl.w;
      if (!near_zero(d)) continue;
      float u, v;
      if (tri_inside(S.tri_aux + 3 * (size_t)k, r.start, u, v) && 0.0f < h.param) {
        h.param = 0.0f;
        h.object = object;
        h.prim = k;
        h.u = u;
        h.v = v;
        improved = true;
      }
    }
    return improved;
  }
  const V3 s = r.start, d = r.dir;
  for (uint32_t i = 0; i < count; i++) {
    uint32_t k = first + i;
    float4 pl = __ldg(S.tri_plane + k);
    float ts = pl.x * d.x + pl.y * d.y + pl.z * d.z;
    if (near_zero(ts)) continue;
    float ns = -(pl.x * s.x + pl.y * s.y + pl.z * s.z + pl.w);
    float t = ns / ts;
    if (!(t >= 0.0f && t <= 1.0f)) continue;
    if (STATS) c.n_bary++;
    // The reference runs the barycentric stage and then drops the candidate unless
    // t < best (mesh.cpp:96); skipping it earlier cannot change the record.
    if (!(t < h.param)) continue;
    float u, v;
    if (tri_inside(S.tri_aux + 3 * (size_t)k, s + d * t, u, v)) {
      h.param = t;
      h.object = object;
      h.prim = k;
      h.u = u;
      h.v = v;
      improved = true;
    }
  }
  return improved;
}

}  // namespace fs2
#include "fs2_lbvh.cuh"
namespace fs2 {

// Out-of-line copy for the rare callers inside the phase machine (zero-length rays, leaves
// beyond the leaf-BVH level cap).  Everything goes in and out by value: a reference
// parameter would pin the caller's ray / record / scene descriptor in local memory for the
// whole kernel (ncu: 5 % of the traversal kernel's instructions were LDL/STL).
struct ColdResult {
  Hit h;
  Counters c;
  bool ret;
};
template <bool STATS>
__device__ __noinline__ ColdResult mesh_leaf_cold(DScene S, uint32_t first, uint32_t count, Ray r,
                                                  uint32_t object, Hit h, Counters c) {
  ColdResult o;
  o.h = h;
  o.c = c;
  o.ret = mesh_leaf<STATS>(S, first, count, r, object, o.h, o.c);
  return o;
}

// ------------------------------------------------------------------ octree walker
// Frame of one internal node whose children are being visited (bvh.h:154-231).
struct Frame {
  uint32_t node;
  uint32_t bits;  // bits 0-2 current child, 3-4 visits done, 5 some child improved the hit
  float tx, ty, tz;  // split-plane parameters: <=1 pending, 2.0 never hit, 1e15 consumed
  float sx, sy, sz;  // start point the plane parameters are measured from
};

struct NodeRec {
  V3 mn, mx;
  uint32_t first, count;
  uint32_t cmask;  // internal nodes: bit c set = child c is not an empty leaf
};

// Device copy of fs_node (tag_nodes() in fs2_cabi.cu): the `first` word of an internal node
// carries the occupancy of its eight children in its top byte (node indices are < 2^24).
#define FS2_NODE_INDEX_BITS 24
__device__ __forceinline__ NodeRec load_node(const DScene& S, uint32_t i) {
  float4 a = ldg_keep(S.nodes + 2 * (size_t)i), b = ldg_keep(S.nodes + 2 * (size_t)i + 1);
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

// Split-plane hit of reference bvh.h:171-181 for the axis plane through the node centre:
// calculate_plane(axis, centre) gives w = -centre.axis, and the dot products with a unit
// axis reduce to the component itself.  2.0 = "not hit" (collision ctor, trace.cpp:18-21).
__device__ __forceinline__ float split_param(float start_c, float dir_c, float centre_c) {
  if (near_zero(dir_c)) return 2.0f;
  float ns = -(start_c + (-centre_c));
  float t = ns / dir_c;
  return (t >= 0.0f && t <= 1.0f) ? t : 2.0f;
}

// Walks the octree under `root`.  leaf(node, first, count, best) returns true when it improved
// `best`; the walker only needs best.param (culling, early exit).  Returns true when any
// leaf improved the record, i.e. the reference's node->Trace() result.
template <int MAXD, bool STATS, class Leaf>
__device__ __forceinline__ bool walk_octree(const DScene& S, uint32_t root, const Ray& r,
                                            Hit& best, Counters& c, Leaf leaf) {
  Frame stack[MAXD];
  int sp = 0;
  uint32_t cur = root;
  bool ret = false;
  for (;;) {
    // ---- "call" node cur (mesh.cpp:78-84 / scene.cpp:76-83)
    {
      NodeRec n = load_node(S, cur);
      float nparam;
      V3 entry = r.start;
      bool inside;
      if (STATS) c.n_node++;
      if (!ray_box(n.mn, n.mx, r, nparam, entry, inside) || nparam > best.param) {
        ret = false;
      } else if (n.count != FS_NODE_INTERNAL) {
        ret = n.count ? leaf(cur, n.first, n.count, best) : false;
      } else if (sp < MAXD) {
        // bvh.h:160-186 — the zero-length ray has entry == start as well
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
        ret = false;  // deeper than any tree the host builds
      }
    }
    // ---- "return" ret to the enclosing frames
    for (;;) {
      if (sp == 0) return ret;
      Frame& f = stack[sp - 1];
      NodeRec pn = load_node(S, f.node);
      if (ret) {
        f.bits |= 32u;
        NodeRec ch = load_node(S, pn.first + (f.bits & 7u));
        if (point_in_box(ch.mn, ch.mx, r.start + r.dir * best.param)) {
          sp--;
          ret = true;
          continue;
        }
      }
      uint32_t visits = ((f.bits >> 3) & 3u) + 1u;
      bool x_hit = f.tx <= 1.0f, y_hit = f.ty <= 1.0f, z_hit = f.tz <= 1.0f;
      float t;
      uint32_t flip;
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
      } else {
        flip = 0u;
        t = 0.f;
      }
      bool go = flip != 0u && visits < 4u;
      if (go) {
        V3 cross_pt = v3(f.sx, f.sy, f.sz) + r.dir * t;
        go = point_in_box(pn.mn, pn.mx, cross_pt);
      }
      if (!go) {
        ret = (f.bits & 32u) != 0u;
        sp--;
        continue;
      }
      f.bits = ((f.bits ^ flip) & ~(3u << 3)) | (visits << 3);
      cur = pn.first + (f.bits & 7u);
      break;
    }
  }
}

// ------------------------------------------------------------------ objects
__device__ __forceinline__ float4 obj_word(const DScene& S, uint32_t object, int w) {
  return __ldg(S.objects + 9 * (size_t)object + w);
}
// float f[k..k+3] of fs_object: the float payload starts at byte 24.
__device__ __forceinline__ void obj_f8(const DScene& S, uint32_t o, float4& w1, float4& w2,
                                       float4& w3) {
  w1 = obj_word(S, o, 1);
  w2 = obj_word(S, o, 2);
  w3 = obj_word(S, o, 3);
}
// f[4*i..4*i+3] straddles two 16-byte words (payload offset 24 = 16 + 8).
__device__ __forceinline__ float4 obj_f4(const DScene& S, uint32_t o, int i) {
  float4 a = obj_word(S, o, 1 + i), b = obj_word(S, o, 2 + i);
  return make_float4(a.z, a.w, b.x, b.y);
}

// reference source/math/intersect.cpp:421-488 + source/object.cpp:23-36
__device__ __forceinline__ bool sphere_test(V3 centre, float radius, const Ray& r, float& param) {
  if (r.zero_length) return false;  // the reference reports a hit with param 2.0, never < best
  V3 o = r.start - centre;
  float a = dot(r.dir, r.dir);
  float b = 2.0f * dot(o, r.dir);
  float cc = dot(o, o) - radius * radius;
  float d = b * b - 4.0f * a * cc;
  if (d < 0.0f) return false;
  float t;
  if (length(o) <= radius) {
    t = (-b + approx_sqrt(d)) / (2.0f * a);
  } else {
    t = (-b - approx_sqrt(d)) / (2.0f * a);
  }
  if (t < 0.0f || t > 1.0f) return false;
  param = t;
  return true;
}

template <bool STATS>
__device__ __forceinline__ bool primitive_trace(const DScene& S, uint32_t oi, uint32_t type,
                                                const Ray& r, Hit& h, Counters& c);

// Per-segment context of the leaf-BVH variant (LB = true); unused otherwise.
struct LbCtx {
  LbRay q;
  uint2* stack;  // this lane's column of the shared-memory stack
  bool far;      // lb_origin_far(ray start): linear leaf scans for this segment
};

template <bool STATS, bool LB = false>
__device__ __forceinline__ bool object_trace(const DScene& S, uint32_t oi, const Ray& r, Hit& h,
                                             Counters& c, const LbCtx* lb = nullptr) {
  float4 w0 = obj_word(S, oi, 0);
  uint32_t type = __float_as_uint(w0.x);
  if (type == FS_OBJ_MESH) {
    uint32_t root = __float_as_uint(w0.z);
    if (root == 0xFFFFFFFFu) return false;
    return walk_octree<FS2_MESH_STACK, STATS>(
        S, root, r, h, c, [&](uint32_t node, uint32_t first, uint32_t count, Hit& best) {
          if (LB && !r.zero_length && !lb->far)
            return lbvh_leaf<STATS>(S, node, first, count, r, lb->q, oi, best, c, lb->stack);
          return mesh_leaf<STATS>(S, first, count, r, oi, best, c);
        });
  }
  return primitive_trace<STATS>(S, oi, type, r, h, c);
}

// The analytic primitives of reference source/object.cpp:23-237 (everything but meshes).
template <bool STATS>
__device__ __forceinline__ bool primitive_trace(const DScene& S, uint32_t oi, uint32_t type,
                                                const Ray& r, Hit& h, Counters& c) {
  if (STATS) c.n_prim++;
  float4 w1 = obj_word(S, oi, 1), w2 = obj_word(S, oi, 2);
  // payload f[0..] = w1.z, w1.w, w2.x, w2.y, w2.z, w2.w, w3.x ...
  switch (type) {
    case FS_OBJ_SPHERE: {
      float t;
      if (sphere_test(v3(w1.z, w1.w, w2.x), w2.y, r, t) && t < h.param) {
        h.param = t;
        h.object = oi;
        h.prim = 0xFFFFFFFFu;
        return true;
      }
      return false;
    }
    case FS_OBJ_PLANE: {
      float t = plane_param(make_float4(w1.z, w1.w, w2.x, w2.y), r.start, r.dir);
      if (t >= 0.0f && t < h.param) {
        h.param = t;
        h.object = oi;
        h.prim = 0xFFFFFFFFu;
        return true;
      }
      return false;
    }
    case FS_OBJ_DISC: {  // object.cpp:97-113
      float t = plane_param(make_float4(w1.z, w1.w, w2.x, w2.y), r.start, r.dir);
      if (t < 0.0f) return false;
      float4 w3 = obj_word(S, oi, 3);
      V3 p = r.start + r.dir * t;
      if (length(v3(w2.z, w2.w, w3.x) - p) <= w3.y && t < h.param) {
        h.param = t;
        h.object = oi;
        h.prim = 0xFFFFFFFFu;
        return true;
      }
      return false;
    }
    case FS_OBJ_QUAD: {  // object.cpp:213-237
      float t = plane_param(make_float4(w1.z, w1.w, w2.x, w2.y), r.start, r.dir);
      if (t < 0.0f) return false;
      float4 w3 = obj_word(S, oi, 3), w4 = obj_word(S, oi, 4), w5 = obj_word(S, oi, 5);
      V3 ph = (r.start + r.dir * t) - v3(w2.z, w2.w, w3.x);
      float tangent_dist = dot(v3(w3.y, w3.z, w3.w), ph);
      float bitangent_dist = dot(v3(w4.x, w4.y, w4.z), ph);
      if (fabsf(bitangent_dist) > w4.w || fabsf(tangent_dist) > w5.x) return false;
      if (t < h.param) {
        h.param = t;
        h.object = oi;
        h.prim = 0xFFFFFFFFu;
        return true;
      }
      return false;
    }
    case FS_OBJ_CUBOID: {  // object.cpp:126-165
#if FS2_CUBOID_REJECT
      {
        // Conservative early out (ncu on the 303-object scene: the six-plane loop below ran at 2-4
        // of 32 lanes and was 45 % of the kernel's instructions although one object in ten is a
        // cuboid).  A face is only accepted at a point p = start + dir * t that passes the other
        // four plane tests, i.e. within rounding (~1e-6 (|centre| + R)) of the solid itself, which
        // lies inside the sphere around the centre of Object::GetBounds() through its corners.  A
        // segment whose supporting line stays outside that sphere — radius^2 padded by 5 % plus
        // 4e-6 (|start - c|^2 + |c|^2), which also covers the cancellation in oo a - b^2 — or that
        // starts outside it pointing away cannot reach any face.
        const float4 w7 = obj_word(S, oi, 7), w8 = obj_word(S, oi, 8);
        const V3 bmn = v3(w7.z, w7.w, w8.x), bmx = v3(w8.y, w8.z, w8.w);
        const V3 cen = (bmn + bmx) * 0.5f, hd = (bmx - bmn) * 0.5f;
        const V3 o = r.start - cen;
        const float oo = dot(o, o), a = dot(r.dir, r.dir), b = dot(o, r.dir);
        const float lim = 1.05f * dot(hd, hd) + 4.0e-6f * (oo + dot(cen, cen));
        if (oo > lim && (b >= 0.0f || oo * a - b * b > lim * a)) return false;
      }
#endif
      bool detected = false;
#pragma unroll 1
      for (int i = 0; i < 6; i++) {
        float t = plane_param(obj_f4(S, oi, i), r.start, r.dir);
        if (t < 0.0f || !(t < h.param)) continue;
        V3 p = r.start + r.dir * t;
        bool on_face = true;
        for (int j = 0; j < 6; j++) {
          if ((i >> 1) == (j >> 1)) continue;
          float4 q = obj_f4(S, oi, j);
          if (q.x * p.x + q.y * p.y + q.z * p.z + q.w > 0.0f) {
            on_face = false;
            break;
          }
        }
        if (on_face) {
          detected = true;
          h.param = t;
          h.object = oi;
          h.prim = (uint32_t)i;
        }
      }
      return detected;
    }
  }
  return false;
}

// reference source/scene.cpp:227-237 (the normal flip of :238-247 is applied when the
// surface record is rebuilt).
template <bool STATS, bool LB = false>
__device__ __forceinline__ Hit scene_trace(const DScene& S, const Ray& r, Counters& c,
                                           uint2* lb_stack = nullptr) {
  Hit best = no_hit();
  LbCtx lbc;
  if (LB) {
    lbc.q = make_lbray(r);
    lbc.stack = lb_stack;
    lbc.far = lb_origin_far(S, r.start);
  }
  const LbCtx* lb = LB ? &lbc : nullptr;
  if (S.scene_root < 0) {
    for (uint32_t i = 0; i < S.n_objects; i++) object_trace<STATS, LB>(S, i, r, best, c, lb);
  } else {
    walk_octree<FS2_SCENE_STACK, STATS>(
        S, (uint32_t)S.scene_root, r, best, c,
        [&](uint32_t, uint32_t first, uint32_t count, Hit& glob) {
          // scene.cpp:84-99: one temporary record (param 2.0) shared by the leaf's objects
          bool improved = false;
          Hit temp = no_hit();
          for (uint32_t i = 0; i < count; i++) {
            uint32_t oi = __ldg(S.scene_refs + first + i);
            if (object_trace<STATS, LB>(S, oi, r, temp, c, lb) && temp.param < glob.param) {
              glob = temp;
              improved = true;
            }
          }
          return improved;
        });
  }
  return best;
}

// ------------------------------------------------------------------ surface record
struct Surface {
  V3 point, normal;
  float u, v;
  uint32_t material;  // index into materials[]
  uint32_t face;      // global face index (mesh), plane index (cuboid) or 0xFFFFFFFF
  bool internal;      // the segment starts behind the surface (ObjectCollision::is_internal)
};

// reference source/math/intersect.cpp:769-778 — signed compares
__device__ __forceinline__ void planar_uv(V3 p, V3 n, float& u, float& v) {
  if (n.x > n.y && n.x > n.z) {
    u = p.y;
    v = p.z;
  } else if (n.y > n.x && n.y > n.z) {
    u = p.x;
    v = p.z;
  } else {
    u = p.x;
    v = p.y;
  }
}

// reference source/math/intersect.cpp:780-784 (float atan2, `+ 0.5` and `1.0 -` in double)
__device__ __forceinline__ void sphere_uv(V3 n, float& u, float& v) {
  u = (float)((double)(atan2f(n.x, n.z) / (2.0f * FS2_PI)) + 0.5);
  float vv = (float)((double)n.y * 0.5 + 0.5);
  v = (float)(1.0 - (double)vv);
}

// Rebuilds ObjectCollision (object.h:47-61) for the winning primitive and applies the
// back-face flip of Scene::Trace (scene.cpp:238-247).  `want_uv` skips the atan2 of
// sphere texture coordinates when the material has no texture.
__device__ __forceinline__ Surface make_surface(const DScene& S, const Ray& r, const Hit& h,
                                                bool want_uv) {
  Surface sf;
  float4 w0 = obj_word(S, h.object, 0);
  uint32_t type = __float_as_uint(w0.x);
  sf.material = __float_as_uint(w0.y);
  sf.point = r.start + r.dir * h.param;
  sf.face = 0xFFFFFFFFu;
  sf.u = 0.f;
  sf.v = 0.f;
  if (type == FS_OBJ_MESH) {
    float4 pl = __ldg(S.tri_plane + h.prim);
    sf.normal = v3(pl.x, pl.y, pl.z);
    if (r.zero_length) sf.point = r.start;
    uint32_t face = __float_as_uint(__ldg(S.tri_aux + 3 * (size_t)h.prim).w) & 0x3FFFFFFFu;
    sf.face = face;
    uint32_t flags = __float_as_uint(obj_word(S, h.object, 1).x);  // u[2]
    const float4* fa = S.faces + 4 * (size_t)face;
    if (flags & 1u) {
      // triangle_interpolate_barycentric_coeff (intersect.cpp:131-144), un-normalised
      float4 a = __ldg(fa), b = __ldg(fa + 1), cc = __ldg(fa + 2);
      V3 n0 = v3(a.x, a.y, a.z), n1 = v3(a.w, b.x, b.y), n2 = v3(b.z, b.w, cc.x);
      sf.normal = n0 + (n1 - n0) * h.u + (n2 - n0) * h.v;
    }
    if (flags & 2u) {
      float4 cc = __ldg(fa + 2), d = __ldg(fa + 3);
      // uv[0] = (cc.y, cc.z), uv[1] = (cc.w, d.x), uv[2] = (d.y, d.z)
      sf.u = cc.y + (cc.w - cc.y) * h.u + (d.y - cc.y) * h.v;
      sf.v = cc.z + (d.x - cc.z) * h.u + (d.z - cc.z) * h.v;
    }
  } else if (type == FS_OBJ_SPHERE) {
    float4 w1 = obj_word(S, h.object, 1), w2 = obj_word(S, h.object, 2);
    sf.normal = normalize(sf.point - v3(w1.z, w1.w, w2.x));
    if (want_uv) sphere_uv(sf.normal, sf.u, sf.v);
  } else if (type == FS_OBJ_CUBOID) {
    float4 pl = obj_f4(S, h.object, (int)h.prim);
    sf.normal = v3(pl.x, pl.y, pl.z);
    sf.face = h.prim;
    planar_uv(sf.point, sf.normal, sf.u, sf.v);
    sf.u = sf.u * 0.1f;
    sf.v = sf.v * 0.1f;
  } else {
    float4 pl = obj_f4(S, h.object, 0);
    sf.normal = v3(pl.x, pl.y, pl.z);
    planar_uv(sf.point, sf.normal, sf.u, sf.v);
  }
  // scene.cpp:238-247: plane through the hit with the hit normal; origin behind it -> flip
  float w = -dot(sf.normal, sf.point);
  float side = sf.normal.x * r.start.x + sf.normal.y * r.start.y + sf.normal.z * r.start.z + w;
  sf.internal = side < 0.0f;
  if (sf.internal) sf.normal = neg(sf.normal);
  return sf;
}

}  // namespace fs2
