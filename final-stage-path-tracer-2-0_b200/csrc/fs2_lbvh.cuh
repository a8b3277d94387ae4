// fs2_lbvh.cuh — per-lane walk of the leaf BVH (fs2_lbvh_build.h) that replaces the linear
// leaf loop of reference source/mesh.cpp:89-110 with an exact equivalent.
//
// Exactness argument.  The reference's loop returns, among the leaf's triangles that pass
// ray_intersect_triangle (source/math/intersect.cpp:357-419) with a parameter strictly below
// the incoming best, the one with the smallest parameter, the first in leaf order on ties.
// Here (a) every triangle test is the same float expression sequence as mesh_leaf()
// (fs2_traverse.cuh), (b) candidates are ranked by (parameter, original leaf position), and
// (c) a group of triangles is skipped only when the segment provably stays outside its
// padded box or enters it beyond the current best — both decided with slack that dominates
// the rounding of the reference's own hit point.  Box arithmetic itself need not be
// bit-reproducible, only conservative.
#pragma once
// Included from the middle of fs2_traverse.cuh (after mesh_leaf, before the octree walker).

namespace fs2 {

#define FS2_LB_STACK 20        // 6 inner levels x 3 pending siblings + slack
#define FS2_LB_THREADS 128     // block size of the kernels that own a shared stack

// Ray constants of one segment for the box tests.
struct LbRay {
  float idx, idy, idz;  // 1 / dir, zero components replaced by a huge finite value
};

__device__ __forceinline__ float lb_inv(float d) {
  const float tiny = 1.0e-30f;
  if (fabsf(d) < tiny) d = (d < 0.0f) ? -tiny : tiny;
  return 1.0f / d;
}

__device__ __forceinline__ LbRay make_lbray(const Ray& r) {
  LbRay o;
  o.idx = lb_inv(r.dir.x);
  o.idy = lb_inv(r.dir.y);
  o.idz = lb_inv(r.dir.z);
  return o;
}

// Conservative entry parameter of the segment into child c of a node, or 3.0f when the
// segment misses the box or enters it beyond `lim`.
__device__ __forceinline__ float lb_child(float lox, float loy, float loz, float hix, float hiy,
                                          float hiz, V3 s, const LbRay& q, float lim) {
  float ax = (lox - s.x) * q.idx, bx = (hix - s.x) * q.idx;
  float ay = (loy - s.y) * q.idy, by = (hiy - s.y) * q.idy;
  float az = (loz - s.z) * q.idz, bz = (hiz - s.z) * q.idz;
  float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
  float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
  float lo = tn * 0.99998f;
  bool hit = lo <= tf * 1.00002f && lo <= lim;
  return hit ? lo : 3.0f;
}

// Tests the triangles [k0, k0+n) of the permuted arrays against the segment; same float
// sequence as mesh_leaf().  (bt, border, bk, bu, bv) is the running best of this leaf:
// border = original packed index of the best, 0 while nothing was found in this leaf.
template <bool STATS>
__device__ __forceinline__ void lb_test_tris(const DScene& S, uint32_t k0, uint32_t n, V3 s, V3 d,
                                             float& bt, uint32_t& border, bool& found, float& bu,
                                             float& bv, Counters& c) {
  if (STATS) c.n_tritest += n;
#pragma unroll 1
  for (uint32_t j = 0; j < n; j++) {
    uint32_t k = k0 + j;
    float4 pl = __ldg(S.lb_plane + k);
    float ts = pl.x * d.x + pl.y * d.y + pl.z * d.z;
    if (near_zero(ts)) continue;
    float ns = -(pl.x * s.x + pl.y * s.y + pl.z * s.z + pl.w);
    float t = ns / ts;
    if (!(t >= 0.0f && t <= 1.0f)) continue;
    uint32_t ord = 0;
    if (!(t < bt)) {
      // equal parameters: the sequential loop keeps the record it met first
      if (!(t == bt) || !found) continue;
      ord = __ldg(S.lb_order + k);
      if (!(ord < border)) continue;
    } else {
      ord = __ldg(S.lb_order + k);
    }
    float u, v;
    if (STATS) c.n_bary++;
    if (tri_inside(S.lb_aux + 3 * (size_t)k, s + d * t, u, v)) {
      bt = t;
      border = ord;
      found = true;
      bu = u;
      bv = v;
    }
  }
}

// L2 -> L1 prefetch hint: the phase machine knows a lane's next node / triangle group one
// loop iteration before it reads it, and other lanes' phases run in between.
__device__ __forceinline__ void prefetch_l1(const void* p) {
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}

// lb_test_tris for one group of n <= 4 records, restructured for memory-level parallelism:
// the four planes are fetched together and the plane stage runs on all of them before the
// (sequential, order-dependent) acceptance stage; the barycentric constants of the records
// that pass the plane stage are prefetched while the others are still being looked at.
// Same float expressions and the same acceptance order as lb_test_tris.
template <bool STATS>
__device__ __forceinline__ void lb_test_group(const DScene& S, uint32_t k0, uint32_t n, V3 s, V3 d,
                                              float& bt, uint32_t& border, bool& found, float& bu,
                                              float& bv, Counters& c) {
  if (STATS) c.n_tritest += n;
  float tt[4];
  unsigned cand = 0;
  {
    float4 pl[4];
#pragma unroll
    for (uint32_t j = 0; j < 4; j++) pl[j] = __ldg(S.lb_plane + k0 + (j < n ? j : n - 1u));
#pragma unroll
    for (uint32_t j = 0; j < 4; j++) {
      float ts = pl[j].x * d.x + pl[j].y * d.y + pl[j].z * d.z;
      float ns = -(pl[j].x * s.x + pl[j].y * s.y + pl[j].z * s.z + pl[j].w);
      bool ok = j < n && !near_zero(ts);
      float t = ns / (ok ? ts : 1.0f);
      ok = ok && t >= 0.0f && t <= 1.0f && !(t > bt);
      tt[j] = t;
      if (ok) {
        cand |= 1u << j;
        prefetch_l1(S.lb_aux + 3 * (size_t)(k0 + j));
      }
    }
  }
#pragma unroll 1
  while (cand) {
    uint32_t j = (uint32_t)__ffs((int)cand) - 1u;
    cand &= cand - 1u;
    float t = j == 0u ? tt[0] : j == 1u ? tt[1] : j == 2u ? tt[2] : tt[3];
    uint32_t k = k0 + j;
    uint32_t ord = 0;
    if (!(t < bt)) {
      // equal parameters: the sequential loop keeps the record it met first
      if (!(t == bt) || !found) continue;
      ord = __ldg(S.lb_order + k);
      if (!(ord < border)) continue;
    } else {
      ord = __ldg(S.lb_order + k);
    }
    float u, v;
    if (STATS) c.n_bary++;
    if (tri_inside(S.lb_aux + 3 * (size_t)k, s + d * t, u, v)) {
      bt = t;
      border = ord;
      found = true;
      bu = u;
      bv = v;
    }
  }
}

// Leaf query.  `stack` is this lane's column of a shared-memory stack (stride
// FS2_LB_THREADS entries).  Returns true when the leaf improved `h`.
template <bool STATS>
__device__ __forceinline__ bool lbvh_leaf(const DScene& S, uint32_t node, uint32_t first,
                                          uint32_t count, const Ray& r, const LbRay& q,
                                          uint32_t object, Hit& h, Counters& c, uint2* stack) {
  if (STATS) c.n_tri += count;  // reference-equivalent work: the loop it replaces tests them all
  const V3 s = r.start, d = r.dir;
  float bt = h.param, bu = 0.f, bv = 0.f;
  uint32_t border = 0;
  bool found = false;
  uint32_t root = __ldg(&S.lb_leaf[node].x);
  if (root == 0xFFFFFFFFu) {
    lb_test_tris<STATS>(S, first, count, s, d, bt, border, found, bu, bv, c);
  } else {
    int sp = 0;
    uint32_t cur = root;
    for (;;) {
      const float4* nd = S.lb_nodes + 8 * (size_t)cur;
      float4 lx = __ldg(nd), ly = __ldg(nd + 1), lz = __ldg(nd + 2);
      float4 hx = __ldg(nd + 3), hy = __ldg(nd + 4), hz = __ldg(nd + 5);
      uint4 meta = __ldg((const uint4*)(nd + 6));
      if (STATS) c.n_lbnode++;
      float lim = fminf(bt, 1.0f);
      float t0 = lb_child(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, s, q, lim);
      float t1 = lb_child(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, s, q, lim);
      float t2 = lb_child(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, s, q, lim);
      float t3 = lb_child(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, s, q, lim);
      if (meta.y) {
        // children are groups of <= 4 triangles: meta.x = first triangle, meta.z = total
        if (t0 < 3.0f) lb_test_tris<STATS>(S, meta.x, min(4u, meta.z), s, d, bt, border, found, bu, bv, c);
        if (t1 < 3.0f && t1 <= fminf(bt, 1.0f))
          lb_test_tris<STATS>(S, meta.x + 4, min(4u, meta.z - 4), s, d, bt, border, found, bu, bv, c);
        if (t2 < 3.0f && t2 <= fminf(bt, 1.0f))
          lb_test_tris<STATS>(S, meta.x + 8, min(4u, meta.z - 8), s, d, bt, border, found, bu, bv, c);
        if (t3 < 3.0f && t3 <= fminf(bt, 1.0f))
          lb_test_tris<STATS>(S, meta.x + 12, min(4u, meta.z - 12), s, d, bt, border, found, bu, bv, c);
      } else {
        // sort the (entry, child) pairs, nearest first; misses (3.0) sink to the end
        uint32_t c0 = meta.x, c1 = meta.x + 1, c2 = meta.x + 2, c3 = meta.x + 3;
#define FS2_CSWAP(ta, ca, tb, cb)  \
  if (tb < ta) {                   \
    float tt = ta;                 \
    ta = tb;                       \
    tb = tt;                       \
    uint32_t cc = ca;              \
    ca = cb;                       \
    cb = cc;                       \
  }
        FS2_CSWAP(t0, c0, t1, c1)
        FS2_CSWAP(t2, c2, t3, c3)
        FS2_CSWAP(t0, c0, t2, c2)
        FS2_CSWAP(t1, c1, t3, c3)
        FS2_CSWAP(t1, c1, t2, c2)
#undef FS2_CSWAP
        if (t3 < 3.0f) stack[FS2_LB_THREADS * sp++] = make_uint2(c3, __float_as_uint(t3));
        if (t2 < 3.0f) stack[FS2_LB_THREADS * sp++] = make_uint2(c2, __float_as_uint(t2));
        if (t1 < 3.0f) stack[FS2_LB_THREADS * sp++] = make_uint2(c1, __float_as_uint(t1));
        if (t0 < 3.0f) {
          cur = c0;
          continue;
        }
      }
      // pop the next pending node that can still matter
      bool more = false;
      while (sp > 0) {
        uint2 e = stack[FS2_LB_THREADS * --sp];
        if (__uint_as_float(e.y) <= fminf(bt, 1.0f)) {
          cur = e.x;
          more = true;
          break;
        }
      }
      if (!more) break;
    }
  }
  if (!found) return false;
  h.param = bt;
  h.object = object;
  h.prim = border;  // original packed index: make_surface() reads the unpermuted arrays
  h.u = bu;
  h.v = bv;
  return true;
}

}  // namespace fs2
