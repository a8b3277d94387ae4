// fs2_lbvh_build.h — builds the per-leaf 4-wide BVH ("leaf BVH") at scene upload.
//
// The reference's mesh octree stops at depth 4 (source/mesh.cpp:15-16,39-44), so its leaves
// hold hundreds to thousands of face references that MeshBvhNode::Trace scans linearly
// (source/mesh.cpp:89-110).  That loop is a pure function of the leaf: the accepted triangle
// with the smallest parameter below the incoming best, first in leaf order on ties.  This
// builder re-orders each leaf's packed records along a Morton curve and stacks an implicit
// 4-ary tree of conservative boxes on top, so a ray only runs the reference's triangle test
// on the groups of four it can actually touch.  Device-layout detail, not part of the ABI.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>

#include "../../include/fs2_cabi.h"

namespace fs2 {

// Leaves with fewer triangles than this are scanned directly (one node, no tree).
#define FS2_LBVH_MAX_LEVELS 7  // 4^(7+1) = 65 536 triangles per octree leaf
#ifndef FS2_LBVH_SAH_SLACK
#define FS2_LBVH_SAH_SLACK 2  // order 3: inner levels beyond the fewest that hold the leaf (room for uneven cuts)
#endif
#ifndef FS2_LBVH_SAH_STEP
#define FS2_LBVH_SAH_STEP 1u  // order 3: cut positions are multiples of this many records
#endif
#ifndef FS2_LBVH_SAH_MAX
#define FS2_LBVH_SAH_MAX 4096u  // order 2: ranges up to this size try all three cut axes
#endif
#ifndef FS2_LBVH_ORDER
#define FS2_LBVH_ORDER 3  // leaf build: 0 Morton curve, 1 median cuts on the longest axis, 2 median cuts on the
                          // axis with the smallest half areas (implicit complete trees), 3 surface-area cuts
#endif

// 128-byte node: boxes of its 4 children as SoA, then the link word.
struct LbNode {
  float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];
  uint32_t child;  // inner: index of child node 0 (children contiguous); quad level: first triangle
  uint32_t kind;   // 0 = children are nodes, 1 = children are groups of <= 4 triangles
  uint32_t count;  // quad level: number of triangles under this node (<= 16); inner: children
  uint32_t pad;
  uint32_t spare[4];
};
static_assert(sizeof(LbNode) == 128, "LbNode must be 128 bytes");

struct LbLeaf {
  uint32_t root;   // LbNode index of the leaf's root, 0xFFFFFFFF = no tree (scan first/count)
  uint32_t first;  // first record in the permuted arrays (== fs_node.first)
};

struct LbvhHost {
  std::vector<float> plane;        // permuted copy of tri_plane (4 floats per record)
  std::vector<fs_tri_aux> aux;     // permuted copy of tri_aux
  std::vector<uint32_t> order;     // permuted index -> original packed index
  std::vector<LbNode> nodes;
  std::vector<LbLeaf> leaf;        // one per octree node
  std::vector<uint8_t> mesh_leaf;  // 1 for nodes that are leaves of a mesh tree
  float m_floor = 1.0f;            // scene coordinate scale the record pads were sized for
  float origin_limit() const { return 3.0f * m_floor; }
};

namespace lbvh_detail {

struct Box3 {
  float lo[3], hi[3];
  void reset() {
    lo[0] = lo[1] = lo[2] = INFINITY;
    hi[0] = hi[1] = hi[2] = -INFINITY;
  }
  void add(const float p[3]) {
    for (int k = 0; k < 3; k++) {
      if (p[k] < lo[k]) lo[k] = p[k];
      if (p[k] > hi[k]) hi[k] = p[k];
    }
  }
  void add(const Box3& b) {
    for (int k = 0; k < 3; k++) {
      if (b.lo[k] < lo[k]) lo[k] = b.lo[k];
      if (b.hi[k] > hi[k]) hi[k] = b.hi[k];
    }
  }
};

// The three vertices of a packed record.  fs_tri_aux keeps p1 and the two kept components
// of the edges (include/fs2_pack.h); the dropped component of p2/p3 is recovered from the
// face plane, whose dominant axis is the dropped one (|n_drop| >= |n|/sqrt(3)).
inline void record_vertices(const float pl[4], const fs_tri_aux& a, float v[4][3]) {
  int drop = (int)(a.face >> 30);
  int ia = (drop == 0) ? 1 : 0, ib = (drop == 2) ? 1 : 2;
  for (int k = 0; k < 3; k++) v[0][k] = v[1][k] = v[2][k] = v[3][k] = a.p1[k];
  v[1][ia] = a.p1[ia] + a.v0a;
  v[1][ib] = a.p1[ib] + a.v0b;
  v[2][ia] = a.p1[ia] + a.v1a;
  v[2][ib] = a.p1[ib] + a.v1b;
  double nd = pl[drop];
  if (nd != 0.0) {
    // v[3] = p1 moved onto the stored plane: an accepted hit lies on the plane, so its dropped
    // coordinate is a blend of the plane heights at the three projected corners
    for (int j = 1; j < 4; j++)
      v[j][drop] = (float)(-((double)pl[3] + (double)pl[ia] * v[j][ia] + (double)pl[ib] * v[j][ib]) / nd);
  }
}

// Conservative box of one record: it must contain every point the reference's triangle test
// (source/math/intersect.cpp:357-419 on the packed constants) can accept, for every segment the
// walk is used for.  Three sources of slack, each bounded here:
//
//  (a) The accepted region is NOT the geometric triangle but {u' >= 0, v' >= 0, u' + v' <= 1}
//      of the float barycentrics u', v' computed from the stored dot00/01/11 and inv_denom.
//      For well-shaped triangles the two coincide to rounding; for needles (aspect ratio r) the
//      stored denominator carries a relative error ~ r^2 * 2^-24, so the region is a scaled /
//      sheared copy of the triangle, or — beyond r ~ 10^3 — has nothing to do with it.  The
//      region's corners are therefore solved (in double) from the stored constants themselves,
//      widened by the evaluation error eta of u', v'; when eta is not small or the linear map
//      is singular the record gets an unbounded box (it is then tested by every segment that
//      reaches its leaf: slow, exact).
//  (b) The hit point p = start + dir * t is rounded: it lies within ~2^-19 (|start| + |p|) of
//      the exact segment and of the stored plane (the plane residual is bounded by the same
//      ulp-scale terms, independent of the grazing angle, and is divided by the dominant normal
//      component: factor <= sqrt 3).  |p| is bounded by the record's own coordinates; |start|
//      is not known here, so the pad is 2^-17 * max(record magnitude, m_floor) with m_floor the
//      scene's coordinate scale, and the kernels leave the leaf BVH for the reference's linear
//      scan whenever |start|_inf > 3 * m_floor (DScene::lb_origin_limit; 4 * 2^-19 = 2^-17).
//      fs_trace / fs_first_hit rebuild the leaf BVHs with a larger m_floor when the camera itself
//      lies beyond the limit, so only bounce rays that start on far-away points of unbounded
//      primitives ever take the linear path.
//  (c) 0.1 % of the extent on top, which also absorbs the vp = p - p1 rounding.
inline Box3 record_box(const float pl[4], const fs_tri_aux& a, float m_floor) {
  float v[4][3];
  record_vertices(pl, a, v);
  Box3 b;
  b.reset();
  for (int j = 0; j < 4; j++) b.add(v[j]);
  auto everything = [&]() {
    for (int k = 0; k < 3; k++) {
      b.lo[k] = -INFINITY;
      b.hi[k] = INFINITY;
    }
    return b;
  };
  // ---- (a) corners of the reference's accepted region, from the stored constants
  const int drop = (int)(a.face >> 30);
  const int ia = (drop == 0) ? 1 : 0, ib = (drop == 2) ? 1 : 2;
  const double inv = a.inv_denom, d00 = a.dot00, d01 = a.dot01, d11 = a.dot11;
  const double au = (d11 * a.v0a - d01 * a.v1a) * inv, bu = (d11 * a.v0b - d01 * a.v1b) * inv;
  const double av = (d00 * a.v1a - d01 * a.v0a) * inv, bv = (d00 * a.v1b - d01 * a.v0b) * inv;
  const double det = au * bv - bu * av;
  if (!(det == det) || isinf(det) || det == 0.0 || !(inv == inv) || isinf(inv)) {
    // inv_denom = inf / nan: u', v' are inf / nan wherever they are not exactly 0, so nothing but
    // a measure-zero set is accepted; a singular map means an unbounded accepted strip
    if (!(inv == inv) || isinf(inv)) {
      // keep the vertex hull: harmless and finite
    } else {
      return everything();
    }
  } else {
    const double q1x = bv / det, q1y = -av / det;  // (u', v') = (1, 0)
    const double q2x = -bu / det, q2y = au / det;  // (u', v') = (0, 1)
    // evaluation error of u' / v' at the far corners (products and sums of the float expression)
    double eta = 0.0;
    const double qs[2][2] = {{q1x, q1y}, {q2x, q2y}};
    for (int c = 0; c < 2; c++) {
      const double x = fabs(qs[c][0]), y = fabs(qs[c][1]);
      const double m02 = fabs((double)a.v0a) * x + fabs((double)a.v0b) * y;
      const double m12 = fabs((double)a.v1a) * x + fabs((double)a.v1b) * y;
      const double eu = (fabs(d11) * m02 + fabs(d01) * m12) * fabs(inv);
      const double ev = (fabs(d00) * m12 + fabs(d01) * m02) * fabs(inv);
      eta = fmax(eta, fmax(eu, ev));
    }
    eta *= 4.76837158203125e-07;  // 2^-21
    if (!(eta == eta) || eta > 0.125) return everything();
    const double e = 2.0 * eta;
    const double uv[3][2] = {{-e, -e}, {1.0 + 3.0 * e, -e}, {-e, 1.0 + 3.0 * e}};
    const double nd = pl[drop];
    for (int c = 0; c < 3; c++) {
      float p[3];
      const double x = uv[c][0] * q1x + uv[c][1] * q2x, y = uv[c][0] * q1y + uv[c][1] * q2y;
      p[ia] = (float)((double)a.p1[ia] + x);
      p[ib] = (float)((double)a.p1[ib] + y);
      p[drop] = a.p1[drop];
      if (nd != 0.0) p[drop] = (float)(-((double)pl[3] + (double)pl[ia] * p[ia] + (double)pl[ib] * p[ib]) / nd);
      for (int k = 0; k < 3; k++)
        if (!(p[k] == p[k]) || isinf(p[k])) return everything();
      b.add(p);
    }
  }
  // ---- (b), (c)
  float m = m_floor, ext = 0.f;
  for (int k = 0; k < 3; k++) {
    m = fmaxf(m, fmaxf(fabsf(b.lo[k]), fabsf(b.hi[k])));
    ext = fmaxf(ext, b.hi[k] - b.lo[k]);
  }
  float pad = m * 7.62939453125e-06f /* 2^-17 */ + ext * 1.0e-3f + 1.0e-30f;
  if (!(pad == pad) || isinf(pad)) return everything();
  for (int k = 0; k < 3; k++) {
    b.lo[k] -= pad;
    b.hi[k] += pad;
  }
  return b;
}

// Coordinate scale of the scene: the largest finite |coordinate| of any bounded object's bounds
// or mesh-tree root box (at least 1; PlanarObject's synthetic +-1000 box is skipped).  Ray
// origins — the camera, or points on objects — are expected within a few multiples of it;
// DScene::lb_origin_limit = 3 * m_floor.
inline float scene_coordinate_scale(const fs_scene_desc* s) {
  float m = 1.0f;
  auto take = [&](float c) {
    c = fabsf(c);
    if (c == c && c <= 1.0e6f && c > m) m = c;
  };
  for (uint32_t o = 0; o < s->n_objects; o++) {
    const fs_object& ob = s->objects[o];
    if (ob.type == FS_OBJ_PLANE) continue;
    for (int k = 0; k < 3; k++) {
      take(ob.bmin[k]);
      take(ob.bmax[k]);
    }
    if (ob.type == FS_OBJ_MESH && ob.u[0] < s->n_nodes) {
      const fs_node& nd = s->nodes[ob.u[0]];
      for (int k = 0; k < 3; k++) {
        take(nd.bmin[k]);
        take(nd.bmax[k]);
      }
    }
  }
  return m;
}

inline uint32_t spread3(uint32_t x) {
  x &= 0x3FFu;
  x = (x | (x << 16)) & 0x030000FFu;
  x = (x | (x << 8)) & 0x0300F00Fu;
  x = (x | (x << 4)) & 0x030C30C3u;
  x = (x | (x << 2)) & 0x09249249u;
  return x;
}

}  // namespace lbvh_detail

namespace lbvh_detail {

struct LeafJob {
  uint32_t node;        // octree node index
  uint32_t first, count;
  uint32_t node_base;   // first LbNode of this leaf (root)
  uint32_t n_nodes;
};

// Number of LbNodes of a leaf with n records: levels of ceil-div by 4 down to one node.
inline uint32_t leaf_node_count(uint32_t n) {
  uint32_t q = (n + 3) / 4, total = 0;
  uint32_t m = q;
  do {
    m = (m + 3) / 4;
    total += m;
  } while (m > 1);
  return total;
}

// Builds one leaf: permutes records [first, first+count) and writes its nodes, root first,
// then level by level towards the quad level (children of a node are contiguous).
inline void build_leaf(const fs_scene_desc* s, const LeafJob& job, LbvhHost* out, float m_floor) {
  const uint32_t n = job.count, first = job.first;
  std::vector<Box3> boxes(n);
  std::vector<uint64_t> keys(n);
  Box3 cb;
  cb.reset();
  std::vector<float> cen(3 * (size_t)n);
  for (uint32_t i = 0; i < n; i++) {
    const float* pl = s->tri_plane + 4 * (size_t)(first + i);
    boxes[i] = record_box(pl, s->tri_aux[first + i], m_floor);
    for (int k = 0; k < 3; k++) {
      float c = 0.5f * (boxes[i].lo[k] + boxes[i].hi[k]);
      if (!(c == c) || isinf(c)) c = 0.f;
      cen[3 * (size_t)i + k] = c;
    }
    cb.add(&cen[3 * (size_t)i]);
  }
#if FS2_LBVH_ORDER == 0
  // Morton order of the record centres (10 bits per axis)
  float scale[3];
  for (int k = 0; k < 3; k++) {
    float e = cb.hi[k] - cb.lo[k];
    scale[k] = e > 0.f ? 1023.0f / e : 0.f;
  }
  for (uint32_t i = 0; i < n; i++) {
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
      float f = (cen[3 * (size_t)i + k] - cb.lo[k]) * scale[k];
      q[k] = (uint32_t)fminf(fmaxf(f, 0.f), 1023.f);
    }
    uint32_t code = spread3(q[0]) | (spread3(q[1]) << 1) | (spread3(q[2]) << 2);
    keys[i] = ((uint64_t)code << 32) | i;
  }
  std::sort(keys.begin(), keys.end());
#else
  // Median-split order: the implicit tree's node i of level l covers the records
  // [4^(l+1) i, 4^(l+1) (i+1)), so the permutation is chosen top down — the range of a node is cut
  // at its children's block boundaries along the longest axis of its centres (two binary cuts
  // per node) — and every aligned block ends up a compact patch of the surface.  Measured against
  // the Morton order on config 2: see DESIGN.md 4.3.
  {
    std::vector<uint32_t> perm(n);
    for (uint32_t i = 0; i < n; i++) perm[i] = i;
    auto cut = [&](uint32_t lo, uint32_t mid, uint32_t hi) {
      if (mid <= lo || mid >= hi) return;
      float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
      for (uint32_t i = lo; i < hi; i++)
        for (int k = 0; k < 3; k++) {
          float v = cen[3 * (size_t)perm[i] + k];
          mn[k] = fminf(mn[k], v);
          mx[k] = fmaxf(mx[k], v);
        }
      auto less_on = [&](int ax) {
        return [&cen, ax](uint32_t a, uint32_t b) {
          float va = cen[3 * (size_t)a + ax], vb = cen[3 * (size_t)b + ax];
          return va < vb || (va == vb && a < b);
        };
      };
      int ax = 0;
      if (mx[1] - mn[1] > mx[ax] - mn[ax]) ax = 1;
      if (mx[2] - mn[2] > mx[ax] - mn[ax]) ax = 2;
#if FS2_LBVH_ORDER == 2
      // try the three axes, keep the cut whose halves have the smallest total surface area
      if (hi - lo <= FS2_LBVH_SAH_MAX) {
        double best = INFINITY;
        std::vector<uint32_t> tmp(perm.begin() + lo, perm.begin() + hi), keep;
        for (int a = 0; a < 3; a++) {
          std::vector<uint32_t> t2 = tmp;
          std::nth_element(t2.begin(), t2.begin() + (mid - lo), t2.end(), less_on(a));
          Box3 l, r;
          l.reset();
          r.reset();
          for (uint32_t i = 0; i < t2.size(); i++) (i < mid - lo ? l : r).add(boxes[t2[i]]);
          auto area = [](const Box3& b) {
            double x = (double)b.hi[0] - b.lo[0], y = (double)b.hi[1] - b.lo[1], z = (double)b.hi[2] - b.lo[2];
            double v = x * y + y * z + z * x;
            return v == v && !isinf(v) ? v : 1e300;
          };
          double cost = area(l) * (mid - lo) + area(r) * (hi - mid);
          if (cost < best) {
            best = cost;
            keep.swap(t2);
          }
        }
        std::copy(keep.begin(), keep.end(), perm.begin() + lo);
        return;
      }
#endif
      std::nth_element(perm.begin() + lo, perm.begin() + mid, perm.begin() + hi, less_on(ax));
    };
    uint32_t top = 4;  // records under one child of the root
    while ((uint64_t)4 * top < n) top *= 4;
    struct Range {
      uint32_t lo, hi, block;
    };
    std::vector<Range> todo;
    todo.push_back({0u, n, top});
    while (!todo.empty()) {
      Range r = todo.back();
      todo.pop_back();
      const uint32_t b = r.block;
      cut(r.lo, r.lo + 2 * b, r.hi);
      cut(r.lo, r.lo + b, r.lo + 2 * b < r.hi ? r.lo + 2 * b : r.hi);
      if (r.lo + 2 * b < r.hi) cut(r.lo + 2 * b, r.lo + 3 * b, r.hi);
      if (b > 4)
        for (uint32_t c = 0; c < 4 && r.lo + c * b < r.hi; c++)
          todo.push_back({r.lo + c * b, r.lo + (c + 1) * b < r.hi ? r.lo + (c + 1) * b : r.hi, b / 4});
    }
    for (uint32_t i = 0; i < n; i++) keys[i] = perm[i];
  }
#endif
  std::vector<Box3> cur(n);
  for (uint32_t i = 0; i < n; i++) {
    uint32_t src = (uint32_t)(keys[i] & 0xFFFFFFFFu);
    memcpy(&out->plane[4 * (size_t)(first + i)], s->tri_plane + 4 * (size_t)(first + src), 16);
    out->aux[first + i] = s->tri_aux[first + src];
    out->order[first + i] = first + src;
    cur[i] = boxes[src];
  }
  // level sizes, bottom (quad level) to top
  uint32_t q = (n + 3) / 4;
  std::vector<Box3> item(q);  // boxes of the items the next level groups: quads first
  for (uint32_t g = 0; g < q; g++) {
    item[g].reset();
    for (uint32_t j = 4 * g; j < 4 * g + 4 && j < n; j++) item[g].add(cur[j]);
  }
  std::vector<uint32_t> sizes;
  for (uint32_t m = q;;) {
    m = (m + 3) / 4;
    sizes.push_back(m);
    if (m <= 1) break;
  }
  // offsets: root level first
  const int L = (int)sizes.size();
  std::vector<uint32_t> base(L);
  uint32_t acc = job.node_base;
  for (int l = L - 1; l >= 0; l--) {
    base[l] = acc;
    acc += sizes[l];
  }
  for (int l = 0; l < L; l++) {
    std::vector<Box3> next(sizes[l]);
    const uint32_t n_items = (uint32_t)item.size();
    for (uint32_t i = 0; i < sizes[l]; i++) {
      LbNode& nd = out->nodes[base[l] + i];
      memset(&nd, 0, sizeof nd);
      next[i].reset();
      uint32_t kids = 0;
      for (uint32_t c = 0; c < 4; c++) {
        uint32_t j = 4 * i + c;
        if (j < n_items) {
          nd.lox[c] = item[j].lo[0];
          nd.loy[c] = item[j].lo[1];
          nd.loz[c] = item[j].lo[2];
          nd.hix[c] = item[j].hi[0];
          nd.hiy[c] = item[j].hi[1];
          nd.hiz[c] = item[j].hi[2];
          next[i].add(item[j]);
          kids++;
        } else {  // empty slot: never hit
          nd.lox[c] = nd.loy[c] = nd.loz[c] = INFINITY;
          nd.hix[c] = nd.hiy[c] = nd.hiz[c] = INFINITY;
        }
      }
      if (l == 0) {
        nd.kind = 1;
        nd.child = first + 16 * i;
        uint32_t rem = n - 16 * i;
        nd.count = rem < 16 ? rem : 16;
      } else {
        nd.kind = 0;
        nd.child = base[l - 1] + 4 * i;
        nd.count = kids;
      }
    }
    item.swap(next);
  }
}


// ------------------------------------------------------------------ order 3: surface-area build
// Top-down build with explicit child links (the kernels only ever follow LbNode::child, so the
// tree need not be the implicit complete one): every inner node cuts its records twice — first
// in two, then each half in two — at the position and axis with the smallest
// area(L) * groups(L) + area(R) * groups(R) (full sweep over the three centre orders), subject to
// the depth budget: a node with `levels` inner levels left below it holds at most 16 * 4^levels
// records, so the walk's stack bound (3 pending siblings per inner level, FS2_PH_LB_STACK) holds
// for any shape.  Ranges above FS2_LBVH_SAH_MAX are cut at the median of the longest axis.
struct SahLeafBuilder {
  const std::vector<Box3>& boxes;
  const std::vector<float>& cen;
  std::vector<uint32_t>& perm;   // positions -> record (local index); = ord[0] when the build ends
  std::vector<LbNode>& nodes;    // local node array, root = 0, child links local
  uint32_t first;                // first record of the leaf in the permuted arrays
  // The records in each of the three centre orders (ties by record index).  Every range [lo, hi)
  // of a node holds the same records in all three; a cut marks its left side and partitions the
  // other two orders stably, so they stay sorted and no split ever sorts again.
  std::vector<uint32_t> ord[3];
  std::vector<uint32_t> tmp;
  std::vector<uint8_t> left;
  std::vector<double> rarea;
  std::vector<Box3> gathered;

  // The records in the three centre orders, ties by record index.  Leaves hold up to 65 536
  // records: a stable least-significant-digit radix sort over the order-preserving integer image
  // of the key (-0 folded onto +0, as the float comparison sees them) gives the same permutation
  // as a comparison sort by (key, record) in a third of the time.
  void presort() {
    const uint32_t n = (uint32_t)perm.size();
    std::vector<uint32_t> key(n), a(n), b(n);
    uint32_t count[2048];
    for (int ax = 0; ax < 3; ax++) {
      for (uint32_t i = 0; i < n; i++) {
        const float f = cen[3 * (size_t)i + ax] + 0.0f;  // -0.0f + 0.0f = +0.0f
        uint32_t u;
        memcpy(&u, &f, 4);
        key[i] = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
        a[i] = i;
      }
      if (n < 1024) {  // (three 2 048-bucket histograms cost more than the sort of a small leaf)
        std::sort(a.begin(), a.end(), [&](uint32_t x, uint32_t y) { return key[x] < key[y] || (key[x] == key[y] && x < y); });
      } else {
        const int shift[3] = {0, 11, 22}, bits[3] = {11, 11, 10};
        for (int pass = 0; pass < 3; pass++) {
          const uint32_t mask = (1u << bits[pass]) - 1u;
          memset(count, 0, sizeof(uint32_t) << bits[pass]);
          for (uint32_t i = 0; i < n; i++) count[(key[a[i]] >> shift[pass]) & mask]++;
          uint32_t sum = 0;
          for (uint32_t d = 0; d <= mask; d++) {
            const uint32_t c = count[d];
            count[d] = sum;
            sum += c;
          }
          for (uint32_t i = 0; i < n; i++) b[count[(key[a[i]] >> shift[pass]) & mask]++] = a[i];
          a.swap(b);
        }
      }
      ord[ax] = a;
    }
    left.assign(n, 0);
    tmp.resize(n);
    gathered.resize(n);
  }
  static double area(const Box3& b) {
    double x = (double)b.hi[0] - b.lo[0], y = (double)b.hi[1] - b.lo[1], z = (double)b.hi[2] - b.lo[2];
    double v = x * y + y * z + z * x;
    return (v == v && !isinf(v)) ? v : 1.0e300;
  }
  Box3 range_box(uint32_t lo, uint32_t hi) const {
    Box3 b;
    b.reset();
    for (uint32_t i = lo; i < hi; i++) b.add(boxes[ord[0][i]]);
    return b;
  }
  // Cuts the range [lo, hi) into [lo, mid) and [mid, hi) with kmin <= mid - lo <= kmax (and both
  // sides non-empty); `step` restricts the left size to multiples of it (group alignment).
  uint32_t split(uint32_t lo, uint32_t hi, uint32_t kmin, uint32_t kmax, uint32_t step) {
    const uint32_t n = hi - lo;
    if (kmin < 1) kmin = 1;
    if (kmax > n - 1) kmax = n - 1;
    if (step > 1) {
      kmin = (kmin + step - 1) / step * step;
      kmax = kmax / step * step;
    }
    if (kmin > kmax) return lo + (n + 1) / 2;  // cannot happen for feasible budgets
    int best_ax = 0;
    uint32_t best_k = kmin;
    if (n > FS2_LBVH_SAH_MAX) {
      // median of the longest axis of the centres
      float ext[3];
      for (int ax = 0; ax < 3; ax++)
        ext[ax] = cen[3 * (size_t)ord[ax][hi - 1] + ax] - cen[3 * (size_t)ord[ax][lo] + ax];
      if (ext[1] > ext[best_ax]) best_ax = 1;
      if (ext[2] > ext[best_ax]) best_ax = 2;
      uint32_t k = (n / 2 + step - 1) / step * step;
      best_k = k < kmin ? kmin : k > kmax ? kmax : k;
    } else {
      double best = INFINITY;
      rarea.resize(n + 1);
      for (int ax = 0; ax < 3; ax++) {
        // one gather of the boxes in this order, then both passes run over contiguous memory
        const uint32_t* o = ord[ax].data() + lo;
        Box3* g = gathered.data();
        for (uint32_t i = 0; i < n; i++) g[i] = boxes[o[i]];
        Box3 b;
        b.reset();
        for (uint32_t i = n; i-- > kmin;) {  // rarea[k] is read for k >= kmin only
          b.add(g[i]);
          rarea[i] = area(b);
        }
        b.reset();
        for (uint32_t k = 1; k <= kmax; k++) {
          b.add(g[k - 1]);
          if (k < kmin || (step > 1 && k % step)) continue;
          double cost = area(b) * ((k + 3) / 4) + rarea[k] * ((n - k + 3) / 4);
          if (cost < best) {
            best = cost;
            best_k = k;
            best_ax = ax;
          }
        }
      }
    }
    // the chosen order's first best_k records go left; the other orders follow, stably
    for (uint32_t i = 0; i < best_k; i++) left[ord[best_ax][lo + i]] = 1;
    for (int ax = 0; ax < 3; ax++) {
      if (ax == best_ax) continue;
      uint32_t* o = ord[ax].data() + lo;
      uint32_t l = 0, r = 0;
      for (uint32_t i = 0; i < n; i++) {
        if (left[o[i]])
          o[l++] = o[i];
        else
          tmp[r++] = o[i];
      }
      std::copy(tmp.begin(), tmp.begin() + r, o + l);
    }
    for (uint32_t i = 0; i < best_k; i++) left[ord[best_ax][lo + i]] = 0;
    return lo + best_k;
  }
  // orders the <= 16 records of a quad-level node into groups of four
  void split_groups(uint32_t lo, uint32_t hi) {
    if (hi - lo <= 4) return;
    uint32_t mid = split(lo, hi, 4, hi - lo - 1, 4);
    split_groups(lo, mid);
    split_groups(mid, hi);
  }
  static uint64_t cap(int levels) { return (uint64_t)16 << (2 * levels); }

  void build(uint32_t self, uint32_t lo, uint32_t hi, int levels) {
    const uint32_t n = hi - lo;
    LbNode nd;
    memset(&nd, 0, sizeof nd);
    for (int c = 0; c < 4; c++) {
      nd.lox[c] = nd.loy[c] = nd.loz[c] = INFINITY;
      nd.hix[c] = nd.hiy[c] = nd.hiz[c] = INFINITY;
    }
    auto set_box = [&](int c, const Box3& b) {
      nd.lox[c] = b.lo[0];
      nd.loy[c] = b.lo[1];
      nd.loz[c] = b.lo[2];
      nd.hix[c] = b.hi[0];
      nd.hiy[c] = b.hi[1];
      nd.hiz[c] = b.hi[2];
    };
    if (n <= 16) {
      split_groups(lo, hi);
      nd.kind = 1;
      nd.child = first + lo;
      nd.count = n;
      for (uint32_t g = 0; 4 * g < n; g++) set_box((int)g, range_box(lo + 4 * g, lo + 4 * g + 4 < hi ? lo + 4 * g + 4 : hi));
      nodes[self] = nd;
      return;
    }
    const uint64_t cc = cap(levels - 1);  // most records a child may hold
    auto lim = [](uint64_t v) { return (uint32_t)(v > 0xFFFFFFFFull ? 0xFFFFFFFFull : v); };
    // first cut: both halves must fit two children
    uint32_t mid = split(lo, hi, n > 2 * cc ? lim(n - 2 * cc) : 1, lim(2 * cc), FS2_LBVH_SAH_STEP);
    uint32_t r[5] = {lo, lo, mid, mid, hi};
    auto half = [&](uint32_t a, uint32_t b, uint32_t& m) {
      uint32_t h = b - a;
      if (h < 2) {
        m = b;  // one child only
        return;
      }
      m = split(a, b, h > cc ? lim(h - cc) : 1, lim(cc), FS2_LBVH_SAH_STEP);
    };
    half(lo, mid, r[1]);
    half(mid, hi, r[3]);
    // children: [r0,r1) [r1,r2) [r2,r3) [r3,r4); empty ranges keep the never-hit box
    nd.kind = 0;
    nd.child = (uint32_t)nodes.size();
    nd.count = 0;
    nodes.resize(nodes.size() + 4);
    for (int c = 0; c < 4; c++) memset(&nodes[nd.child + c], 0, sizeof(LbNode));
    for (int c = 0; c < 4; c++) {
      if (r[c + 1] <= r[c]) continue;
      set_box(c, range_box(r[c], r[c + 1]));
      nd.count++;
    }
    nodes[self] = nd;
    const uint32_t child = nd.child;
    for (int c = 0; c < 4; c++)
      if (r[c + 1] > r[c]) build(child + (uint32_t)c, r[c], r[c + 1], levels - 1);
  }
};

// Leaf build of order 3.  Returns the leaf's nodes (root first, child links relative to the root)
// and permutes the records like build_leaf.  Empty result: the leaf stays linear.
inline std::vector<LbNode> build_leaf_sah(const fs_scene_desc* s, const LeafJob& job, LbvhHost* out,
                                          float m_floor) {
  const uint32_t n = job.count, first = job.first;
  std::vector<Box3> boxes(n);
  std::vector<float> cen(3 * (size_t)n);
  for (uint32_t i = 0; i < n; i++) {
    boxes[i] = record_box(s->tri_plane + 4 * (size_t)(first + i), s->tri_aux[first + i], m_floor);
    for (int k = 0; k < 3; k++) {
      float c = 0.5f * (boxes[i].lo[k] + boxes[i].hi[k]);
      if (!(c == c) || isinf(c)) c = 0.f;
      cen[3 * (size_t)i + k] = c;
    }
  }
  std::vector<uint32_t> perm(n);
  for (uint32_t i = 0; i < n; i++) perm[i] = i;
  std::vector<LbNode> nodes(1);
  // levels of inner nodes above the quad level: the fewest that hold n, plus two of slack for
  // uneven cuts, never more than the stack allows
  int levels = 0;
  while (SahLeafBuilder::cap(levels) < n) levels++;
  levels = levels + FS2_LBVH_SAH_SLACK < FS2_LBVH_MAX_LEVELS - 1 ? levels + FS2_LBVH_SAH_SLACK : FS2_LBVH_MAX_LEVELS - 1;
  SahLeafBuilder b{boxes, cen, perm, nodes, first, {}, {}, {}, {}, {}};
  b.presort();
  b.build(0, 0, n, levels);
  perm = b.ord[0];
  if (nodes.size() >= 65536) return {};  // stack entries carry 16-bit node offsets
  for (uint32_t i = 0; i < n; i++) {
    uint32_t src = perm[i];
    memcpy(&out->plane[4 * (size_t)(first + i)], s->tri_plane + 4 * (size_t)(first + src), 16);
    out->aux[first + i] = s->tri_aux[first + src];
    out->order[first + i] = first + src;
  }
  return nodes;
}

}  // namespace lbvh_detail

// Builds the leaf BVHs of every mesh-tree leaf of `s`.  min_leaf: leaves with fewer records
// keep root = 0xFFFFFFFF (scanned linearly).  Leaves beyond the level cap stay linear too.
// m_floor_min: lower bound for the coordinate scale the pads are sized for (the camera's distance
// from the origin when it lies outside the scene's own scale).
inline void build_lbvh(const fs_scene_desc* s, LbvhHost* out, uint32_t min_leaf = 9,
                       unsigned threads = 0, float m_floor_min = 0.0f) {
  using namespace lbvh_detail;
  out->plane.assign(s->tri_plane, s->tri_plane + 4 * (size_t)s->n_tris);
  out->aux.assign(s->tri_aux, s->tri_aux + s->n_tris);
  out->order.resize(s->n_tris);
  for (uint32_t i = 0; i < s->n_tris; i++) out->order[i] = i;
  out->leaf.assign(s->n_nodes, LbLeaf{0xFFFFFFFFu, 0});
  out->mesh_leaf.assign(s->n_nodes, 0);
  // enumerate mesh-tree leaves
  std::vector<LeafJob> jobs;
  std::vector<uint32_t> stack;
  std::vector<uint8_t> seen(s->n_nodes, 0);
  uint32_t total_nodes = 0;
  for (uint32_t o = 0; o < s->n_objects; o++) {
    const fs_object& ob = s->objects[o];
    if (ob.type != FS_OBJ_MESH || ob.u[0] == 0xFFFFFFFFu) continue;
    stack.push_back(ob.u[0]);
    while (!stack.empty()) {
      uint32_t ni = stack.back();
      stack.pop_back();
      if (ni >= s->n_nodes || seen[ni]) continue;
      seen[ni] = 1;
      const fs_node& nd = s->nodes[ni];
      if (nd.count == FS_NODE_INTERNAL) {
        for (uint32_t c = 0; c < 8; c++) stack.push_back(nd.first + c);
        continue;
      }
      out->leaf[ni].first = nd.first;
      out->mesh_leaf[ni] = 1;
      if ((uint64_t)nd.first + nd.count > s->n_tris) continue;  // malformed: stay linear
      if (nd.count < min_leaf || nd.count > (1u << (2 * (FS2_LBVH_MAX_LEVELS + 1)))) continue;
      LeafJob j;
      j.node = ni;
      j.first = nd.first;
      j.count = nd.count;
      j.n_nodes = leaf_node_count(nd.count);
      j.node_base = total_nodes;
      total_nodes += j.n_nodes;
      jobs.push_back(j);
    }
  }
  const float m_floor = out->m_floor = fmaxf(scene_coordinate_scale(s), m_floor_min);
  if (threads == 0) threads = std::thread::hardware_concurrency();
  if (threads == 0) threads = 1;
  if (threads > 32) threads = 32;
  auto run = [&](auto&& body) {
    std::atomic<size_t> next(0);
    auto work = [&]() {
      for (;;) {
        size_t i = next.fetch_add(1);
        if (i >= jobs.size()) break;
        body(i);
      }
    };
    if (threads <= 1 || jobs.size() < 8) {
      work();
    } else {
      std::vector<std::thread> pool;
      for (unsigned t = 0; t < threads; t++) pool.emplace_back(work);
      for (auto& t : pool) t.join();
    }
  };
#if FS2_LBVH_ORDER == 3
  // node counts are known only after the build: leaves are built into their own arrays (fat
  // leaves first, they take longest), then laid out one after the other and their links rebased
  std::vector<std::vector<LbNode>> built(jobs.size());
  std::vector<size_t> by_size(jobs.size());
  for (size_t i = 0; i < jobs.size(); i++) by_size[i] = i;
  std::sort(by_size.begin(), by_size.end(), [&](size_t a, size_t b) {
    return jobs[a].count > jobs[b].count || (jobs[a].count == jobs[b].count && a < b);
  });
  run([&](size_t i) { built[by_size[i]] = build_leaf_sah(s, jobs[by_size[i]], out, m_floor); });
  total_nodes = 0;
  for (size_t i = 0; i < jobs.size(); i++) {
    jobs[i].node_base = total_nodes;
    jobs[i].n_nodes = (uint32_t)built[i].size();
    total_nodes += jobs[i].n_nodes;
  }
  out->nodes.resize(total_nodes ? total_nodes : 1);
  run([&](size_t i) {
    for (uint32_t k = 0; k < jobs[i].n_nodes; k++) {
      LbNode nd = built[i][k];
      if (nd.kind == 0) nd.child += jobs[i].node_base;
      out->nodes[jobs[i].node_base + k] = nd;
    }
    if (jobs[i].n_nodes) out->leaf[jobs[i].node].root = jobs[i].node_base;
    std::vector<LbNode>().swap(built[i]);
  });
#else
  out->nodes.resize(total_nodes ? total_nodes : 1);
  for (const LeafJob& j : jobs) out->leaf[j.node].root = j.node_base;
  run([&](size_t i) { build_leaf(s, jobs[i], out, m_floor); });
#endif
}

}  // namespace fs2
