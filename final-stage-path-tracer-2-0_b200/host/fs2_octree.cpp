// fs2_octree.cpp — host-side construction of the reference's two octrees directly in the
// flattened node layout of include/fs2_cabi.h.
//
//   mesh tree   reference source/mesh.cpp:39-76,117-152 (depth < 4, > 16 faces subdivides,
//               faces duplicated into every child their SAT test touches)
//   scene tree  reference source/scene.cpp:33-74,105-126,216-225
//   children    reference source/bvh.h:108-137 (child i = x | z<<1 | y<<2)
//   SAT         reference source/math/intersect.cpp:193-311 (plane pre-test + 9 edge axes)
//
// Node numbering = the flattening convention of this project: a subdividing node reserves
// 8 consecutive slots, then its children are completed depth-first in order 0..7; leaf
// references are appended in that same depth-first order.
#include "fs2_octree.h"

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <atomic>
#include <thread>

namespace fs2h {

// GPU build switch: fs2h_set_gpu_build(device) / env FS2_GPU_BUILD=<device>; -1 = host build.
int& gpu_build_device() {
  static int dev = [] {
    const char* e = getenv("FS2_GPU_BUILD");
    return e && *e ? atoi(e) : -1;
  }();
  return dev;
}
float& gpu_build_ms() {
  static float ms = 0.f;
  return ms;
}

namespace {

const uint32_t kMaxFaceCountPerNode = 16;   // mesh.cpp:15
const uint32_t kMaxMeshDepth = 4;           // mesh.cpp:16
const uint32_t kMaxObjectCountPerNode = 2;  // scene.cpp:11

// bounds_intersect_plane(vertices, plane), intersect.cpp:211-226
bool box_touches_plane(const Vec3* bv, const Plane& pl) {
  uint32_t count = 0;
  for (int i = 0; i < 8; i++) {
    float d = plane_distance(pl, bv[i]);
    if (near_zero(d)) return true;
    count += (d > 0.0f);
  }
  return count != 0 && count != 8;
}

}  // namespace

// triangle_intersect_bounds, intersect.cpp:228-311
bool triangle_touches_box(Vec3 p1, Vec3 p2, Vec3 p3, const Box& b) {
  static const Vec3 axes[3] = {Vec3(1, 0, 0), Vec3(0, 1, 0), Vec3(0, 0, 1)};
  Vec3 bv[8];
  for (int i = 0; i < 8; i++) {
    int xm = i & 1, ym = (i >> 1) & 1, zm = (i >> 2) & 1;
    bv[i] = Vec3(b.mn.x * (!xm) + b.mx.x * (xm), b.mn.y * (!ym) + b.mx.y * (ym),
                 b.mn.z * (!zm) + b.mx.z * (zm));
  }
  Vec3 tv[3] = {p1, p2, p3};
  Vec3 edges[3] = {p1 - p2, p2 - p3, p3 - p1};
  Plane tp = plane_from(p1, p2, p3);
  if (!box_touches_plane(bv, tp)) return false;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      Vec3 axis = cross(edges[i], axes[j]);
      float tmin = 9999999.0f, tmax = -tmin;
      for (int k = 0; k < 3; k++) {
        float p = dot(tv[k], axis);
        if (p < tmin) tmin = p;
        if (p > tmax) tmax = p;
      }
      float bmin = 9999999.0f, bmax = -bmin;
      for (int k = 0; k < 8; k++) {
        float p = dot(bv[k], axis);
        if (p < bmin) bmin = p;
        if (p > bmax) bmax = p;
      }
      if (bmax < tmin || bmin > tmax) return false;
    }
  return true;
}

// bounds_intersect_bounds, intersect.cpp:26-43 (zero-volume boxes never intersect)
bool box_touches_box(const Box& a, const Box& b) {
  if (0 == a.volume() || 0 == b.volume()) return false;
  Vec3 ac = a.center(), bc = b.center();
  Vec3 ae = a.mx - ac, be = b.mx - bc, d = bc - ac;
  return fabs(d.x) <= (ae.x + be.x) && fabs(d.y) <= (ae.y + be.y) && fabs(d.z) <= (ae.z + be.z);
}

// BaseBvhNode::ConfigureChildren, bvh.h:108-137
void child_boxes(const Box& parent, Box out[8]) {
  Vec3 mn = parent.mn, mx = parent.mx;
  Vec3 c = parent.center();
  Vec3 hx = Vec3(c.x, mn.y, mn.z) - mn;
  Vec3 hy = Vec3(mn.x, c.y, mn.z) - mn;
  Vec3 hz = Vec3(mn.x, mn.y, c.z) - mn;
  Vec3 span = hx + hy + hz;
  (void)mx;
  for (int i = 0; i < 8; i++) {
    Vec3 nmin = mn + hx * (float)(i % 2) + hy * (float)(i >> 2) + hz * (float)((i % 4) >> 1);
    Vec3 nmax = nmin + span;
    Box b;
    b.add(nmin);
    b.add(nmax);
    out[i] = b;
  }
}

static void put_box(fs_node* n, const Box& b) {
  n->bmin[0] = b.mn.x;
  n->bmin[1] = b.mn.y;
  n->bmin[2] = b.mn.z;
  n->bmax[0] = b.mx.x;
  n->bmax[1] = b.mx.y;
  n->bmax[2] = b.mx.z;
}

// ---------------------------------------------------------------- mesh tree
namespace {

struct MeshBuild {
  const std::vector<Vec3>* verts;
  const std::vector<MeshFaceRec>* faces;
  uint32_t face_base;
  FlatArrays* out;
  unsigned threads;
};

// Distributes `list` over the 8 children; faces keep their relative order in every child.
void classify(const MeshBuild& mb, const std::vector<uint32_t>& list, const Box kids[8],
              std::vector<uint32_t> child[8]) {
  size_t n = list.size();
  std::vector<uint8_t> mask(n);
  auto work = [&](size_t lo, size_t hi) {
    for (size_t j = lo; j < hi; j++) {
      const MeshFaceRec& f = (*mb.faces)[list[j]];
      Vec3 v0 = (*mb.verts)[f.v[0]], v1 = (*mb.verts)[f.v[1]], v2 = (*mb.verts)[f.v[2]];
      uint8_t m = 0;
      for (int i = 0; i < 8; i++)
        if (triangle_touches_box(v0, v1, v2, kids[i])) m |= (uint8_t)(1u << i);
      mask[j] = m;
    }
  };
  unsigned t = (n >= 8192) ? mb.threads : 1;
  if (t <= 1) {
    work(0, n);
  } else {
    std::vector<std::thread> pool;
    size_t chunk = (n + t - 1) / t;
    for (unsigned k = 0; k < t; k++) {
      size_t lo = k * chunk, hi = std::min(n, lo + chunk);
      if (lo < hi) pool.emplace_back(work, lo, hi);
    }
    for (auto& th : pool) th.join();
  }
  for (size_t j = 0; j < n; j++)
    for (int i = 0; i < 8; i++)
      if (mask[j] & (1u << i)) child[i].push_back(list[j]);
}

void build_mesh_node(const MeshBuild& mb, uint32_t self, const Box& box, uint32_t depth,
                     std::vector<uint32_t>& list) {
  FlatArrays& F = *mb.out;
  put_box(&F.nodes[self], box);
  if (depth >= kMaxMeshDepth || list.size() <= kMaxFaceCountPerNode) {
    F.nodes[self].first = (uint32_t)F.tri_aux.size();
    F.nodes[self].count = (uint32_t)list.size();
    for (uint32_t fi : list) {
      const MeshFaceRec& f = (*mb.faces)[fi];
      const Vec3& a = (*mb.verts)[f.v[0]];
      const Vec3& b = (*mb.verts)[f.v[1]];
      const Vec3& c = (*mb.verts)[f.v[2]];
      float pa[3] = {a.x, a.y, a.z}, pb[3] = {b.x, b.y, b.z}, pc[3] = {c.x, c.y, c.z};
      float plane[4] = {f.plane.x, f.plane.y, f.plane.z, f.plane.w};
      float out_plane[4];
      fs_tri_aux aux;
      fs_pack_triangle(pa, pb, pc, plane, mb.face_base + fi, out_plane, &aux);
      F.tri_plane.insert(F.tri_plane.end(), out_plane, out_plane + 4);
      F.tri_aux.push_back(aux);
    }
    return;
  }
  Box kids[8];
  child_boxes(box, kids);
  std::vector<uint32_t> child[8];
  classify(mb, list, kids, child);
  std::vector<uint32_t>().swap(list);  // the parent hands its faces over (mesh.cpp:66)
  uint32_t first = (uint32_t)F.nodes.size();
  F.nodes.resize(first + 8);
  F.nodes[self].first = first;
  F.nodes[self].count = FS_NODE_INTERNAL;
  for (int i = 0; i < 8; i++) build_mesh_node(mb, first + i, kids[i], depth + 1, child[i]);
}

}  // namespace

uint32_t build_mesh_tree(const std::vector<Vec3>& verts, std::vector<MeshFaceRec>& faces,
                         uint32_t face_base, FlatArrays* out, unsigned threads) {
  if (verts.empty() || faces.empty()) return 0xFFFFFFFFu;  // MeshBvh::BuildBvh early return
  Box root;
  for (const Vec3& v : verts) root.add(v);
  std::vector<uint32_t> all(faces.size());
  parallel_for(faces.size(), threads, [&](size_t i) {
    MeshFaceRec& f = faces[i];
    if (f.plane.x == 0 && f.plane.y == 0 && f.plane.z == 0 && f.plane.w == 0) {
      Vec3 p0 = verts[f.v[0]], p1 = verts[f.v[1]], p2 = verts[f.v[2]];
      f.plane = plane_from(face_normal(p0, p1, p2), p0);
    }
    all[i] = (uint32_t)i;
  });
  uint32_t root_index = (uint32_t)out->nodes.size();
  if (gpu_build_device() >= 0) {
    // same tree, built by the CUDA library (csrc/fs2_build.cuh); byte-identical output
    std::vector<float> vbuf(3 * verts.size()), pbuf(4 * faces.size());
    std::vector<uint32_t> ibuf(3 * faces.size());
    parallel_for(verts.size(), threads, [&](size_t i) {
      vbuf[3 * i] = verts[i].x;
      vbuf[3 * i + 1] = verts[i].y;
      vbuf[3 * i + 2] = verts[i].z;
    });
    parallel_for(faces.size(), threads, [&](size_t i) {
      for (int k = 0; k < 3; k++) ibuf[3 * i + k] = faces[i].v[k];
      pbuf[4 * i] = faces[i].plane.x;
      pbuf[4 * i + 1] = faces[i].plane.y;
      pbuf[4 * i + 2] = faces[i].plane.z;
      pbuf[4 * i + 3] = faces[i].plane.w;
    });
    fs_mesh_tree t;
    int rc = fs_mesh_tree_build(gpu_build_device(), vbuf.data(), (uint32_t)verts.size(), ibuf.data(),
                                pbuf.data(), (uint32_t)faces.size(), face_base, root_index,
                                (uint32_t)out->tri_aux.size(), &t);
    if (rc == FS_OK) {
      out->nodes.insert(out->nodes.end(), t.nodes, t.nodes + t.n_nodes);
      out->tri_plane.insert(out->tri_plane.end(), t.tri_plane, t.tri_plane + 4 * (size_t)t.n_tris);
      out->tri_aux.insert(out->tri_aux.end(), t.tri_aux, t.tri_aux + t.n_tris);
      gpu_build_ms() += t.build_ms;
      fs_mesh_tree_free(&t);
      return root_index;
    }
    fprintf(stderr, "fs2: GPU octree build failed (%s); building on the host\n", fs_last_error(nullptr));
  }
  out->nodes.resize(root_index + 1);
  MeshBuild mb{&verts, &faces, face_base, out, threads ? threads : 1};
  build_mesh_node(mb, root_index, root, 0, all);
  return root_index;
}

// ---------------------------------------------------------------- scene tree
namespace {

struct SceneBuild {
  const std::vector<Box>* boxes;
  uint32_t max_depth;
  FlatArrays* out;
};

void build_scene_node(const SceneBuild& sb, uint32_t self, const Box& box, uint32_t depth,
                      const std::vector<uint32_t>& list) {
  FlatArrays& F = *sb.out;
  put_box(&F.nodes[self], box);
  bool split = true;
  if (depth >= sb.max_depth) split = false;
  // scene.cpp:39-42: unsigned depth against min(unsigned, double), evaluated in double.
  // (The reference leaves max_tree_depth_ of child nodes uninitialised; children inherit
  // the root's value here, which gives the same trees for every N < 11 586 objects.)
  if (split) {
    double a = (double)(uint32_t)(sb.max_depth - depth);
    double b = (log((double)list.size()) / log(8) + 0.5f) - 2.0f;
    double lim = a < b ? a : b;
    if ((double)depth >= lim) split = false;
  }
  if (split && !(list.size() > kMaxObjectCountPerNode)) split = false;
  if (!split) {
    F.nodes[self].first = (uint32_t)F.scene_refs.size();
    F.nodes[self].count = (uint32_t)list.size();
    for (uint32_t oi : list) F.scene_refs.push_back(oi);
    return;
  }
  Box kids[8];
  child_boxes(box, kids);
  std::vector<uint32_t> child[8];
  for (uint32_t oi : list)
    for (int i = 0; i < 8; i++)
      if (box_touches_box((*sb.boxes)[oi], kids[i])) child[i].push_back(oi);
  uint32_t first = (uint32_t)F.nodes.size();
  F.nodes.resize(first + 8);
  F.nodes[self].first = first;
  F.nodes[self].count = FS_NODE_INTERNAL;
  for (int i = 0; i < 8; i++) build_scene_node(sb, first + i, kids[i], depth + 1, child[i]);
}

}  // namespace

int32_t build_scene_tree(const std::vector<Box>& object_boxes, FlatArrays* out) {
  // Scene::Optimize, scene.cpp:216-225
  int32_t ideal_depth = (log(object_boxes.size()) / log(8) + 0.5) - 2;
  if (!(ideal_depth > 0) || object_boxes.empty()) return -1;
  Box root;
  for (const Box& b : object_boxes) root.add(b);
  std::vector<uint32_t> all(object_boxes.size());
  for (uint32_t i = 0; i < all.size(); i++) all[i] = i;
  uint32_t root_index = (uint32_t)out->nodes.size();
  out->nodes.resize(root_index + 1);
  SceneBuild sb{&object_boxes, (uint32_t)ideal_depth, out};
  build_scene_node(sb, root_index, root, 0, all);
  return (int32_t)root_index;
}

}  // namespace fs2h
