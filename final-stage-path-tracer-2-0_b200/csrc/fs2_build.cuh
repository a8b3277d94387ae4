// fs2_build.cuh — the reference's per-mesh octree built on the GPU, level by level.
//
// What it reproduces: MeshBvhNode::Subdivide, reference
// source/mesh.cpp:39-76,117-152 — a node with more than 16 faces and depth < 4 hands every
// face to each of its 8 children whose box the triangle touches (SAT test
// triangle_intersect_bounds, source/math/intersect.cpp:193-311: plane pre-test + 9 edge
// axes), faces keep their relative order in every child.  The reference does this
// recursively on one CPU thread (every face x 8 children x 4 levels: the wall after tracing
// for multi-million-triangle meshes); here one level is
//
//   k_bd_mask      one thread per (node, face) reference: 8 SAT tests -> child mask,
//                  flags[c][i] for the scan
//   cub exclusive scan over the 8 flag rows (one call): rank of every reference in its
//                  (node, child) list, order preserving
//   k_bd_counts    (split node, child) list sizes -> host (<= 8 x 585 numbers per level);
//                  the host lays out the next level (child boxes with the reference's float
//                  expressions, bvh.h:108-137 — exact, and a few thousand boxes at most)
//   k_bd_scatter   writes the next level's reference array (node major, order preserved)
//   k_bd_pack      leaf references -> (tri_plane, fs_tri_aux) records with fs_pack_triangle
//
// Node numbering and leaf order are the host builder's (host/fs2_octree.cpp): depth first,
// 8 consecutive slots reserved per subdividing node.  The result is byte-identical to the
// host build (tests/test_gpu_build.py); arithmetic uses the same float expressions with
// -fmad=false and IEEE division.
#pragma once
#include "fs2_math.cuh"

namespace fs2 {

struct BdBox {
  float mn[3], mx[3];
};

struct BdNode {
  uint32_t start, count;  // references of this node in the level's array
  int32_t split;          // rank among the subdividing nodes of the level, -1 = leaf
  uint32_t out;           // leaf: first output record
};

// bounds_intersect_plane, intersect.cpp:211-226
__device__ __forceinline__ bool bd_box_touches_plane(const V3* bv, float4 pl) {
  uint32_t count = 0;
  for (int i = 0; i < 8; i++) {
    float d = pl.x * bv[i].x + pl.y * bv[i].y + pl.z * bv[i].z + pl.w;
    if (near_zero(d)) return true;
    count += (d > 0.0f) ? 1u : 0u;
  }
  return count != 0 && count != 8;
}

// triangle_intersect_bounds, intersect.cpp:228-311; tp = plane through (p1, p2, p3)
__device__ __forceinline__ bool bd_triangle_touches_box(const V3* tv, const V3* edges, float4 tp,
                                                        const BdBox& b) {
  V3 bv[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float xm = (float)(i & 1), ym = (float)((i >> 1) & 1), zm = (float)((i >> 2) & 1);
    bv[i] = v3(b.mn[0] * (1.0f - xm) + b.mx[0] * xm, b.mn[1] * (1.0f - ym) + b.mx[1] * ym,
               b.mn[2] * (1.0f - zm) + b.mx[2] * zm);
  }
  if (!bd_box_touches_plane(bv, tp)) return false;
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      V3 unit = v3(j == 0 ? 1.f : 0.f, j == 1 ? 1.f : 0.f, j == 2 ? 1.f : 0.f);
      V3 axis = cross(edges[i], unit);
      float tmin = 9999999.0f, tmax = -tmin;
      for (int k = 0; k < 3; k++) {
        float p = dot(tv[k], axis);
        if (p < tmin) tmin = p;
        if (p > tmax) tmax = p;
      }
      float bmin = 9999999.0f, bmax = -bmin;
#pragma unroll
      for (int k = 0; k < 8; k++) {
        float p = dot(bv[k], axis);
        if (p < bmin) bmin = p;
        if (p > bmax) bmax = p;
      }
      if (bmax < tmin || bmin > tmax) return false;
    }
  }
  return true;
}

__device__ __forceinline__ V3 bd_vertex(const float* verts, uint32_t i) {
  return v3(verts[3 * (size_t)i], verts[3 * (size_t)i + 1], verts[3 * (size_t)i + 2]);
}

// Child mask of every reference of a subdividing node; flags[c * n + i] = bit c.
__global__ void __launch_bounds__(128)
k_bd_mask(uint32_t n, const uint32_t* __restrict__ ref_face, const uint32_t* __restrict__ ref_node,
          const BdNode* __restrict__ nodes, const BdBox* __restrict__ child_box,
          const float* __restrict__ verts, const uint32_t* __restrict__ fidx,
          uint32_t* __restrict__ flags) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t split = nodes[ref_node[i]].split;
  uint32_t mask = 0;
  if (split >= 0) {
    uint32_t f = ref_face[i];
    V3 tv[3] = {bd_vertex(verts, fidx[3 * (size_t)f]), bd_vertex(verts, fidx[3 * (size_t)f + 1]),
                bd_vertex(verts, fidx[3 * (size_t)f + 2])};
    V3 edges[3] = {tv[0] - tv[1], tv[1] - tv[2], tv[2] - tv[0]};
    // plane(p1, p2, p3) = plane(normalize(cross(p2 - p1, p3 - p1)), p1), plane.h:68-83
    V3 nrm = normalize(cross(tv[1] - tv[0], tv[2] - tv[0]));
    float4 tp = make_float4(nrm.x, nrm.y, nrm.z, -dot(nrm, tv[0]));
    const BdBox* kids = child_box + 8 * (size_t)split;
    for (int c = 0; c < 8; c++)
      if (bd_triangle_touches_box(tv, edges, tp, kids[c])) mask |= 1u << c;
  }
#pragma unroll
  for (int c = 0; c < 8; c++) flags[(size_t)c * n + i] = (mask >> c) & 1u;
}

// counts[s * 8 + c] = references node s hands to its child c.  scan has 8 n + 1 entries.
__global__ void k_bd_counts(uint32_t n, uint32_t n_nodes, const BdNode* __restrict__ nodes,
                            const uint32_t* __restrict__ scan, uint32_t* __restrict__ counts) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_nodes * 8u) return;
  const BdNode nd = nodes[t >> 3];
  if (nd.split < 0) return;
  uint32_t c = t & 7u;
  size_t row = (size_t)c * n;
  counts[8 * (size_t)nd.split + c] = scan[row + nd.start + nd.count] - scan[row + nd.start];
}

// Next level's reference array: child lists in (split node, child) order.
__global__ void __launch_bounds__(128)
k_bd_scatter(uint32_t n, const uint32_t* __restrict__ ref_face, const uint32_t* __restrict__ ref_node,
             const BdNode* __restrict__ nodes, const uint32_t* __restrict__ flags,
             const uint32_t* __restrict__ scan, const uint32_t* __restrict__ child_start,
             uint32_t* __restrict__ next_face, uint32_t* __restrict__ next_node) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const BdNode nd = nodes[ref_node[i]];
  if (nd.split < 0) return;
  uint32_t f = ref_face[i];
#pragma unroll
  for (int c = 0; c < 8; c++) {
    size_t row = (size_t)c * n;
    if (flags[row + i]) {
      uint32_t child = 8u * (uint32_t)nd.split + c;
      uint32_t pos = child_start[child] + (scan[row + i] - scan[row + nd.start]);
      next_face[pos] = f;
      next_node[pos] = child;
    }
  }
}

}  // namespace fs2

#define FS2_PACK_FN static inline __host__ __device__
#include "fs2_pack.h"

namespace fs2 {

// Leaf references of one level -> output records (mesh.cpp:89-110 reads them in this order).
__global__ void __launch_bounds__(128)
k_bd_pack(uint32_t n, const uint32_t* __restrict__ ref_face, const uint32_t* __restrict__ ref_node,
          const BdNode* __restrict__ nodes, const float* __restrict__ verts,
          const uint32_t* __restrict__ fidx, const float* __restrict__ planes, uint32_t face_base,
          float* __restrict__ out_plane, fs_tri_aux* __restrict__ out_aux) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const BdNode nd = nodes[ref_node[i]];
  if (nd.split >= 0) return;
  uint32_t f = ref_face[i];
  float p[3][3], pl[4], op[4];
  for (int k = 0; k < 3; k++) {
    uint32_t v = fidx[3 * (size_t)f + k];
    p[k][0] = verts[3 * (size_t)v];
    p[k][1] = verts[3 * (size_t)v + 1];
    p[k][2] = verts[3 * (size_t)v + 2];
  }
  for (int k = 0; k < 4; k++) pl[k] = planes[4 * (size_t)f + k];
  fs_tri_aux aux;
  fs_pack_triangle(p[0], p[1], p[2], pl, face_base + f, op, &aux);
  size_t o = (size_t)nd.out + (i - nd.start);
  reinterpret_cast<float4*>(out_plane)[o] = make_float4(op[0], op[1], op[2], op[3]);
  out_aux[o] = aux;
}

}  // namespace fs2
