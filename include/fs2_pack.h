/*
 * fs2_pack.h — host-side packing of one triangle into the (tri_plane, fs_tri_aux)
 * record pair, header-only, plain C.
 *
 * The per-triangle constants are produced with exactly the float operations
 * the reference performs per ray inside ray_intersect_triangle
 * (source/math/intersect.cpp:392-402 → triangle_planar_map_vectors :96-129 and
 * triangle_find_barycentric_coeff :146-177), so hoisting them out of the ray
 * loop does not change a single bit of u, v.  Compile without FP contraction
 * (-ffp-contract=off; x86-64 baseline has no FMA).
 */
#ifndef FS2_PACK_H_
#define FS2_PACK_H_

#include <math.h>
#include "fs2_cabi.h"

/* Function qualifier: a CUDA translation unit defines it as
 * `static inline __host__ __device__` so the GPU octree builder packs records with this
 * very source (csrc/fs2_build.cuh). */
#ifndef FS2_PACK_FN
#define FS2_PACK_FN static inline
#endif

/* Axis dropped by triangle_planar_map_vectors (intersect.cpp:118-128):
 * 0 = x (map to yz), 1 = y (map to xz), 2 = z (map to xy). */
FS2_PACK_FN int fs_pack_drop_axis(const float plane[4]) {
  float ax = fabsf(plane[0]), ay = fabsf(plane[1]), az = fabsf(plane[2]);
  if (ax >= ay) return (ax >= az) ? 0 : 2;
  return (ay >= az) ? 1 : 2;
}

FS2_PACK_FN void fs_pack_triangle(const float p1[3], const float p2[3], const float p3[3],
                                    const float plane[4], uint32_t face_index,
                                    float out_plane[4], fs_tri_aux* aux) {
  float v0[3], v1[3];
  int k, drop = fs_pack_drop_axis(plane), a, b;
  for (k = 0; k < 3; k++) {
    v0[k] = p2[k] - p1[k];
    v1[k] = p3[k] - p1[k];
    out_plane[k] = plane[k];
  }
  out_plane[3] = plane[3];
  v0[drop] = 0.0f;
  v1[drop] = 0.0f;
  a = (drop == 0) ? 1 : 0;
  b = (drop == 2) ? 1 : 2;
  aux->p1[0] = p1[0];
  aux->p1[1] = p1[1];
  aux->p1[2] = p1[2];
  aux->face = (face_index & 0x3FFFFFFFu) | ((uint32_t)drop << 30);
  aux->v0a = v0[a];
  aux->v0b = v0[b];
  aux->v1a = v1[a];
  aux->v1b = v1[b];
  /* vector3::dot order: x*x + y*y + z*z (vector3.h:291-293) */
  aux->dot00 = v0[0] * v0[0] + v0[1] * v0[1] + v0[2] * v0[2];
  aux->dot01 = v0[0] * v1[0] + v0[1] * v1[1] + v0[2] * v1[2];
  aux->dot11 = v1[0] * v1[0] + v1[1] * v1[1] + v1[2] * v1[2];
  aux->inv_denom = 1 / (aux->dot00 * aux->dot11 - aux->dot01 * aux->dot01);
}

#endif /* FS2_PACK_H_ */
