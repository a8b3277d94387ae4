// fs2_oracle.cpp — CPU restatement of Final Stage 2.0's per-sample path-tracing
// hot path, operating on the flattened scene (include/fs2_cabi.h).
//
// TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline leg may load this library; the product (libfs2cuda.so and the
// host mirror) never links, imports or executes anything under oracle/.
//
// Parity status: PINNED.  The reference has no tests or golden vectors
// (SURVEY.md §4), so this restatement is pinned against the reference itself
// compiled unmodified here (oracle/ref → oracle/_ref/fs2_ref): bit-exact
// first-hit records (param, point, normal, uv, material) on every scene of
// scenes/gen_scenes.py, and bit-exact float film after whole TraceScene frames
// when the reference's xorshift64* streams are replayed (tests/test_oracle_vs_ref.py;
// vectors committed under tests/golden/ with tests/golden/make_golden.py).
//
// Deliberately written as plain scalar recursion (the shape of the reference),
// unlike the CUDA path which is an iterative state machine — the two share no
// code.  Every function cites the reference lines it follows.  Build with
//   g++ -std=c++17 -O2 -ffp-contract=off -shared -fPIC   (no -ffast-math, no -march)
// so each float operation rounds once, as in the oracle/_ref build.

#include <math.h>
#include <stdint.h>
#include <string.h>

#include "fs2_cabi.h"

namespace {

// ---------------------------------------------------------------- constants
// reference source/math/base.h:80-84
const float kPi = 3.14159262f;
const float kInfinity = 1.0e15f;
const float kEpsilon = 1.0e-5f;
// reference source/engine.cpp:16-17
const uint32_t kMaximumTraceDepth = 8;
const float kTraceStepObjectOffset = 0.03f;
// reference source/material.cpp:12-13
const float kDiffuseContribThreshold = 0.001f;
const float kDiffuseRoughnessThreshold = 0.95f;

// ---------------------------------------------------------------- scalar math
// reference source/math/scalar.h:233-243 — fast inverse square root, one Newton step.
inline float inv_sqrt(float f) {
  float half = 0.5f * f;
  int32_t i;
  memcpy(&i, &f, 4);
  i = 0x5f3759df - (i >> 1);
  memcpy(&f, &i, 4);
  f = f * (1.5f - half * f * f);
  return f;
}
// reference source/math/scalar.h:245 — hides ::sqrtf inside namespace base.
inline float approx_sqrt(float f) { return 1.0f / inv_sqrt(f); }

// reference source/math/scalar.h:333-343
inline bool compare_epsilon(float a, float b, float ep) { return a > (b - ep) && a < (b + ep); }
inline bool compare_epsilon(float a, float b) { return compare_epsilon(a, b, kEpsilon); }
// reference source/math/scalar.h:164-170
inline float saturate(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }

// ---------------------------------------------------------------- vectors
struct V3 {
  float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
// reference source/math/vector3.h:291-293
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// reference source/math/vector3.h:301
inline float length(V3 a) { return approx_sqrt(dot(a, a)); }
// reference source/math/vector3.h:295-299
inline float distance(V3 from, V3 to) { return length(to - from); }
// reference source/math/vector3.h:166-177
inline V3 normalize(V3 a) {
  float l = length(a);
  if (l == 0.0) l = 1.0f;
  return v3(a.x / l, a.y / l, a.z / l);
}
// reference source/math/vector3.h:179-187
inline V3 cross(V3 a, V3 b) {
  return v3((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x));
}
// reference source/math/vector3.h:200-203 (freflectivity = 1)
inline V3 reflect(V3 v, V3 n) { return v - (n * (dot(n, v) * (1.0f + 1.0f))); }
// reference source/math/vector3.h:205-214 — zero vector on total internal reflection
inline V3 refract(V3 v, V3 n, float index) {
  float n_dot_v = -dot(v, n);
  float sin2 = (index * index) * (1.0 - n_dot_v * n_dot_v);
  if (sin2 >= 1.0) return v3(0, 0, 0);
  V3 r = v * index + n * (index * n_dot_v - approx_sqrt(1.0 - sin2));
  return normalize(r);
}
// reference source/math/vector3.h:315-333 — axis-angle rotation, axis not normalised
inline V3 rotate(V3 p, float angle, V3 a) {
  float c = (float)cos(angle);
  float s = (float)sin(angle);
  V3 o;
  o.x = (c + (1 - c) * a.x * a.x) * p.x;
  o.x += ((1 - c) * a.x * a.y - a.z * s) * p.y;
  o.x += ((1 - c) * a.x * a.z + a.y * s) * p.z;
  o.y = ((1 - c) * a.x * a.y + a.z * s) * p.x;
  o.y += (c + (1 - c) * a.y * a.y) * p.y;
  o.y += ((1 - c) * a.y * a.z - a.x * s) * p.z;
  o.z = ((1 - c) * a.x * a.z - a.y * s) * p.x;
  o.z += ((1 - c) * a.y * a.z + a.x * s) * p.y;
  o.z += (c + (1 - c) * a.z * a.z) * p.z;
  return o;
}

struct Plane {
  float x, y, z, w;
};
// reference source/math/plane.h:73-83
inline Plane calculate_plane(V3 n, V3 p) { return Plane{n.x, n.y, n.z, -dot(n, p)}; }
// reference source/math/plane.h:68-71
inline float plane_distance(const Plane& pl, V3 p) {
  return (float)pl.x * p.x + pl.y * p.y + pl.z * p.z + pl.w;
}

// reference source/math/trace.h:40-72, trace.cpp:6-21 — rays are segments
struct Ray {
  V3 start, stop, dir;
};
inline Ray make_ray(V3 start, V3 stop) { return Ray{start, stop, stop - start}; }
inline bool has_zero_length(const Ray& r) { return compare_epsilon(length(r.dir), 0.0f); }

struct Collision {
  float param = 2.0f;
  V3 point{0, 0, 0};
  V3 normal{0, 0, 0};
};

struct Bounds {
  V3 mn, mx;
};

// ---------------------------------------------------------------- intersectors
// reference source/math/intersect.cpp:16-24
inline bool point_in_bounds(const Bounds& b, V3 p) {
  return p.x >= b.mn.x && p.x <= b.mx.x && p.y >= b.mn.y && p.y <= b.mx.y && p.z >= b.mn.z &&
         p.z <= b.mx.z;
}

// reference source/math/intersect.cpp:728-744 (the live code of ray_intersect_plane)
inline bool ray_intersect_plane(const Plane& p, const Ray& r, Collision* hit) {
  float ts = (p.x * r.dir.x + p.y * r.dir.y + p.z * r.dir.z);
  if (!compare_epsilon(ts, 0.0)) {
    float ns = (p.x * r.start.x + p.y * r.start.y + p.z * r.start.z + p.w) * -1.0;
    float t = ns / ts;
    if (t >= 0 && t <= 1) {
      hit->param = t;
      hit->point = r.start + r.dir * t;
      hit->normal = v3(p.x, p.y, p.z);
      return true;
    }
  }
  return false;
}

// reference source/math/intersect.cpp:421-488
inline bool ray_intersect_sphere(V3 center, float radius, const Ray& r, Collision* hit) {
  if (has_zero_length(r)) {
    return length(r.start - center) <= radius;  // hit left untouched (param stays 2.0)
  }
  V3 o = r.start - center;
  float a = dot(r.dir, r.dir);
  float b = 2 * (dot(o, r.dir));
  float c = dot(o, o) - radius * radius;
  float d = b * b - 4 * a * c;
  float t = 0;
  if (d < 0) return false;
  if (length(o) <= radius) {
    t = (-b + approx_sqrt(d)) / (2 * a);
  } else {
    t = (-b - approx_sqrt(d)) / (2 * a);
  }
  if (t < 0 || t > 1) return false;
  V3 p = r.start + r.dir * t;
  hit->param = t;
  hit->point = p;
  hit->normal = normalize(p - center);
  return true;
}

// reference source/math/intersect.cpp:490-529
inline bool ray_intersect_slab(float start, float vec, float smin, float smax, float* p_near,
                               float* p_far) {
  if (compare_epsilon(vec, 0.0)) {
    if (start < smin || start > smax) return false;
    return true;
  }
  float T1 = (smin - start) / vec;
  float T2 = (smax - start) / vec;
  if (T1 > T2) {
    float tmp = T1;
    T1 = T2;
    T2 = tmp;
  }
  if (T1 > (*p_near)) (*p_near) = T1;
  if (T2 < (*p_far)) (*p_far) = T2;
  return true;
}

// reference source/math/intersect.cpp:531-597 (hit normal omitted: nothing on the path reads it)
inline bool ray_intersect_bounds_worker(const Bounds& b, V3 start, V3 stop, Collision* hit) {
  V3 vec = stop - start;
  float fnear = -kInfinity;
  float ffar = kInfinity;
  if (!ray_intersect_slab(start.x, vec.x, b.mn.x, b.mx.x, &fnear, &ffar)) return false;
  if (!ray_intersect_slab(start.y, vec.y, b.mn.y, b.mx.y, &fnear, &ffar)) return false;
  if (!ray_intersect_slab(start.z, vec.z, b.mn.z, b.mx.z, &fnear, &ffar)) return false;
  if (fnear > ffar) return false;
  if (ffar < 0 || fnear < 0 || fnear > 1) return false;
  hit->param = fnear;
  hit->point = start + vec * fnear;
  return true;
}

// reference source/math/intersect.cpp:599-654
inline bool ray_intersect_bounds(const Bounds& b, const Ray& r, Collision* hit) {
  if (has_zero_length(r)) {
    if (point_in_bounds(b, r.start)) {
      hit->param = 0.0f;
      hit->point = r.start;
      return true;
    }
    return false;
  }
  if (point_in_bounds(b, r.start) && point_in_bounds(b, r.stop)) {
    hit->param = 0.0f;
    hit->point = r.start;
    return true;
  }
  if (point_in_bounds(b, r.start)) {
    if (ray_intersect_bounds_worker(b, r.stop, r.start, hit)) {
      hit->param = 1.0f - hit->param;
      return true;
    }
    return false;
  }
  return ray_intersect_bounds_worker(b, r.start, r.stop, hit);
}

// Barycentric stage shared by ray_intersect_triangle and point_in_triangle
// (reference source/math/intersect.cpp:392-402,96-129,146-177) on a packed
// record: the per-triangle constants were produced by include/fs2_pack.h with
// the same operations.
inline bool barycentric_inside(const fs_tri_aux& a, V3 point, float* u_out, float* v_out) {
  V3 vp = point - v3(a.p1[0], a.p1[1], a.p1[2]);
  uint32_t drop = a.face >> 30;
  float vpa = (drop == 0) ? vp.y : vp.x;
  float vpb = (drop == 2) ? vp.y : vp.z;
  // v0.dot(vp) with the dropped component zero in both operands
  float dot02 = a.v0a * vpa + a.v0b * vpb;
  float dot12 = a.v1a * vpa + a.v1b * vpb;
  float u = (a.dot11 * dot02 - a.dot01 * dot12) * a.inv_denom;
  float v = (a.dot00 * dot12 - a.dot01 * dot02) * a.inv_denom;
  if ((u >= 0) && (v >= 0) && (u + v <= 1)) {
    *u_out = u;
    *v_out = v;
    return true;
  }
  return false;
}

// reference source/math/intersect.cpp:357-419
inline bool ray_intersect_triangle(const float* plane4, const fs_tri_aux& aux, const Ray& r,
                                   bool zero_length, Collision* hit, float* u, float* v) {
  Plane p{plane4[0], plane4[1], plane4[2], plane4[3]};
  if (zero_length) {
    // point_in_triangle → point_in_plane (intersect.cpp:45-47,318-345): the plane is a
    // vector4 and vector4::dot(vector3→vector4(w=1)) adds w once more.
    float d = (p.x * r.start.x + p.y * r.start.y + p.z * r.start.z + p.w * 1.0f) + p.w;
    if (!compare_epsilon(d, 0.0)) return false;
    if (barycentric_inside(aux, r.start, u, v)) {
      hit->param = 0.0;
      hit->normal = v3(p.x, p.y, p.z);
      hit->point = r.start;
      return true;
    }
    return false;
  }
  Collision possible;
  if (!ray_intersect_plane(p, r, &possible)) return false;
  if (possible.param < 0 || possible.param > 1) return false;
  if (barycentric_inside(aux, possible.point, u, v)) {
    *hit = possible;
    return true;
  }
  return false;
}

// reference source/math/intersect.cpp:769-778 — signed compares
inline void planar_map_texcoords(V3 point, V3 n, float uv[2]) {
  if (n.x > n.y && n.x > n.z) {
    uv[0] = point.y;
    uv[1] = point.z;
  } else if (n.y > n.x && n.y > n.z) {
    uv[0] = point.x;
    uv[1] = point.z;
  } else {
    uv[0] = point.x;
    uv[1] = point.y;
  }
}

// reference source/math/intersect.cpp:780-784 — atan2 resolves to the float overload
inline void sphere_map_texcoords(V3 n, float uv[2]) {
  float u = atan2(n.x, n.z) / (2 * kPi) + 0.5;
  float v = n.y * 0.5 + 0.5;
  uv[0] = u;
  uv[1] = 1.0 - v;
}

}  // namespace

#include "fs2_oracle_trace.inc"
