// fs2_math.cuh — device-side scalar / vector arithmetic of the path-tracing hot path.
//
// Every helper reproduces the rounding sequence of the reference's math layer so that
// first-hit records are bit-identical to the CPU renderer:
//   * the translation unit is compiled with -fmad=false (no FMA contraction) and the
//     default IEEE division, so a*b+c rounds twice exactly like the x86-64 baseline;
//   * sqrt is the reference's one-Newton-step fast inverse square root
//     (reference source/math/scalar.h:233-245), NOT sqrtf;
//   * expressions the reference evaluates in double are evaluated in double here.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace fs2 {

// reference source/math/base.h:80-84
#define FS2_PI 3.14159262f
#define FS2_INFINITY 1.0e15f
#define FS2_EPSILON 1.0e-5f

struct V3 {
  float x, y, z;
};

__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 neg(V3 a) { return v3(-a.x, -a.y, -a.z); }

// reference source/math/vector3.h:291-293 — (x*x + y*y) + z*z, three roundings after the products
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// reference source/math/scalar.h:233-243
__device__ __forceinline__ float inv_sqrt(float f) {
  float half = 0.5f * f;
  int32_t i = __float_as_int(f);
  i = 0x5f3759df - (i >> 1);
  float g = __int_as_float(i);
  return g * (1.5f - half * g * g);
}
// reference source/math/scalar.h:245
__device__ __forceinline__ float approx_sqrt(float f) { return 1.0f / inv_sqrt(f); }

// reference source/math/vector3.h:301
__device__ __forceinline__ float length(V3 a) { return approx_sqrt(dot(a, a)); }

// reference source/math/vector3.h:166-177
__device__ __forceinline__ V3 normalize(V3 a) {
  float l = length(a);
  if (l == 0.0f) l = 1.0f;
  return v3(a.x / l, a.y / l, a.z / l);
}

// reference source/math/vector3.h:179-187
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return v3((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x));
}

// reference source/math/scalar.h:333-343 with ep = BASE_EPSILON, b = 0
__device__ __forceinline__ bool near_zero(float a) { return a > -FS2_EPSILON && a < FS2_EPSILON; }
__device__ __forceinline__ bool near_value(float a, float b) {
  return a > (b - FS2_EPSILON) && a < (b + FS2_EPSILON);
}

// reference source/math/vector3.h:200-203
__device__ __forceinline__ V3 reflect(V3 v, V3 n) { return v - (n * (dot(n, v) * 2.0f)); }

// reference source/math/vector3.h:205-214: `1.0 - x` terms are double there
__device__ __forceinline__ V3 refract(V3 v, V3 n, float index) {
  float n_dot_v = -dot(v, n);
  float sin2 = (float)((double)(index * index) * (1.0 - (double)(n_dot_v * n_dot_v)));
  if (sin2 >= 1.0f) return v3(0.f, 0.f, 0.f);
  float root = approx_sqrt((float)(1.0 - (double)sin2));
  V3 r = v * index + n * (index * n_dot_v - root);
  return normalize(r);
}

// sin / cos of a float angle, correctly rounded for all practical purposes: evaluated in double
// (<= 2 ulp of a double) and rounded once.  The reference calls the host libm's float versions
// (engine.cpp:235-236, vector3.h:317-318), which are correctly rounded but for rare cases, so
// path-by-path comparison against the CPU oracle sees the same bits; CUDA's sincosf (2 ulp) does
// not.
__device__ __forceinline__ void sincos_rn(float angle, float& s, float& c) {
  double ds, dc;
  sincos((double)angle, &ds, &dc);
  s = (float)ds;
  c = (float)dc;
}

// reference source/math/vector3.h:315-333 (axis is used as given, not normalised)
__device__ __forceinline__ V3 rotate_axis(V3 p, float angle, V3 a) {
  float s, c;
  sincos_rn(angle, s, c);
  float k = 1.0f - c;
  V3 o;
  o.x = (c + k * a.x * a.x) * p.x;
  o.x += (k * a.x * a.y - a.z * s) * p.y;
  o.x += (k * a.x * a.z + a.y * s) * p.z;
  o.y = (k * a.x * a.y + a.z * s) * p.x;
  o.y += (c + k * a.y * a.y) * p.y;
  o.y += (k * a.y * a.z - a.x * s) * p.z;
  o.z = (k * a.x * a.z - a.y * s) * p.x;
  o.z += (k * a.y * a.z + a.x * s) * p.y;
  o.z += (c + k * a.z * a.z) * p.z;
  return o;
}

__device__ __forceinline__ float saturate01(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }

// A ray is a segment (reference source/math/trace.cpp:6-10): dir = stop - start, un-normalised.
struct Ray {
  V3 start, stop, dir;
  bool zero_length;  // reference trace.h has_zero_length(): |dir| within 1e-5 of 0
};

__device__ __forceinline__ Ray make_ray(V3 start, V3 stop) {
  Ray r;
  r.start = start;
  r.stop = stop;
  r.dir = stop - start;
  r.zero_length = near_zero(length(r.dir));
  return r;
}

// Read-only loads with an L1 policy.  FS2_CACHE_HINTS bit 0: triangle records (read once per
// ray, 64 + 48 + 4 bytes each) do not allocate in L1; bit 1: tree nodes are evicted last.
#ifndef FS2_CACHE_HINTS
#define FS2_CACHE_HINTS 0
#endif
__device__ __forceinline__ float4 ldg_once(const float4* p) {
#if FS2_CACHE_HINTS & 1
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
#else
  return __ldg(p);
#endif
}
__device__ __forceinline__ uint32_t ldg_once(const uint32_t* p) {
#if FS2_CACHE_HINTS & 1
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
#else
  return __ldg(p);
#endif
}
__device__ __forceinline__ float4 ldg_keep(const float4* p) {
#if FS2_CACHE_HINTS & 2
  float4 v;
  asm volatile("ld.global.nc.L1::evict_last.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
#else
  return __ldg(p);
#endif
}

// reference source/math/intersect.cpp:16-24 (inclusive on both sides)
__device__ __forceinline__ bool point_in_box(V3 mn, V3 mx, V3 p) {
  return p.x >= mn.x && p.x <= mx.x && p.y >= mn.y && p.y <= mx.y && p.z >= mn.z && p.z <= mx.z;
}

}  // namespace fs2
