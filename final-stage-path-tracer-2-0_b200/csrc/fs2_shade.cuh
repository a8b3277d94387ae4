// fs2_shade.cuh — random numbers, camera rays, material sampling / weighting and the sky,
// i.e. the device side of reference source/engine.cpp:59-159,205-244,
// source/material.cpp:28-339, source/math/normal.cpp:29-105, source/scene.cpp:157-162.
//
// The reference evaluates radiance on recursion unwind; every Material::Sample is affine
// in the incoming radiance, so the path is evaluated forwards here:
//   L += T_acc * A_d ;  T_acc *= T_d      
// Random streams differ from the reference by design (counter-seeded per pixel+sample),
// so shading parity is statistical; the arithmetic still follows the reference term by term.
#pragma once
#include "fs2_traverse.cuh"

namespace fs2 {

// ------------------------------------------------------------------ RNG
// xorshift64* (the reference's generator family, source/math/random.cpp:14-23) with one
// stream per (seed, pixel, sample): the start state is a splitmix64 hash of the three,
// so a sample's stream does not depend on which GPU or launch renders it.
struct Rng {
  uint64_t s;
  __device__ __forceinline__ uint64_t next() {
    uint64_t x = s;
    x ^= x << 12;
    x ^= x >> 25;
    x ^= x << 27;
    s = x;
    return x * 0x2545F4914F6CDD1Dull;
  }
  // [0,1) with 24 bits (reference: (x mod (2^31-1)) / (2^31-1), random.cpp:25-27)
  __device__ __forceinline__ float uniform() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
  __device__ __forceinline__ uint32_t below(uint32_t n) {
    return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32);
  }
};

__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ Rng seed_rng(uint64_t seed, uint32_t pixel, uint32_t sample) {
  Rng r;
  r.s = splitmix64(splitmix64(seed ^ ((uint64_t)sample << 32 | pixel)) + 0x632BE59BD9B4E019ull);
  if (r.s == 0) r.s = 521288629ull;  // xorshift must not start at 0 (reference default seed)
  return r;
}

// Stream of the depth-0 fog decision (material.cpp:330-337).  The reference's recursion draws it
// after the whole sub-path has returned; a forward path needs it at the next hit, so it comes
// from a stream of its own (same hash, salted seed) and the order of the path's other draws
// stays the reference's.
#define FS2_FOG_STREAM_SALT 0xF06F06F06F06F06Full

// ------------------------------------------------------------------ camera
// Per-frame camera basis, computed on the host exactly as reference
// source/engine.cpp:170-197 / :300-313 does (tanf/atan/tan of the host libm).
struct CamBasis {
  float origin[3], right[3], up[3], forward[3], proj_origin[3];
  float focal_plane[4];
  float half_w, half_h;
  float wm1, hm1;  // (float)width - 1, (float)height - 1
  float aperture, z_far;
  int fast_render;
};

__device__ __forceinline__ V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// reference source/engine.cpp:213-219; jx = jy = 0 gives TraceRange's ray (:315-321).
__device__ __forceinline__ Ray camera_ray(const CamBasis& b, float i, float j, float jx, float jy) {
  float xd = (float)((double)b.half_w * ((double)((i + jx) / b.wm1) * 2.0 - 1.0));
  float yd = (float)((double)b.half_h * ((double)((j + jy) / b.hm1) * 2.0 - 1.0));
  V3 stop = ld3(b.proj_origin) + ld3(b.right) * xd + ld3(b.up) * yd;
  return make_ray(ld3(b.origin), stop);
}

// reference source/engine.cpp:221-244 — thin-lens depth of field
__device__ __forceinline__ void apply_dof(const CamBasis& b, Rng& rng, Ray& r) {
  float4 fp = make_float4(b.focal_plane[0], b.focal_plane[1], b.focal_plane[2], b.focal_plane[3]);
  float t = plane_param(fp, r.start, r.dir);
  if (t < 0.0f) return;
  V3 focal = r.start + r.dir * t;
  float angle = (float)((double)rng.uniform() * 2.0 * (double)FS2_PI);
  float mag = approx_sqrt(rng.uniform()) * b.aperture;
  float sn, cs;
  sincos_rn(angle, sn, cs);
  V3 off = ld3(b.right) * cs * mag + ld3(b.up) * sn * mag;
  V3 start = r.start + off;
  V3 stop = start + normalize(focal - start) * b.z_far;
  r = make_ray(start, stop);
}

// ------------------------------------------------------------------ materials
struct Mat {
  uint32_t type, id;
  V3 diffuse, emissive;
  float roughness, index, reflectivity, frost, density, tex_scale;
  int32_t texture;
};

__device__ __forceinline__ Mat load_material(const DScene& S, uint32_t i) {
  const float4* p = S.materials + 4 * (size_t)i;
  float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3);
  Mat m;
  m.type = __float_as_uint(a.x);
  m.id = __float_as_uint(a.y);
  m.diffuse = v3(a.z, a.w, b.x);
  m.emissive = v3(b.y, b.z, b.w);
  m.roughness = c.x;
  m.index = c.y;
  m.reflectivity = c.z;
  m.frost = c.w;
  m.density = d.x;
  m.tex_scale = d.y;
  m.texture = (int32_t)__float_as_uint(d.z);
  return m;
}

// reference source/material.cpp:107-127: nearest texel, wrap by %, float -> uint32 through
// a 64-bit truncation (x86-64 behaviour for negative inputs).
__device__ __forceinline__ V3 sample_diffuse(const DScene& S, const Mat& m, float u, float v) {
  if (m.texture < 0) return m.diffuse;
  fs_texture t = S.textures[m.texture];
  uint32_t x = (uint32_t)(long long)((u * m.tex_scale * (float)t.width + 0.5f) - 1.0f);
  uint32_t y = (uint32_t)(long long)((v * m.tex_scale * (float)t.height + 0.5f) - 1.0f);
  x %= t.width;
  y %= t.height;
  const float* px = S.texels + t.offset + (size_t)(y * t.width * 3u + x * 3u);
  return v3(__ldg(px), __ldg(px + 1), __ldg(px + 2));
}

__device__ __forceinline__ V3 lut_normal(const DScene& S, Rng& rng, V3 n) {
  float4 q = __ldg(S.lut + rng.below(S.n_lut));
  V3 d = v3(q.x, q.y, q.z);
  if (dot(d, n) < 0.0f) d = neg(d);
  return d;
}

// reference source/math/normal.cpp:36-62
__device__ __forceinline__ V3 random_reflection(const DScene& S, Rng& rng, V3 incident, V3 n,
                                                float solid_angle) {
  V3 reflect_dir = reflect(incident, n);
  V3 diffuse_dir = lut_normal(S, rng, n);
  float interp = solid_angle / FS2_PI;
  float keep = (float)(1.0 - (double)interp);
  V3 r = normalize(diffuse_dir * interp + reflect_dir * keep);
  if (dot(r, n) < 0.0f) r = neg(r);
  return r;
}

// reference source/math/normal.cpp:64-105
__device__ __forceinline__ V3 random_refraction(const DScene& S, Rng& rng, V3 incident, V3 n,
                                                float solid_angle, float index) {
  float delta = (-1.0f + rng.uniform() * (1.0f - -1.0f)) * (solid_angle / 2.0f);
  V3 refl = near_value(index, 1.0f) ? incident : refract(incident, n, index);
  refl = normalize(refl);
  V3 axis = lut_normal(S, rng, n);
  if (near_value(solid_angle, FS2_PI)) return axis;
  if (near_zero(solid_angle)) return refl;
  return rotate_axis(refl, delta, axis);
}

// float32 dot_spec = pow(half_vec.dot(n), 50) — reference material.cpp:280 promotes to double
// pow and rounds the result to float.  x^50 by squaring in double (7 products, <= 4 ulp of a
// double) rounds to the same float as a correctly rounded pow except when the exact value lies
// within 2^-50 of a float rounding boundary.
__device__ __forceinline__ float pow50(float xf) {
  double x = (double)xf;
  double x2 = x * x, x4 = x2 * x2, x8 = x4 * x4, x16 = x8 * x8, x32 = x16 * x16;
  return (float)(x32 * x16 * x2);
}

// Outcome of one surface interaction in forward form.
struct Bounce {
  V3 dir;      // Material::Reflection() result
  V3 add;      // A: radiance added regardless of the incoming light
  V3 weight;   // T: factor on the incoming radiance
  bool recurse;  // Material::WillUseIndirectLight()
};

// Material::Reflection + WillUseIndirectLight + Sample for all nine classes
// (reference source/material.cpp:28-339, table in SURVEY.md App. F).
__device__ __forceinline__ Bounce shade_surface(const DScene& S, const Mat& m, Rng& rng, V3 view,
                                                V3 n, float u, float v) {
  Bounce b;
  b.add = v3(0.f, 0.f, 0.f);
  b.weight = v3(0.f, 0.f, 0.f);
  b.dir = v3(0.f, 0.f, 0.f);
  b.recurse = true;
  switch (m.type) {
    case FS_MAT_LIGHT:
      b.recurse = false;
      b.add = (m.texture >= 0) ? sample_diffuse(S, m, u, v) : m.emissive;
      break;
    case FS_MAT_DIFFUSE: {
      b.dir = random_reflection(S, rng, view, n, FS2_PI);
      float cs = dot(b.dir, n);
      b.recurse = cs > 0.001f;
      b.weight = sample_diffuse(S, m, u, v) * fmaxf(0.0f, dot(n, b.dir));
      break;
    }
    case FS_MAT_METAL: {
      b.dir = random_reflection(S, rng, view, n, FS2_PI * m.roughness);
      b.recurse = m.roughness <= 0.95f || dot(b.dir, n) > 0.001f;
      V3 kd = sample_diffuse(S, m, u, v);
      float cs = fmaxf(0.0f, dot(n, b.dir));
      b.weight = kd * cs * m.roughness + kd * (float)(1.0 - (double)m.roughness);
      break;
    }
    case FS_MAT_MIRROR:
      b.dir = reflect(view, n);
      b.weight = m.diffuse;
      break;
    case FS_MAT_GLASS:
      if (rng.uniform() < m.reflectivity)
        b.dir = random_reflection(S, rng, view, n, FS2_PI * m.frost);
      else
        b.dir = random_refraction(S, rng, view, n, FS2_PI * m.frost, m.index);
      b.weight = m.diffuse;
      break;
    case FS_MAT_LIQUID:
      b.dir = (rng.uniform() < m.reflectivity) ? reflect(view, n) : refract(view, n, m.index);
      b.weight = m.diffuse;
      break;
    case FS_MAT_CERAMIC:
    case FS_MAT_GLOW: {
      if (rng.uniform() < 0.1f)
        b.dir = random_reflection(S, rng, view, n, 0.0f);
      else
        b.dir = random_reflection(S, rng, view, n, (float)((double)FS2_PI * (1.0 - (double)m.roughness)));
      V3 half_vec = normalize(neg(view) + b.dir);
      float spec = pow50(dot(half_vec, n));
      V3 kd = sample_diffuse(S, m, u, v) * fmaxf(0.0f, dot(n, b.dir));
      float rest = (float)(1.0 - (double)spec);
      b.weight = v3(spec, spec, spec) + kd * rest;
      if (m.type == FS_MAT_GLOW) b.add = m.emissive;
      break;
    }
    case FS_MAT_FOG:
      b.dir = view;
      b.weight = v3(1.f, 1.f, 1.f);
      break;
    default:
      b.recurse = false;
      break;
  }
  return b;
}

// reference source/scene.cpp:157-162 + engine.cpp:88-102: sky.Sample(...) * 3 with every
// vector argument zero.  A non-Light sky material therefore evaluates to black.
__device__ __forceinline__ V3 sample_sky(const DScene& S, V3 view) {
  Mat sky = load_material(S, S.sky_material);
  V3 c = v3(0.f, 0.f, 0.f);
  if (sky.type == FS_MAT_LIGHT) {
    if (sky.texture >= 0) {
      float u, v;
      sphere_uv(view, u, v);
      c = sample_diffuse(S, sky, u, v);
    } else {
      c = sky.emissive;
    }
  } else if (sky.type == FS_MAT_GLOW) {
    c = sky.emissive;
  }
  return c * 3.0f;
}

}  // namespace fs2
