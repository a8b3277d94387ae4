// fs2_selftest.cuh — known-answer probe of the device functions (diagnostic entry point
// fs_device_selftest).  One thread evaluates the inputs of SURVEY.md Appendix C — values printed
// by the unmodified reference — with the functions the traversal and shading kernels are built
// from; tests/test_gpu_kat.py compares the words with the reference's bits.
#pragma once
#include "fs2_kernels.cuh"
#include "fs2_build.cuh"  // fs_pack_triangle as a __host__ __device__ function

namespace fs2 {

#define FS2_SELFTEST_WORDS 48

__device__ __forceinline__ void st_put3(uint32_t* o, V3 v) {
  o[0] = __float_as_uint(v.x);
  o[1] = __float_as_uint(v.y);
  o[2] = __float_as_uint(v.z);
}

// Packs the probe triangle the way the scene upload does (include/fs2_pack.h); a separate launch
// because tri_inside reads its record through the read-only cache.
__global__ void k_selftest_pack(float4* plane, fs_tri_aux* aux) {
  const float p0[3] = {-10.f, 0.f, 0.f}, p1[3] = {30.f, -2.f, 10.f}, p2[3] = {5.f, 40.f, 20.f};
  V3 a = v3(p0[0], p0[1], p0[2]), b = v3(p1[0], p1[1], p1[2]), c = v3(p2[0], p2[1], p2[2]);
  V3 n = normalize(cross(b - a, c - a));  // plane(p1, p2, p3), reference source/math/plane.h:68-83
  const float pl[4] = {n.x, n.y, n.z, -dot(n, a)};
  float op[4];
  fs_pack_triangle(p0, p1, p2, pl, 0u, op, aux);
  *plane = make_float4(op[0], op[1], op[2], op[3]);
}

__global__ void k_selftest(const float4* __restrict__ tri_plane, const fs_tri_aux* __restrict__ tri_aux,
                           uint32_t* __restrict__ o) {
  o[0] = __float_as_uint(inv_sqrt(2.0f));
  o[1] = __float_as_uint(approx_sqrt(2.0f));
  o[2] = __float_as_uint(length(v3(3.f, 4.f, 0.f)));
  const V3 eye = v3(-5.8f, 7.05f, -47.06f);
  {  // ray A
    Ray r = make_ray(eye, v3(120.5f, -33.25f, 9000.f));
    float t = plane_param(make_float4(0.f, 1.f, 0.f, 0.f), r.start, r.dir);
    o[3] = __float_as_uint(t);
    st_put3(o + 4, r.start + r.dir * t);
    float4 pl = __ldg(tri_plane);
    float tt = plane_param(pl, r.start, r.dir);
    float u = -1.f, v = -1.f;
    bool in = tt >= 0.f && tri_inside(reinterpret_cast<const float4*>(tri_aux), r.start + r.dir * tt, u, v);
    o[7] = __float_as_uint(in ? tt : -1.f);
    o[8] = __float_as_uint(u);
    o[9] = __float_as_uint(v);
    st_put3(o + 10, v3(pl.x, pl.y, pl.z));
  }
  {  // ray B
    V3 stop = eye + (v3(4.5f, 6.f, 1.f) - eye) * 200.f;
    Ray r = make_ray(eye, stop);
    float t = -1.f;
    bool hit = sphere_test(v3(4.f, 5.f, 0.f), 5.f, r, t);
    o[13] = __float_as_uint(hit ? t : -1.f);
    st_put3(o + 14, r.start + r.dir * t);
    st_put3(o + 17, stop);
    float bp = -1.f;
    V3 entry = r.start;
    bool inside = true;
    bool bh = ray_box(v3(-1.f, 0.f, -1.f), v3(9.f, 10.f, 9.f), r, bp, entry, inside);
    o[20] = __float_as_uint(bh && !inside ? bp : -1.f);
    st_put3(o + 21, entry);
  }
  {
    const V3 up = v3(0.f, 1.f, 0.f);
    V3 v = normalize(v3(0.3f, -0.8f, 0.52f));
    st_put3(o + 24, v);
    st_put3(o + 27, reflect(v, up));
    st_put3(o + 30, refract(v, up, 0.75f));
    st_put3(o + 33, rotate_axis(v, 0.7f, up));
    float su, sv;
    sphere_uv(v, su, sv);
    o[36] = __float_as_uint(su);
    o[37] = __float_as_uint(sv);
    st_put3(o + 38, refract(normalize(v3(0.9f, -0.1f, 0.42f)), up, 1.5f));
  }
  {
    Rng g;
    g.s = 521288629ull;  // the reference's default seed, source/math/random.cpp
    for (int i = 0; i < 3; i++) {
      uint64_t x = g.next();
      o[41 + 2 * i] = (uint32_t)x;
      o[42 + 2 * i] = (uint32_t)(x >> 32);
    }
  }
  {  // DisplayFrame::WritePixel twice on one pixel (frame.cpp:53-76): mean, gamma, quantise
    float s[3] = {0.2f + 0.4f, 0.5f + 0.1f, 1.7f + 0.0f};
    uint32_t w = 0;
    for (int k = 0; k < 3; k++) {
      float m = s[k] / 2.0f;
      w |= (uint32_t)(uint8_t)(255.0 * pow((double)saturate01(m), 1.0 / 2.2) + 0.5) << (8 * k);
    }
    o[47] = w;
  }
}

}  // namespace fs2
