// fs2_post.cuh — film post-processing on the device.
//
// The reference's README describes variance-adaptive denoising, bloom and flare, but its code has
// none of them; what it does have are the guide buffers such a filter needs:
// DisplayFrame's per-pixel normal, depth and material id (reference source/frame.h:86-91, written
// by frame.cpp:78-85).  There is therefore no reference behaviour to be at parity with; the
// definition below is this repo's, restated on the CPU in oracle/post_oracle.py and checked
// against it in tests/test_gpu_post.py (float tolerance stated there).
//
// All stages read the resolved mean image (k_resolve) and the film's guide planes and are
// HBM / L2-bound image passes, one thread per pixel:
//
//   k_post_prep     mean rgb -> float4 (rgb, luminance)
//   k_post_sigma    3x3 same-material luminance standard deviation of the mean image: the noise
//                   estimate the colour weight adapts to
//   k_post_atrous   edge-avoiding a-trous wavelet pass (5x5 B3 spline taps at stride 2^i), weights
//                   = tap x normal x depth x colour, taps of another material id rejected
//   k_post_bright   bright pass: colour x max(0, lum - threshold) / lum
//   k_post_blur     separable Gaussian (run twice: rows, then columns), clamped at the border
//   k_post_compose  image + bloom + flare ghosts (the blurred bright pass mirrored through the
//                   image centre at G scales), float rgb and display rgb8 (frame.cpp:60-69)
#pragma once
#include "fs2_math.cuh"

namespace fs2 {

struct PostParams {
  float sigma_normal, sigma_depth, sigma_colour;
  float bloom_threshold, bloom_strength;
  int bloom_radius;
  float flare_strength;
  int flare_ghosts;
};

__device__ __forceinline__ float post_lum(float r, float g, float b) {
  return 0.2126f * r + 0.7152f * g + 0.0722f * b;
}

__global__ void __launch_bounds__(256)
k_post_prep(const float* __restrict__ mean_rgb, int n, float4* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float r = mean_rgb[3 * i], g = mean_rgb[3 * i + 1], b = mean_rgb[3 * i + 2];
  out[i] = make_float4(r, g, b, post_lum(r, g, b));
}

__global__ void __launch_bounds__(256)
k_post_sigma(const float4* __restrict__ col, const uint32_t* __restrict__ ids, int w, int h,
             float* __restrict__ sigma) {
  int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= w || y >= h) return;
  int p = y * w + x;
  uint32_t id = ids[p];
  float s1 = 0.f, s2 = 0.f, cnt = 0.f;
  for (int dy = -1; dy <= 1; dy++)
    for (int dx = -1; dx <= 1; dx++) {
      int qx = x + dx, qy = y + dy;
      if (qx < 0 || qy < 0 || qx >= w || qy >= h) continue;
      int q = qy * w + qx;
      if (ids[q] != id) continue;
      float l = col[q].w;
      s1 += l;
      s2 += l * l;
      cnt += 1.0f;
    }
  float m1 = s1 / cnt, m2 = s2 / cnt;
  float var = fmaxf(m2 - m1 * m1, 0.0f);
  sigma[p] = sqrtf(var);
}

// One a-trous pass at stride `step`; sigma_scale = 2^-pass (the noise left after each pass).
__global__ void __launch_bounds__(256)
k_post_atrous(const float4* __restrict__ col, const float4* __restrict__ aux,
              const uint32_t* __restrict__ ids, const float* __restrict__ sigma, int w, int h,
              int step, float sigma_scale, PostParams pp, float4* __restrict__ out) {
  int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= w || y >= h) return;
  int p = y * w + x;
  const float4 cp = col[p], gp = aux[p];
  const uint32_t id = ids[p];
  const bool has_normal = (gp.x * gp.x + gp.y * gp.y + gp.z * gp.z) > 0.0f;
  const float inv_z = 1.0f / (pp.sigma_depth * fmaxf(fabsf(gp.w), 1e-3f));
  const float inv_l = 1.0f / (pp.sigma_colour * sigma[p] * sigma_scale + 1e-4f);
  const float tap[3] = {0.375f, 0.25f, 0.0625f};
  float ar = 0.f, ag = 0.f, ab = 0.f, aw = 0.f;
  for (int dy = -2; dy <= 2; dy++) {
    int qy = y + dy * step;
    if (qy < 0 || qy >= h) continue;
    for (int dx = -2; dx <= 2; dx++) {
      int qx = x + dx * step;
      if (qx < 0 || qx >= w) continue;
      int q = qy * w + qx;
      if (ids[q] != id) continue;
      float4 cq = __ldg(col + q), gq = __ldg(aux + q);
      float wt = tap[dx < 0 ? -dx : dx] * tap[dy < 0 ? -dy : dy];
      if (has_normal) {
        float d = fmaxf(gp.x * gq.x + gp.y * gq.y + gp.z * gq.z, 0.0f);
        wt *= powf(d, pp.sigma_normal);
      }
      wt *= expf(-fabsf(gp.w - gq.w) * inv_z);
      wt *= expf(-fabsf(cp.w - cq.w) * inv_l);
      ar += wt * cq.x;
      ag += wt * cq.y;
      ab += wt * cq.z;
      aw += wt;
    }
  }
  float r = ar / aw, g = ag / aw, b = ab / aw;
  out[p] = make_float4(r, g, b, post_lum(r, g, b));
}

__global__ void __launch_bounds__(256)
k_post_bright(const float4* __restrict__ col, int n, float threshold, float4* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 c = col[i];
  float k = fmaxf(c.w - threshold, 0.0f) / fmaxf(c.w, 1e-6f);
  out[i] = make_float4(c.x * k, c.y * k, c.z * k, 0.0f);
}

// Gaussian of radius R (sigma = R / 3) along x (dir = 0) or y (dir = 1), coordinates clamped.
__global__ void __launch_bounds__(256)
k_post_blur(const float4* __restrict__ in, int w, int h, int radius, int dir,
            float4* __restrict__ out) {
  int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= w || y >= h) return;
  float sg = (float)radius / 3.0f;
  float inv = 1.0f / (2.0f * sg * sg);
  float ar = 0.f, ag = 0.f, ab = 0.f, aw = 0.f;
  for (int k = -radius; k <= radius; k++) {
    int qx = dir ? x : min(max(x + k, 0), w - 1);
    int qy = dir ? min(max(y + k, 0), h - 1) : y;
    float wt = expf(-(float)(k * k) * inv);
    float4 c = __ldg(in + qy * w + qx);
    ar += wt * c.x;
    ag += wt * c.y;
    ab += wt * c.z;
    aw += wt;
  }
  out[y * w + x] = make_float4(ar / aw, ag / aw, ab / aw, 0.0f);
}

__global__ void __launch_bounds__(256)
k_post_compose(const float4* __restrict__ col, const float4* __restrict__ glow, int w, int h,
               PostParams pp, float* __restrict__ rgb, uint8_t* __restrict__ rgb8) {
  int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= w || y >= h) return;
  int p = y * w + x;
  float4 c = col[p];
  float r = c.x, g = c.y, b = c.z;
  if (glow) {
    if (pp.bloom_strength > 0.0f) {
      float4 e = glow[p];
      r += pp.bloom_strength * e.x;
      g += pp.bloom_strength * e.y;
      b += pp.bloom_strength * e.z;
    }
    if (pp.flare_strength > 0.0f && pp.flare_ghosts > 0) {
      float cx = 0.5f * (float)(w - 1), cy = 0.5f * (float)(h - 1);
      float k = pp.flare_strength / (float)pp.flare_ghosts;
      for (int i = 1; i <= pp.flare_ghosts; i++) {
        float t = (float)i / (float)pp.flare_ghosts;
        int qx = (int)floorf(cx + (cx - (float)x) * t + 0.5f);
        int qy = (int)floorf(cy + (cy - (float)y) * t + 0.5f);
        if (qx < 0 || qy < 0 || qx >= w || qy >= h) continue;
        float4 e = __ldg(glow + qy * w + qx);
        r += k * e.x;
        g += k * e.y;
        b += k * e.z;
      }
    }
  }
  if (rgb) {
    rgb[3 * p] = r;
    rgb[3 * p + 1] = g;
    rgb[3 * p + 2] = b;
  }
  if (rgb8) {
    rgb8[3 * p + 0] = (uint8_t)(255.0 * pow((double)saturate01(r), 1.0 / 2.2) + 0.5);
    rgb8[3 * p + 1] = (uint8_t)(255.0 * pow((double)saturate01(g), 1.0 / 2.2) + 0.5);
    rgb8[3 * p + 2] = (uint8_t)(255.0 * pow((double)saturate01(b), 1.0 / 2.2) + 0.5);
  }
}

}  // namespace fs2
