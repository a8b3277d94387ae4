// fs2_kernels.cuh — the sm_100a kernels of the hot path.
//
//   k_first_hit   TraceRange for every pixel (reference source/engine.cpp:298-329)
//   k_trace_mega  n samples of TraceScene per pixel (source/engine.cpp:59-159,166-251),
//                 persistent per-pixel thread with path regeneration
//   k_resolve     DisplayFrame::WritePixel's mean / gamma / quantise (source/frame.cpp:53-76)
#pragma once
#include "fs2_shade.cuh"

namespace fs2 {

struct Film {
  float4* sum;       // rgb sum, w = sample count
  float4* aux;       // normal xyz, depth (last sample wins, frame.cpp:78-85)
  uint32_t* mat_id;  // material id of the last sample
  int width, height;
};

struct TraceParams {
  uint32_t first_sample, n_samples, sample_stride;
  uint64_t seed;
  int flags;
};

// stats slots (device array of unsigned long long)
enum { ST_PATHS = 0, ST_SEGMENTS, ST_NODE, ST_TRI, ST_BARY, ST_PRIM, ST_SHADE, ST_COUNT };

__device__ __forceinline__ void warp_add(unsigned long long* slot, uint32_t v) {
  v = __reduce_add_sync(0xFFFFFFFFu, v);
  if ((threadIdx.x & 31) == 0 && v) atomicAdd(slot, (unsigned long long)v);
}

// Warp = 8x4 pixel tile, block = 4 warps = 16x8 pixels: neighbouring lanes trace
// neighbouring camera rays (coherent octree walks and leaf loads on the first segment).
__device__ __forceinline__ bool pixel_of_thread(int width, int height, int& px, int& py) {
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  px = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  py = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  return px < width && py < height;
}

__global__ void __launch_bounds__(128)
k_first_hit(DScene S, CamBasis cam, int width, int height, float* __restrict__ out_t,
            uint32_t* __restrict__ out_object, uint32_t* __restrict__ out_face,
            uint32_t* __restrict__ out_mat, float* __restrict__ out_range,
            unsigned long long* __restrict__ stats) {
  int px, py;
  bool valid = pixel_of_thread(width, height, px, py);
  Counters c = {0, 0, 0, 0};
  if (valid) {
    Ray r = camera_ray(cam, (float)px, (float)py, 0.0f, 0.0f);
    Hit h = scene_trace<true>(S, r, c);
    size_t p = (size_t)py * width + px;
    bool got = h.object != 0xFFFFFFFFu;
    uint32_t face = 0xFFFFFFFFu, mat = S.sky_material;
    float range = cam.z_far;
    if (got) {
      Surface sf = make_surface(S, r, h, false);
      face = sf.face;
      mat = sf.material;
      range = length(sf.point - r.start);
    }
    if (out_t) out_t[p] = got ? h.param : FS_NO_HIT_PARAM;
    if (out_object) out_object[p] = h.object;
    if (out_face) out_face[p] = face;
    if (out_mat) out_mat[p] = __float_as_uint(__ldg(S.materials + 4 * (size_t)mat).y);
    if (out_range) out_range[p] = range;
  }
  warp_add(stats + ST_PATHS, valid ? 1u : 0u);
  warp_add(stats + ST_SEGMENTS, valid ? 1u : 0u);
  warp_add(stats + ST_NODE, c.n_node);
  warp_add(stats + ST_TRI, c.n_tri);
  warp_add(stats + ST_BARY, c.n_bary);
  warp_add(stats + ST_PRIM, c.n_prim);
}

// One thread owns one pixel and renders its n_samples paths back to back; a finished
// path is replaced by the pixel's next sample inside the same loop iteration, so the
// warp only diverges inside a segment, never across path lengths.
template <bool STATS>
__global__ void __launch_bounds__(128)
k_trace_mega(DScene S, CamBasis cam, Film film, TraceParams tp,
             unsigned long long* __restrict__ stats) {
  int px, py;
  bool alive = pixel_of_thread(film.width, film.height, px, py) && tp.n_samples > 0;
  const uint32_t pixel = alive ? (uint32_t)(py * film.width + px) : 0u;
  Counters c = {0, 0, 0, 0};
  uint32_t n_seg = 0, n_shade = 0;

  V3 sum = v3(0.f, 0.f, 0.f);
  V3 L = v3(0.f, 0.f, 0.f), T = v3(1.f, 1.f, 1.f);
  uint32_t s = 0, depth = 0;
  Rng rng;
  Ray ray;
  // depth-0 fog decision waits for the next hit position (material.cpp:330-337)
  bool fog_pending = false;
  V3 fog_pos = v3(0.f, 0.f, 0.f);
  uint32_t fog_mat = 0;
  // aux of the current path (TraceResult, frame.h:40-47)
  V3 aux_n = v3(0.f, 0.f, 0.f);
  float aux_depth = 0.f;
  uint32_t aux_mat = 0;

  auto begin_path = [&]() {
    uint32_t sample = tp.first_sample + s * tp.sample_stride;
    rng = seed_rng(tp.seed, pixel, sample);
    float jx = 0.f, jy = 0.f;
    if (!(tp.flags & FS_TRACE_NO_JITTER)) {
      jx = rng.uniform() - 0.5f;
      jy = rng.uniform() - 0.5f;
    }
    ray = camera_ray(cam, (float)px, (float)py, jx, jy);
    if (cam.aperture > 0.0f && !(tp.flags & FS_TRACE_NO_DOF)) apply_dof(cam, rng, ray);
    L = v3(0.f, 0.f, 0.f);
    T = v3(1.f, 1.f, 1.f);
    depth = 0;
    fog_pending = false;
  };
  if (alive) begin_path();

  while (alive) {
    Hit h = scene_trace<STATS>(S, ray, c);
    n_seg++;
    bool got = h.object != 0xFFFFFFFFu;
    bool done = false;

    if (fog_pending) {
      // FogMaterial::Sample at depth 0 (material.cpp:330-337): light_pos is the next hit
      // position, or the segment end on a miss (engine.cpp:89-91).
      fog_pending = false;
      V3 light_pos = got ? (ray.start + ray.dir * h.param) : ray.stop;
      float dist = length(fog_pos - light_pos);
      Mat fm = load_material(S, fog_mat);
      float a = (dist * dist) * fm.density * 0.00005f;
      float threshold = saturate01(0.0f > a ? 0.0f : a);
      if (rng.uniform() < threshold) {
        L = fm.diffuse;
        done = true;
      }
    }

    if (!done && !got) {
      // engine.cpp:88-102
      V3 view = normalize(ray.stop - ray.start);
      L = L + T * sample_sky(S, view);
      n_shade++;
      if (depth == 0) {
        aux_n = normalize(ray.dir);
        aux_depth = cam.z_far;
        aux_mat = S.sky_material;
      }
      done = true;
    } else if (!done) {
      Mat m = load_material(S, __float_as_uint(obj_word(S, h.object, 0).y));
      Surface sf = make_surface(S, ray, h, m.texture >= 0);
      V3 view = normalize(sf.point - ray.start);
      Bounce b = shade_surface(S, m, rng, view, sf.normal, sf.u, sf.v);
      n_shade++;
      V3 add = b.add;
      if (depth == 0) {
        // engine.cpp:147-156
        if (m.type == FS_MAT_LIGHT && length(add) > 10.0f) add = normalize(add) * 10.0f;
        aux_n = sf.normal;
        aux_depth = length(ray.start - sf.point);
        aux_mat = sf.material;
        if (m.type == FS_MAT_FOG) {
          fog_pending = true;
          fog_pos = sf.point;
          fog_mat = sf.material;
        }
      }
      L = L + T * add;
      if (!b.recurse) {
        done = true;
      } else {
        T = T * b.weight;
        depth++;
        if (depth >= 8u) {
          done = true;  // engine.cpp:62-64: black beyond kMaximumTraceDepth
        } else if (cam.fast_render && depth > 1u) {
          L = L + T;  // engine.cpp:68-70
          done = true;
        } else {
          // engine.cpp:120-127
          V3 start = sf.point + b.dir * 0.03f;
          V3 stop = sf.point + b.dir * cam.z_far;
          Ray nr;
          nr.start = start;
          nr.stop = stop;
          nr.dir = (stop - sf.point) - (b.dir * 0.03f);
          nr.zero_length = near_zero(length(nr.dir));
          ray = nr;
        }
      }
    }

    if (done) {
      sum = sum + L;
      s++;
      if (s == tp.n_samples) {
        alive = false;
      } else {
        begin_path();
      }
    }
  }

  if (pixel_of_thread(film.width, film.height, px, py) && tp.n_samples > 0) {
    float4 acc = film.sum[pixel];
    acc.x += sum.x;
    acc.y += sum.y;
    acc.z += sum.z;
    acc.w += (float)tp.n_samples;
    film.sum[pixel] = acc;
    film.aux[pixel] = make_float4(aux_n.x, aux_n.y, aux_n.z, aux_depth);
    film.mat_id[pixel] = __float_as_uint(__ldg(S.materials + 4 * (size_t)aux_mat).y);
  }
  warp_add(stats + ST_PATHS, s);
  warp_add(stats + ST_SEGMENTS, n_seg);
  warp_add(stats + ST_SHADE, n_shade);
  if (STATS) {
    warp_add(stats + ST_NODE, c.n_node);
    warp_add(stats + ST_TRI, c.n_tri);
    warp_add(stats + ST_BARY, c.n_bary);
    warp_add(stats + ST_PRIM, c.n_prim);
  }
}

// reference source/frame.cpp:56-69: mean and display RGB8 = 255*pow(sat(mean), 1/2.2) + 0.5
__global__ void __launch_bounds__(256)
k_resolve(const float4* __restrict__ sum, int n, float* __restrict__ mean_rgb,
          uint32_t* __restrict__ count, uint8_t* __restrict__ rgb8) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 a = sum[i];
  float cnt = a.w;
  V3 m = v3(0.f, 0.f, 0.f);
  if (cnt > 0.0f) m = v3(a.x / cnt, a.y / cnt, a.z / cnt);
  mean_rgb[3 * i + 0] = m.x;
  mean_rgb[3 * i + 1] = m.y;
  mean_rgb[3 * i + 2] = m.z;
  count[i] = (uint32_t)cnt;
  rgb8[3 * i + 0] = (uint8_t)(255.0 * pow((double)saturate01(m.x), 1.0 / 2.2) + 0.5);
  rgb8[3 * i + 1] = (uint8_t)(255.0 * pow((double)saturate01(m.y), 1.0 / 2.2) + 0.5);
  rgb8[3 * i + 2] = (uint8_t)(255.0 * pow((double)saturate01(m.z), 1.0 / 2.2) + 0.5);
}

}  // namespace fs2
