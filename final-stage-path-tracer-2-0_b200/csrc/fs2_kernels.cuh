// fs2_kernels.cuh — the sm_100a kernels of the hot path.
//
//   k_first_hit   TraceRange for every pixel (reference source/engine.cpp:298-329)
//   k_trace_mega  n samples of TraceScene per pixel (source/engine.cpp:59-159,166-251),
//                 persistent per-pixel thread with path regeneration
//   k_resolve     DisplayFrame::WritePixel's mean / gamma / quantise (source/frame.cpp:53-76)
#pragma once
#include "fs2_coop.cuh"
#include "fs2_shade.cuh"
#include "fs2_phase.cuh"

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
  uint32_t tune;  // experiment knob (env FS2_TUNE), see the kernels that read it
};

// stats slots (device array of unsigned long long)
enum { ST_PATHS = 0, ST_SEGMENTS, ST_NODE, ST_TRI, ST_BARY, ST_PRIM, ST_SHADE, ST_LBNODE, ST_TRITEST, ST_COUNT };

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

// LB: mesh leaves are searched through the leaf BVH (default); false = the reference's linear
// leaf loop (kept for A/B checks).
template <bool LB>
__global__ void __launch_bounds__(FS2_LB_THREADS)
k_first_hit(DScene S, CamBasis cam, int width, int height, float* __restrict__ out_t,
            uint32_t* __restrict__ out_object, uint32_t* __restrict__ out_face,
            uint32_t* __restrict__ out_mat, float* __restrict__ out_range,
            unsigned long long* __restrict__ stats) {
  __shared__ uint2 lb_stack[LB ? FS2_LB_STACK * FS2_LB_THREADS : 1];
  int px, py;
  bool valid = pixel_of_thread(width, height, px, py);
  Counters c = {0, 0, 0, 0, 0, 0};
  if (valid) {
    Ray r = camera_ray(cam, (float)px, (float)py, 0.0f, 0.0f);
    Hit h = scene_trace<true, LB>(S, r, c, lb_stack + (LB ? threadIdx.x : 0));
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
  warp_add(stats + ST_LBNODE, c.n_lbnode);
  warp_add(stats + ST_TRITEST, c.n_tritest);
}

// Scene::Trace for a list of caller-supplied segments (fs_trace_rays): one thread per ray.
__global__ void __launch_bounds__(FS2_LB_THREADS)
k_probe_rays(DScene S, const float* __restrict__ seg, int n, fs_ray_hit* __restrict__ out) {
  __shared__ uint2 lb_stack[FS2_LB_STACK * FS2_LB_THREADS];
  const int i = blockIdx.x * FS2_LB_THREADS + threadIdx.x;
  if (i >= n) return;
  Counters c = {0, 0, 0, 0, 0, 0};
  const float* g = seg + 6 * (size_t)i;
  Ray r = make_ray(v3(g[0], g[1], g[2]), v3(g[3], g[4], g[5]));
  Hit h = scene_trace<false, true>(S, r, c, lb_stack + threadIdx.x);
  fs_ray_hit o;
  o.param = FS_NO_HIT_PARAM;
  o.point[0] = o.point[1] = o.point[2] = 0.f;
  o.normal[0] = o.normal[1] = o.normal[2] = 0.f;
  o.uv[0] = o.uv[1] = 0.f;
  o.object = 0xFFFFFFFFu;
  o.face = 0xFFFFFFFFu;
  o.material = S.sky_material;
  o.is_internal = 0;
  if (h.object != 0xFFFFFFFFu) {
    Surface sf = make_surface(S, r, h, true);
    o.param = h.param;
    o.point[0] = sf.point.x;
    o.point[1] = sf.point.y;
    o.point[2] = sf.point.z;
    o.normal[0] = sf.normal.x;
    o.normal[1] = sf.normal.y;
    o.normal[2] = sf.normal.z;
    o.uv[0] = sf.u;
    o.uv[1] = sf.v;
    o.object = h.object;
    o.face = sf.face;
    o.material = sf.material;
    o.is_internal = sf.internal ? 1u : 0u;
  }
  out[i] = o;
}

// Scene::SampleSky for a list of view directions (fs_sample_sky).
__global__ void k_probe_sky(DScene S, const float* __restrict__ views, int n, float* __restrict__ rgb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  V3 c = sample_sky(S, v3(views[3 * i], views[3 * i + 1], views[3 * i + 2]));
  rgb[3 * i] = c.x;
  rgb[3 * i + 1] = c.y;
  rgb[3 * i + 2] = c.z;
}

// ------------------------------------------------------------------ path state
struct PathState {
  V3 sum;  // radiance of the finished paths of this launch
  V3 L, T;
  uint32_t s, depth;
  Rng rng;
  Ray ray;
  // depth-0 fog decision waits for the next hit position (material.cpp:330-337)
  bool fog_pending;
  V3 fog_pos;
  uint32_t fog_mat;
};

// reference source/engine.cpp:205-244: jittered camera ray (+ thin lens) of the next sample
__device__ __forceinline__ void begin_path(const CamBasis& cam, const TraceParams& tp,
                                           uint32_t pixel, int px, int py, PathState& ps) {
  uint32_t sample = tp.first_sample + ps.s * tp.sample_stride;
  ps.rng = seed_rng(tp.seed, pixel, sample);
  float jx = 0.f, jy = 0.f;
  if (!(tp.flags & FS_TRACE_NO_JITTER)) {
    jx = ps.rng.uniform() - 0.5f;
    jy = ps.rng.uniform() - 0.5f;
  }
  ps.ray = camera_ray(cam, (float)px, (float)py, jx, jy);
  if (cam.aperture > 0.0f && !(tp.flags & FS_TRACE_NO_DOF)) apply_dof(cam, ps.rng, ps.ray);
  ps.L = v3(0.f, 0.f, 0.f);
  ps.T = v3(1.f, 1.f, 1.f);
  ps.depth = 0;
  ps.fog_pending = false;
}

// Everything TraceStep does around Scene::Trace (reference source/engine.cpp:59-159) for the
// segment whose nearest hit is `h`, in forward form.  Returns true when the path is finished
// (its radiance is ps.L); otherwise the next segment's ray is in ps.ray.  The aux buffers of
// the film are written by the last sample of the launch (frame.cpp:78-85: last sample wins).
__device__ __forceinline__ bool shade_segment(const DScene& S, const CamBasis& cam,
                                              const Film& film, const TraceParams& tp,
                                              uint32_t pixel, bool last, PathState& ps,
                                              const Hit& h, uint32_t& n_shade) {
  const bool got = h.object != 0xFFFFFFFFu;
  bool done = false;

  if (ps.fog_pending) {
    // FogMaterial::Sample at depth 0: light_pos is the next hit position, or the segment
    // end on a miss (engine.cpp:89-91).
    ps.fog_pending = false;
    V3 light_pos = got ? (ps.ray.start + ps.ray.dir * h.param) : ps.ray.stop;
    float dist = length(ps.fog_pos - light_pos);
    Mat fm = load_material(S, ps.fog_mat);
    float a = (dist * dist) * fm.density * 0.00005f;
    float threshold = saturate01(0.0f > a ? 0.0f : a);
    Rng fog_rng = seed_rng(tp.seed ^ FS2_FOG_STREAM_SALT, pixel,
                           tp.first_sample + ps.s * tp.sample_stride);
    if (fog_rng.uniform() < threshold) {
      ps.L = fm.diffuse;
      done = true;
    }
  }

  if (!done && !got) {
    // engine.cpp:88-102
    V3 view = normalize(ps.ray.stop - ps.ray.start);
    ps.L = ps.L + ps.T * sample_sky(S, view);
    n_shade++;
    if (ps.depth == 0 && last) {
      V3 n = normalize(ps.ray.dir);
      film.aux[pixel] = make_float4(n.x, n.y, n.z, cam.z_far);
      film.mat_id[pixel] = __float_as_uint(__ldg(S.materials + 4 * (size_t)S.sky_material).y);
    }
    done = true;
  } else if (!done) {
    Mat m = load_material(S, __float_as_uint(obj_word(S, h.object, 0).y));
    Surface sf = make_surface(S, ps.ray, h, m.texture >= 0);
    V3 view = normalize(sf.point - ps.ray.start);
    Bounce b = shade_surface(S, m, ps.rng, view, sf.normal, sf.u, sf.v);
    n_shade++;
    V3 add = b.add;
    if (ps.depth == 0) {
      // engine.cpp:147-156
      if (m.type == FS_MAT_LIGHT && length(add) > 10.0f) add = normalize(add) * 10.0f;
      if (last) {
        film.aux[pixel] = make_float4(sf.normal.x, sf.normal.y, sf.normal.z,
                                      length(ps.ray.start - sf.point));
        film.mat_id[pixel] = m.id;
      }
      if (m.type == FS_MAT_FOG) {
        ps.fog_pending = true;
        ps.fog_pos = sf.point;
        ps.fog_mat = sf.material;
      }
    }
    ps.L = ps.L + ps.T * add;
    if (!b.recurse) {
      done = true;
    } else {
      ps.T = ps.T * b.weight;
      ps.depth++;
      if (ps.depth >= 8u) {
        done = true;  // engine.cpp:62-64: black beyond kMaximumTraceDepth
      } else if (cam.fast_render && ps.depth > 1u) {
        ps.L = ps.L + ps.T;  // engine.cpp:68-70
        done = true;
      } else {
        // engine.cpp:120-127
        V3 off = b.dir * 0.03f;
        V3 stop = sf.point + b.dir * cam.z_far;
        ps.ray.start = sf.point + off;
        ps.ray.stop = stop;
        ps.ray.dir = (stop - sf.point) - off;
        ps.ray.zero_length = near_zero(length(ps.ray.dir));
      }
    }
  }

  return done;
}

// shade_segment, then — when the path is finished — the pixel's next sample (if any) is
// started.  Returns false when the lane has rendered all its samples.
__device__ __forceinline__ bool finish_segment(const DScene& S, const CamBasis& cam,
                                               const Film& film, const TraceParams& tp,
                                               uint32_t pixel, int px, int py, PathState& ps,
                                               const Hit& h, uint32_t& n_shade) {
  if (shade_segment(S, cam, film, tp, pixel, ps.s + 1 == tp.n_samples, ps, h, n_shade)) {
    ps.sum = ps.sum + ps.L;
    ps.s++;
    if (ps.s == tp.n_samples) return false;
    begin_path(cam, tp, pixel, px, py, ps);
  }
  return true;
}

__device__ __forceinline__ void film_add(const Film& film, uint32_t pixel, V3 sum, uint32_t n) {
  float4 acc = film.sum[pixel];
  acc.x += sum.x;
  acc.y += sum.y;
  acc.z += sum.z;
  acc.w += (float)n;
  film.sum[pixel] = acc;
}

// Variant A — one thread owns one pixel, walks its own leaves (kept for the ncu comparison
// and as the FS_TRACE_PER_THREAD path).  A finished path is replaced by the pixel's next
// sample inside the same loop iteration, so the warp only diverges inside a segment.
template <bool STATS, bool LB>
__global__ void __launch_bounds__(FS2_LB_THREADS)
k_trace_mega(DScene S, CamBasis cam, Film film, TraceParams tp,
             unsigned long long* __restrict__ stats) {
  // LB: per-lane stacks of the leaf-BVH walk, one column per thread (conflict-free)
  __shared__ uint2 lb_stack[LB ? FS2_LB_STACK * FS2_LB_THREADS : 1];
  uint2* my_stack = lb_stack + (LB ? threadIdx.x : 0);
  int px, py;
  const bool mine = pixel_of_thread(film.width, film.height, px, py) && tp.n_samples > 0;
  const uint32_t pixel = mine ? (uint32_t)(py * film.width + px) : 0u;
  Counters c = {0, 0, 0, 0, 0, 0};
  uint32_t n_seg = 0, n_shade = 0;
  PathState ps;
  ps.sum = v3(0.f, 0.f, 0.f);
  ps.s = 0;
  bool alive = mine;
  if (alive) begin_path(cam, tp, pixel, px, py, ps);
  while (alive) {
    Hit h = scene_trace<STATS, LB>(S, ps.ray, c, my_stack);
    n_seg++;
    alive = finish_segment(S, cam, film, tp, pixel, px, py, ps, h, n_shade);
  }
  if (mine) film_add(film, pixel, ps.sum, tp.n_samples);
  warp_add(stats + ST_PATHS, mine ? ps.s : 0u);
  warp_add(stats + ST_SEGMENTS, n_seg);
  warp_add(stats + ST_SHADE, n_shade);
  if (STATS) {
    warp_add(stats + ST_NODE, c.n_node);
    warp_add(stats + ST_TRI, c.n_tri);
    warp_add(stats + ST_BARY, c.n_bary);
    warp_add(stats + ST_PRIM, c.n_prim);
    warp_add(stats + ST_LBNODE, c.n_lbnode);
    warp_add(stats + ST_TRITEST, c.n_tritest);
  }
}

// Variant B (default) — same per-pixel ownership and path regeneration, but the traversal
// of every lane is a resumable state machine and mesh leaves are scanned by the whole warp
// (fs2_coop.cuh).
template <bool STATS>
__global__ void __launch_bounds__(128)
k_trace_coop(DScene S, CamBasis cam, Film film, TraceParams tp,
             unsigned long long* __restrict__ stats) {
  int px, py;
  const bool mine = pixel_of_thread(film.width, film.height, px, py) && tp.n_samples > 0;
  const uint32_t pixel = mine ? (uint32_t)(py * film.width + px) : 0u;
  Counters c = {0, 0, 0, 0, 0, 0};
  uint32_t n_seg = 0, n_shade = 0;
  PathState ps;
  ps.sum = v3(0.f, 0.f, 0.f);
  ps.s = 0;
  ps.ray = make_ray(v3(0.f, 0.f, 0.f), v3(0.f, 0.f, 0.f));
  Frame sstack[FS2_SCENE_STACK];
  Frame mstack[FS2_MESH_STACK];
  LaneWalk w;
  w.pc = PC_IDLE;
  w.rec = no_hit();
  w.glob = no_hit();
  w.ret = false;
  w.resume = false;
  w.leaf_first = w.leaf_count = 0;
  w.cur_obj = 0;
  if (mine) {
    begin_path(cam, tp, pixel, px, py, ps);
    w.pc = PC_SEG_START;
  }
  while (__any_sync(0xFFFFFFFFu, w.pc != PC_IDLE)) {
    lane_advance<STATS>(S, w, ps.ray, sstack, mstack, c);
    warp_scan_leaves<STATS>(S, w, ps.ray, c);
    if (w.pc == PC_SEG_DONE) {
      n_seg++;
      bool alive = finish_segment(S, cam, film, tp, pixel, px, py, ps, w.rec, n_shade);
      w.pc = alive ? PC_SEG_START : PC_IDLE;
    }
  }
  if (mine) film_add(film, pixel, ps.sum, tp.n_samples);
  warp_add(stats + ST_PATHS, mine ? ps.s : 0u);
  warp_add(stats + ST_SEGMENTS, n_seg);
  warp_add(stats + ST_SHADE, n_shade);
  if (STATS) {
    warp_add(stats + ST_NODE, c.n_node);
    warp_add(stats + ST_TRI, c.n_tri);
    warp_add(stats + ST_BARY, c.n_bary);
    warp_add(stats + ST_PRIM, c.n_prim);
  }
}

// Variant C (default) — warp-synchronous phase machine (fs2_phase.cuh) over the leaf BVH.
// Same per-pixel ownership and path regeneration as the other variants.
#define FS2_PHASE_N(STATE, MAXREP, MINLANES, CALL)                       \
  _Pragma("unroll 1") for (int rep_ = 0; rep_ < (MAXREP); ++rep_) {      \
    unsigned m_ = __ballot_sync(0xFFFFFFFFu, w.st == (STATE));           \
    if (!m_ || (rep_ && __popc(m_) < (MINLANES))) break;                 \
    if (w.st == (STATE)) {                                               \
      CALL;                                                              \
    }                                                                    \
  }
#define FS2_PHASE(STATE, MAXREP, CALL) FS2_PHASE_N(STATE, MAXREP, FS2_PH_REPEAT_MIN, CALL)

template <bool STATS>
__global__ void __launch_bounds__(FS2_PH_THREADS, 4)
k_trace_phase(DScene S, CamBasis cam, Film film, TraceParams tp,
              unsigned long long* __restrict__ stats) {
  extern __shared__ uint32_t ph_smem[];
  PhSmem sm;
  sm.col = ph_smem + threadIdx.x;
  int px, py;
  const bool mine = pixel_of_thread(film.width, film.height, px, py) && tp.n_samples > 0;
  const uint32_t pixel = mine ? (uint32_t)(py * film.width + px) : 0u;
  Counters c = {0, 0, 0, 0, 0, 0};
  uint32_t n_seg = 0, n_shade = 0;
  PathState ps;
  ps.sum = v3(0.f, 0.f, 0.f);
  ps.s = 0;
  ps.ray = make_ray(v3(0.f, 0.f, 0.f), v3(0.f, 0.f, 0.f));
  PhLane w;
  w.st = PH_IDLE;
  w.rec = no_hit();
  w.glob = no_hit();
  w.obj_i = w.cur_obj = w.sleaf_first = w.sleaf_count = w.sleaf_i = w.cur = 0;
  w.msp = w.ssp = 0;
  w.flags = 0;
  w.lb_cur = w.lb_root = 0;
  w.lb_sp = 0;
  w.qbase = w.qcount = w.qmask = 0;
  LbRay q = make_lbray(ps.ray);
  if (mine) {
    begin_path(cam, tp, pixel, px, py, ps);
    w.st = PH_SEG_START;
  }
  for (;;) {
    if (w.st == PH_SEG_START) {  // scene.cpp:227-237
      w.rec = no_hit();
      w.flags = lb_origin_far(S, ps.ray.start) ? PH_F_FAR : 0u;
      w.ssp = 0;
      q = make_lbray(ps.ray);
      if (S.scene_root < 0) {
        w.obj_i = 0;
        w.st = PH_OBJ;
      } else {
        w.cur = (uint32_t)S.scene_root;
        w.st = PH_ENTER;
      }
    }
    FS2_PHASE(PH_OBJ, 3, ph_object<STATS>(S, w, ps.ray, c))
    FS2_PHASE(PH_ENTER, 2, ph_enter<STATS>(S, w, ps.ray, sm, c))
    FS2_PHASE(PH_LB_NODE, 2, ph_lb_node<STATS>(S, w, ps.ray, q, sm, c))
    FS2_PHASE(PH_LB_TRIS, 2, ph_lb_tris<STATS>(S, w, ps.ray, sm, c))
    FS2_PHASE(PH_RETURN, 4, ph_return<STATS>(S, w, ps.ray, sm, c))
    unsigned m_shade = __ballot_sync(0xFFFFFFFFu, w.st == PH_SHADE);
    unsigned m_busy = __ballot_sync(0xFFFFFFFFu, w.st != PH_SHADE && w.st != PH_IDLE);
    if (m_shade && (__popc(m_shade) >= FS2_PH_SHADE_BATCH || !m_busy)) {
      if (w.st == PH_SHADE) {
        n_seg++;
        bool alive = finish_segment(S, cam, film, tp, pixel, px, py, ps, w.rec, n_shade);
        w.st = alive ? PH_SEG_START : PH_IDLE;
      }
    } else if (!m_busy) {
      break;  // every lane is idle
    }
  }
  if (mine) film_add(film, pixel, ps.sum, tp.n_samples);
  warp_add(stats + ST_PATHS, mine ? ps.s : 0u);
  warp_add(stats + ST_SEGMENTS, n_seg);
  warp_add(stats + ST_SHADE, n_shade);
  if (STATS) {
    warp_add(stats + ST_NODE, c.n_node);
    warp_add(stats + ST_TRI, c.n_tri);
    warp_add(stats + ST_BARY, c.n_bary);
    warp_add(stats + ST_PRIM, c.n_prim);
    warp_add(stats + ST_LBNODE, c.n_lbnode);
    warp_add(stats + ST_TRITEST, c.n_tritest);
  }
}

}  // namespace fs2
#include "fs2_wavefront.cuh"
namespace fs2 {

// aux planes -> the layouts DisplayFrame exposes (frame.h:86-91): normal[3], depth, 64-bit id
__global__ void __launch_bounds__(256)
k_unpack_aux(const float4* __restrict__ aux, const uint32_t* __restrict__ ids, int n,
             float* __restrict__ normal, float* __restrict__ depth,
             unsigned long long* __restrict__ id64) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 a = aux[i];
  normal[3 * i + 0] = a.x;
  normal[3 * i + 1] = a.y;
  normal[3 * i + 2] = a.z;
  depth[i] = a.w;
  id64[i] = ids[i];
}

// Multi-device film combine (fs_create_multi): runs on the lead device and reads the other
// devices' films through NVLink peer pointers (or staged copies of them), in device order, so
// the N-GPU film is a deterministic function of the per-device films.
//   sample sharding: every device holds partial sums of every pixel: sum += sum of peers; the
//                    aux planes stay the lead's (it owns the last sample, frame.cpp:78-85)
//   tile sharding:   pixel (x, y) lives on device (tx + ty) mod n only: add that one film's
//                    value and take its aux planes (traffic: one film in total, not n - 1)
#define FS2_MAX_DEVICES 16
struct PeerFilms {
  const float4* sum[FS2_MAX_DEVICES - 1];
  const float4* aux[FS2_MAX_DEVICES - 1];
  const uint32_t* mat_id[FS2_MAX_DEVICES - 1];
  int n_peers;
  int tile_n;  // 0: sample sharding
};

__global__ void __launch_bounds__(256)
k_film_combine(float4* __restrict__ sum, float4* __restrict__ aux, uint32_t* __restrict__ mat_id,
               PeerFilms P, int width, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (P.tile_n > 1) {
    int x = i % width, y = i / width;
    int owner = ((x >> 4) + (y >> 3)) % P.tile_n;
    if (owner == 0) return;
    float4 a = sum[i], b = __ldcs(P.sum[owner - 1] + i);
    sum[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
    if (b.w > 0.0f) {
      aux[i] = __ldcs(P.aux[owner - 1] + i);
      mat_id[i] = __ldcs(P.mat_id[owner - 1] + i);
    }
    return;
  }
  float4 a = sum[i];
  for (int p = 0; p < P.n_peers; p++) {
    float4 b = __ldcs(P.sum[p] + i);
    a = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  }
  sum[i] = a;
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
