// fs2_wavefront.cuh — wavefront form of the path tracer: traversal and shading are separate
// kernels that exchange rays / hits through path-state planes in HBM.
//
// Why (ncu, profiles/r1_d_*, r1_e_*): the megakernel is ~100 KB of SASS (9 material classes,
// double-precision islands, IEEE divisions, the traversal machine).  Blackwell's instruction
// caches are ~6 KB L0 / 32 KB L1.5 per SM, so warps that sit in different phases keep
// refetching code from L2: 34-56 % of all warp stall samples are "no instruction", issue
// slots are 27-34 % busy, and DRAM moves 0.01 % of peak.  Splitting at the segment boundary
// makes the traversal loop a small kernel and pays ~300 B of state traffic per segment —
// bandwidth this path otherwise leaves idle.
//
//   k_wf_begin   one thread per pixel slot: first camera ray, compacted list of live slots
//   k_wf_flip    device-side counter hand-over between waves (no host round trip)
//   k_wf_trace   persistent warps; lanes pull rays from the live list in chunks and walk
//                them with the phase machine of fs2_phase.cuh; writes the nearest hit
//   k_wf_shade   one thread per live slot: TraceStep's bookkeeping for the finished segment
//                (finish_segment, fs2_kernels.cuh), next ray or next sample, film add when the
//                pixel's samples are done; live slots are re-compacted with ballot + one
//                atomic per warp
//
// Slot = pixel (tile order of pixel_of_thread), so a slot always has exactly one ray in
// flight until its samples are exhausted; arithmetic is the megakernel's, bit for bit.
#pragma once
#include "fs2_phase.cuh"

namespace fs2 {

#define FS2_WF_PLANES 9      // float4 planes of per-slot state
#define FS2_WF_CHUNK 128     // ray indices a warp reserves per atomic

struct WfBuffers {
  float4* planes;      // [FS2_WF_PLANES][cap]
  uint32_t* list_in;   // live slots of this wave
  uint32_t* list_out;  // live slots of the next wave
  uint32_t* counters;  // 0: n_in, 1: n_out, 2: fetch cursor
  uint32_t cap;        // slots = padded pixel tiles
  uint32_t tiles_x;
};

// plane layout
//  0: ray.start.xyz, ray.dir.x      1: ray.stop.xyz, ray.dir.y     2: ray.dir.z, zero_length, meta, s
//  3: L.xyz, T.x                    4: T.y, T.z, sum.x, sum.y      5: sum.z, rng lo, rng hi, fog_mat
//  6: fog_pos.xyz, -                7: hit param, object, prim, u  8: hit v, -, -, -
// meta = depth | fog_pending << 8

__device__ __forceinline__ void wf_slot_pixel(const WfBuffers& B, uint32_t slot, int& px, int& py) {
  uint32_t block = slot >> 7, t = slot & 127u;
  uint32_t lane = t & 31u, warp = t >> 5;
  px = (int)((block % B.tiles_x) * 16u + (warp & 1u) * 8u + (lane & 7u));
  py = (int)((block / B.tiles_x) * 8u + (warp >> 1) * 4u + (lane >> 3));
}

__device__ __forceinline__ void wf_store_ray(const WfBuffers& B, uint32_t slot, const PathState& ps) {
  float4* p = B.planes + slot;
  p[0 * (size_t)B.cap] = make_float4(ps.ray.start.x, ps.ray.start.y, ps.ray.start.z, ps.ray.dir.x);
  p[1 * (size_t)B.cap] = make_float4(ps.ray.stop.x, ps.ray.stop.y, ps.ray.stop.z, ps.ray.dir.y);
  uint32_t meta = ps.depth | (ps.fog_pending ? 256u : 0u);
  p[2 * (size_t)B.cap] = make_float4(ps.ray.dir.z, __uint_as_float(ps.ray.zero_length ? 1u : 0u),
                                     __uint_as_float(meta), __uint_as_float(ps.s));
}

__device__ __forceinline__ void wf_store_path(const WfBuffers& B, uint32_t slot, const PathState& ps) {
  wf_store_ray(B, slot, ps);
  float4* p = B.planes + slot;
  p[3 * (size_t)B.cap] = make_float4(ps.L.x, ps.L.y, ps.L.z, ps.T.x);
  p[4 * (size_t)B.cap] = make_float4(ps.T.y, ps.T.z, ps.sum.x, ps.sum.y);
  p[5 * (size_t)B.cap] = make_float4(ps.sum.z, __uint_as_float((uint32_t)ps.rng.s),
                                     __uint_as_float((uint32_t)(ps.rng.s >> 32)),
                                     __uint_as_float(ps.fog_mat));
  if (ps.fog_pending) p[6 * (size_t)B.cap] = make_float4(ps.fog_pos.x, ps.fog_pos.y, ps.fog_pos.z, 0.f);
}

__device__ __forceinline__ Ray wf_load_ray(const WfBuffers& B, uint32_t slot) {
  const float4* p = B.planes + slot;
  float4 a = p[0], b = p[1 * (size_t)B.cap], c = p[2 * (size_t)B.cap];
  Ray r;
  r.start = v3(a.x, a.y, a.z);
  r.stop = v3(b.x, b.y, b.z);
  r.dir = v3(a.w, b.w, c.x);
  r.zero_length = __float_as_uint(c.y) != 0u;
  return r;
}

__device__ __forceinline__ void wf_load_path(const WfBuffers& B, uint32_t slot, PathState& ps, Hit& h) {
  const float4* p = B.planes + slot;
  float4 a = p[0], b = p[1 * (size_t)B.cap], c = p[2 * (size_t)B.cap];
  float4 d = p[3 * (size_t)B.cap], e = p[4 * (size_t)B.cap], f = p[5 * (size_t)B.cap];
  float4 hh = p[7 * (size_t)B.cap], hi = p[8 * (size_t)B.cap];
  ps.ray.start = v3(a.x, a.y, a.z);
  ps.ray.stop = v3(b.x, b.y, b.z);
  ps.ray.dir = v3(a.w, b.w, c.x);
  ps.ray.zero_length = __float_as_uint(c.y) != 0u;
  uint32_t meta = __float_as_uint(c.z);
  ps.depth = meta & 255u;
  ps.fog_pending = (meta & 256u) != 0u;
  ps.s = __float_as_uint(c.w);
  ps.L = v3(d.x, d.y, d.z);
  ps.T = v3(d.w, e.x, e.y);
  ps.sum = v3(e.z, e.w, f.x);
  ps.rng.s = (uint64_t)__float_as_uint(f.y) | ((uint64_t)__float_as_uint(f.z) << 32);
  ps.fog_mat = __float_as_uint(f.w);
  ps.fog_pos = v3(0.f, 0.f, 0.f);
  if (ps.fog_pending) {
    float4 g = p[6 * (size_t)B.cap];
    ps.fog_pos = v3(g.x, g.y, g.z);
  }
  h.param = hh.x;
  h.object = __float_as_uint(hh.y);
  h.prim = __float_as_uint(hh.z);
  h.u = hh.w;
  h.v = hi.x;
}

__device__ __forceinline__ void wf_store_hit(const WfBuffers& B, uint32_t slot, const Hit& h) {
  float4* p = B.planes + slot;
  p[7 * (size_t)B.cap] = make_float4(h.param, __uint_as_float(h.object), __uint_as_float(h.prim), h.u);
  p[8 * (size_t)B.cap] = make_float4(h.v, 0.f, 0.f, 0.f);
}

// Appends the slots of the lanes with `alive` to list_out: ballot + one atomic per warp.
__device__ __forceinline__ void wf_append(const WfBuffers& B, bool alive, uint32_t slot) {
  unsigned m = __ballot_sync(0xFFFFFFFFu, alive);
  if (!m) return;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(m) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(B.counters + 1, (uint32_t)__popc(m));
  base = __shfl_sync(0xFFFFFFFFu, base, leader);
  if (alive) B.list_out[base + __popc(m & ((1u << lane) - 1u))] = slot;
}

// ------------------------------------------------------------------ kernels
// (PathState, begin_path, finish_segment, film_add, warp_add: fs2_kernels.cuh, which
// includes this file after defining them.)

__global__ void __launch_bounds__(128)
k_wf_begin(CamBasis cam, Film film, TraceParams tp, WfBuffers B) {
  uint32_t slot = blockIdx.x * 128u + threadIdx.x;
  int px, py;
  wf_slot_pixel(B, slot, px, py);
  bool mine = slot < B.cap && px < film.width && py < film.height && tp.n_samples > 0;
  if (mine) {
    PathState ps;
    ps.sum = v3(0.f, 0.f, 0.f);
    ps.s = 0;
    ps.fog_mat = 0;
    ps.fog_pos = v3(0.f, 0.f, 0.f);
    begin_path(cam, tp, (uint32_t)(py * film.width + px), px, py, ps);
    wf_store_path(B, slot, ps);
  }
  wf_append(B, mine, slot);
}

// n_in <- n_out, n_out <- 0, cursor <- 0; the lists are swapped by the host (pointer swap).
__global__ void k_wf_flip(uint32_t* counters) {
  counters[0] = counters[1];
  counters[1] = 0;
  counters[2] = 0;
}

template <bool STATS>
__global__ void __launch_bounds__(FS2_PH_THREADS, 4)
k_wf_trace(DScene S, WfBuffers B, TraceParams tp, unsigned long long* __restrict__ stats) {
  extern __shared__ uint32_t ph_smem[];
  PhSmem sm;
  sm.col = ph_smem + threadIdx.x;
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const uint32_t n_in = B.counters[0];
  Counters c = {0, 0, 0, 0, 0, 0};
  Ray ray = make_ray(v3(0.f, 0.f, 0.f), v3(0.f, 0.f, 0.f));
  LbRay q = make_lbray(ray);
  uint32_t slot = 0;
  PhLane w;
  w.st = PH_IDLE;
  w.rec = no_hit();
  w.glob = no_hit();
  w.obj_i = w.cur_obj = w.sleaf_first = w.sleaf_count = w.sleaf_i = w.cur = 0;
  w.msp = w.ssp = 0;
  w.flags = 0;
  w.lb_cur = 0;
  w.lb_sp = 0;
  w.bt = w.bu = w.bv = 0.f;
  w.border = w.qbase = w.qcount = w.qmask = 0;
  uint32_t chunk_next = 0, chunk_end = 0;
  bool exhausted = false;
  const uint32_t fetch_min = max(1u, tp.tune & 31u);
  for (;;) {
    // ---- fetch: idle lanes take the next rays of the warp's reserved chunk
    unsigned idle = __ballot_sync(FULL, w.st == PH_IDLE);
    if (idle && ((uint32_t)__popc(idle) >= fetch_min || idle == FULL || exhausted)) {
      if (chunk_next == chunk_end && !exhausted) {
        uint32_t b = 0;
        if (lane == 0) b = atomicAdd(B.counters + 2, (uint32_t)FS2_WF_CHUNK);
        b = __shfl_sync(FULL, b, 0);
        if (b >= n_in) {
          exhausted = true;
        } else {
          chunk_next = b;
          chunk_end = min(b + (uint32_t)FS2_WF_CHUNK, n_in);
        }
      }
      uint32_t avail = chunk_end - chunk_next;
      if (avail) {
        uint32_t rank = (uint32_t)__popc(idle & ((1u << lane) - 1u));
        if (w.st == PH_IDLE && rank < avail) {
          slot = B.list_in[chunk_next + rank];
          ray = wf_load_ray(B, slot);
          q = ph_segment_start(S, w, ray);
        }
        chunk_next += min((uint32_t)__popc(idle), avail);
      } else if (exhausted && idle == FULL) {
        break;
      }
    }
    FS2_PHASE(PH_OBJ, 3, ph_object<STATS>(S, w, ray, c))
    FS2_PHASE(PH_ENTER, 2, ph_enter<STATS>(S, w, ray, sm, c))
    FS2_PHASE(PH_LB_NODE, 2, ph_lb_node<STATS>(S, w, ray, q, sm, c))
    FS2_PHASE(PH_LB_TRIS, 2, ph_lb_tris<STATS>(S, w, ray, sm, c))
    FS2_PHASE(PH_RETURN, 4, ph_return(S, w, ray, sm))
    if (w.st == PH_SHADE) {  // nearest hit of the segment is final
      wf_store_hit(B, slot, w.rec);
      w.st = PH_IDLE;
    }
  }
  if (STATS) {
    warp_add(stats + ST_NODE, c.n_node);
    warp_add(stats + ST_TRI, c.n_tri);
    warp_add(stats + ST_BARY, c.n_bary);
    warp_add(stats + ST_PRIM, c.n_prim);
    warp_add(stats + ST_LBNODE, c.n_lbnode);
    warp_add(stats + ST_TRITEST, c.n_tritest);
  }
}

__global__ void __launch_bounds__(128)
k_wf_shade(DScene S, CamBasis cam, Film film, TraceParams tp, WfBuffers B,
           unsigned long long* __restrict__ stats) {
  const uint32_t n_in = B.counters[0];
  uint32_t idx = blockIdx.x * 128u + threadIdx.x;
  if (blockIdx.x * 128u >= n_in) return;  // whole block past the live list
  bool have = idx < n_in;
  bool alive = false;
  uint32_t slot = 0, n_shade = 0, finished = 0;
  if (have) {
    slot = B.list_in[idx];
    int px, py;
    wf_slot_pixel(B, slot, px, py);
    const uint32_t pixel = (uint32_t)(py * film.width + px);
    PathState ps;
    Hit h;
    wf_load_path(B, slot, ps, h);
    uint32_t s_before = ps.s;
    alive = finish_segment(S, cam, film, tp, pixel, px, py, ps, h, n_shade);
    finished = ps.s - s_before;
    if (alive) wf_store_path(B, slot, ps);
    else film_add(film, pixel, ps.sum, tp.n_samples);
  }
  wf_append(B, alive, slot);
  warp_add(stats + ST_PATHS, finished);
  warp_add(stats + ST_SEGMENTS, have ? 1u : 0u);
  warp_add(stats + ST_SHADE, n_shade);
}

}  // namespace fs2
