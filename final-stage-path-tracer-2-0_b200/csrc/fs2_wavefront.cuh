// fs2_wavefront.cuh — wavefront form of the path tracer: traversal and shading are separate
// kernels that exchange rays / hits through path-state planes in HBM.
//
// Why (ncu, profiles/r1_d_*, r1_e_*): the megakernel is ~100 KB of SASS (9 material classes,
// double-precision islands, IEEE divisions, the traversal machine).  Blackwell's instruction
// caches are ~6 KB L0 / 32 KB L1.5 per SM, so warps that sit in different phases keep
// refetching code from L2: 34-56 % of all warp stall samples are "no instruction", issue
// slots are 27-34 % busy, and DRAM moves 0.01 % of peak.  Splitting at the segment boundary
// makes the traversal loop a small kernel and pays ~250 B of state traffic per segment —
// bandwidth this path otherwise leaves idle.
//
// A *pool* of `cap` paths is in flight (not one per pixel: profiles/r1_f showed that with one
// ray per pixel the wave count is the longest pixel's segment count, 8 x spp, and the sparse
// tail waves cost as much as full ones).  Path ids = (sample, pixel slot) are handed out from
// one 64-bit counter, sample-major and in pixel-tile order, so a warp that regenerates lanes
// gets neighbouring pixels of the same sample.
//
//   k_wf_shade<true>   fills the pool with the first `cap` camera paths
//   k_wf_flip          device-side counter hand-over between waves (no host round trip)
//   k_wf_trace         persistent warps; lanes pull rays from the current state array in
//                      chunks and walk them with the phase machine of fs2_phase.cuh; writes
//                      the nearest hit
//   k_wf_shade<false>  one thread per live path: TraceStep's bookkeeping for the finished
//                      segment (shade_segment, fs2_kernels.cuh).  A finished path adds its
//                      radiance to the film (one 16-byte vector atomic) and the lane takes
//                      the next path id; live lanes write their state *compacted* into the
//                      other state array (ballot + one atomic per warp), so every wave reads
//                      and writes dense, coalesced planes and there are no index lists.
//
// The per-path arithmetic is the megakernel's, bit for bit, and so is every path's RNG stream
// (keyed by seed, pixel, sample); only the order in which a pixel's samples are added to the
// film differs (float atomics), so films agree to rounding, not bit for bit.
#pragma once
#include "fs2_phase.cuh"

namespace fs2 {

#define FS2_WF_PLANES 6          // float4 planes of per-path state
#define FS2_WF_CHUNK 128         // most ray indices a warp reserves per atomic
#ifndef FS2_EXP_ITERS
#define FS2_EXP_ITERS 0  // diagnostic: per-ray loop iteration statistics in the stats build
#endif
#ifndef FS2_WF_CHUNK_MIN
#define FS2_WF_CHUNK_MIN 8       // fewest (a power of two): the end of a wave and sparse waves; 1080p @ 8 spp: 32 -> 26.16, 16 -> 25.90, 8 -> 25.83 ms
#endif
#define FS2_WF_MAX_PATHS (1u << 25)  // pool size bound (state: 2 x 96 B + 20 B per path = 7.1 GB); measured on
                                     // config 2 at r1_l: 4 M -> 492, 8 M -> 520, 16 M -> 537, 32 M -> 544, 64 M -> 543 Mpaths/s

// counters (uint32): 0 continuing paths of this wave, 1 continuing paths of the next wave,
//                    2 fetch cursor of k_wf_trace, 3 current state array, 6 waves run,
//                    7 new camera paths of the next wave, 8 new camera paths of this wave
// (Tried and dropped: tracing the bounce rays in direction-octant order through per-octant index
// lists changed nothing, 383.5 vs 386.3 Mpaths/s: lanes diverge by traversal stage, not by the
// order in which they visit octree children.)
// A wave's array holds the continuing (bounce) paths from the front and the camera paths
// started by the last shade kernel from the back, so warps of k_wf_trace pull either
// coherent camera rays or bounce rays, not a mix of both.
struct WfBuffers {
  float4* state[2];   // [FS2_WF_PLANES][cap], ping-pong
  float4* hit;        // [cap] param, object, prim, u   (same index as the ray of this wave)
  float* hit_v;       // [cap]
  uint32_t* counters;
  unsigned long long* next_path;  // next path id; shared by the pipelines of one fs_trace
  uint32_t* host_live;  // pinned host word: live paths at the start of the newest wave
  uint32_t cap;         // paths in flight
  uint32_t pix_cap;     // pixel slots = padded 16x8 tiles x 128 (this shard's tiles only)
  uint32_t tiles_x;     // tiles per row of this shard's slot space
  unsigned long long total;  // path ids: pix_cap x n_samples (ids of padding pixels are skipped)
  // interleaved tile sharding (fs_set_tile_shard): this context renders the 16x8-pixel tiles
  // (tx, ty) with (tx + ty) mod tile_n == tile_r; slot space row ty holds tiles_x of them
  uint32_t tile_n, tile_r;
};

// plane layout
//  0: ray.start.xyz, ray.dir.x      1: ray.stop.xyz, ray.dir.y
//  2: ray.dir.z, meta, pixel slot, sample      meta = zero_length | depth << 8 | fog_pending << 16
//  3: L.xyz, T.x                    4: T.y, T.z, rng lo, rng hi
//  5: fog_pos.xyz, fog_mat          (written / read only while a fog decision is pending)

__device__ __forceinline__ void wf_slot_pixel(const WfBuffers& B, uint32_t slot, int& px, int& py) {
  uint32_t block = slot >> 7, t = slot & 127u;
  uint32_t lane = t & 31u, warp = t >> 5;
  uint32_t tx = block % B.tiles_x, ty = block / B.tiles_x;
  if (B.tile_n > 1u)  // k-th tile of row ty owned by this shard (diagonal interleave)
    tx = tx * B.tile_n + (B.tile_r + B.tile_n - ty % B.tile_n) % B.tile_n;
  px = (int)(tx * 16u + (warp & 1u) * 8u + (lane & 7u));
  py = (int)(ty * 8u + (warp >> 1) * 4u + (lane >> 3));
}

// physical index of the i-th path of a wave with n_sec continuing paths
__device__ __forceinline__ uint32_t wf_phys(uint32_t i, uint32_t n_sec, uint32_t cap) {
  return i < n_sec ? i : cap - 1u - (i - n_sec);
}

__device__ __forceinline__ void wf_store_path(float4* st, uint32_t cap, uint32_t i,
                                              const PathState& ps, uint32_t pixslot) {
  float4* p = st + i;
  const size_t c = cap;
  uint32_t meta = (ps.ray.zero_length ? 1u : 0u) | (ps.depth << 8) | (ps.fog_pending ? 65536u : 0u);
  __stcs(p, make_float4(ps.ray.start.x, ps.ray.start.y, ps.ray.start.z, ps.ray.dir.x));
  __stcs(p + c, make_float4(ps.ray.stop.x, ps.ray.stop.y, ps.ray.stop.z, ps.ray.dir.y));
  __stcs(p + 2 * c, make_float4(ps.ray.dir.z, __uint_as_float(meta), __uint_as_float(pixslot),
                                __uint_as_float(ps.s)));
  __stcs(p + 3 * c, make_float4(ps.L.x, ps.L.y, ps.L.z, ps.T.x));
  __stcs(p + 4 * c, make_float4(ps.T.y, ps.T.z, __uint_as_float((uint32_t)ps.rng.s),
                                __uint_as_float((uint32_t)(ps.rng.s >> 32))));
  if (ps.fog_pending)
    __stcs(p + 5 * c, make_float4(ps.fog_pos.x, ps.fog_pos.y, ps.fog_pos.z, __uint_as_float(ps.fog_mat)));
}

__device__ __forceinline__ Ray wf_load_ray(const float4* st, uint32_t cap, uint32_t i) {
  const float4* p = st + i;
  const size_t c = cap;
  float4 a = __ldcs(p), b = __ldcs(p + c), d = __ldcs(p + 2 * c);
  Ray r;
  r.start = v3(a.x, a.y, a.z);
  r.stop = v3(b.x, b.y, b.z);
  r.dir = v3(a.w, b.w, d.x);
  r.zero_length = (__float_as_uint(d.y) & 1u) != 0u;
  return r;
}

__device__ __forceinline__ void wf_load_path(const WfBuffers& B, const float4* st, uint32_t i,
                                             PathState& ps, uint32_t& pixslot, Hit& h) {
  const float4* p = st + i;
  const size_t c = B.cap;
  float4 a = __ldcs(p), b = __ldcs(p + c), d = __ldcs(p + 2 * c);
  float4 e = __ldcs(p + 3 * c), f = __ldcs(p + 4 * c);
  float4 hh = __ldcs(B.hit + i);
  float hv = __ldcs(B.hit_v + i);
  ps.ray.start = v3(a.x, a.y, a.z);
  ps.ray.stop = v3(b.x, b.y, b.z);
  ps.ray.dir = v3(a.w, b.w, d.x);
  uint32_t meta = __float_as_uint(d.y);
  ps.ray.zero_length = (meta & 1u) != 0u;
  ps.depth = (meta >> 8) & 255u;
  ps.fog_pending = (meta & 65536u) != 0u;
  pixslot = __float_as_uint(d.z);
  ps.s = __float_as_uint(d.w);
  ps.L = v3(e.x, e.y, e.z);
  ps.T = v3(e.w, f.x, f.y);
  ps.rng.s = (uint64_t)__float_as_uint(f.z) | ((uint64_t)__float_as_uint(f.w) << 32);
  ps.sum = v3(0.f, 0.f, 0.f);
  ps.fog_pos = v3(0.f, 0.f, 0.f);
  ps.fog_mat = 0;
  if (ps.fog_pending) {
    float4 g = __ldcs(p + 5 * c);
    ps.fog_pos = v3(g.x, g.y, g.z);
    ps.fog_mat = __float_as_uint(g.w);
  }
  h.param = hh.x;
  h.object = __float_as_uint(hh.y);
  h.prim = __float_as_uint(hh.z);
  h.u = hh.w;
  h.v = hv;
}

__device__ __forceinline__ void wf_store_hit(const WfBuffers& B, uint32_t i, const Hit& h) {
  __stcs(B.hit + i, make_float4(h.param, __uint_as_float(h.object), __uint_as_float(h.prim), h.u));
  __stcs(B.hit_v + i, h.v);
}

// ------------------------------------------------------------------ kernels
// (PathState, begin_path, shade_segment, warp_add: fs2_kernels.cuh, which includes this file
// after defining them.)

// next wave's counts become this wave's, cursor <- 0, swap the state arrays; tells the host how
// many paths the wave it starts has.
__global__ void k_wf_flip(uint32_t* counters, uint32_t* host_live) {
  uint32_t n_sec = counters[1], n_pri = counters[7];
  counters[0] = n_sec;
  counters[8] = n_pri;
  counters[1] = 0;
  counters[7] = 0;
  counters[2] = 0;
  counters[3] ^= 1u;
  counters[6] += 1u;
  *(volatile uint32_t*)host_live = n_sec + n_pri;
  __threadfence_system();
}

// most consecutive executions of a phase per loop iteration; a repeat needs FS2_WF_REPEAT_MIN
// lanes still in the phase.  Measured on config 2 (Mpaths/s): min 12 / reps 3,2,2,2,4 -> 352;
// min 20 -> 342; no repeats -> 369; min 6 -> 382; min 6 / reps 3,3,3,3,4 -> 386; min 4 -> 376;
// min 3 -> 368; min 1 -> 359.
#ifndef FS2_WF_REPEAT_MIN
#define FS2_WF_REPEAT_MIN 6
#endif
#ifndef FS2_WF_REP_OBJ
#define FS2_WF_REP_OBJ 3
#endif
#ifndef FS2_WF_REP_OBJ_TREE
#define FS2_WF_REP_OBJ_TREE 64  // scene-tree leaves list tens of objects: with runs of 4 spheres per step, 16 -> 383,
                                // 32 -> 404, 64 / 128 / 512 -> 423 Mpaths/s on the 303-object scene
#endif
#ifndef FS2_WF_REP_ENTER
#define FS2_WF_REP_ENTER 3
#endif
#ifndef FS2_WF_REP_NODE
#define FS2_WF_REP_NODE 3
#endif
#ifndef FS2_WF_REP_TRIS
#define FS2_WF_REP_TRIS 3
#endif
#ifndef FS2_WF_REP_RETURN
#define FS2_WF_REP_RETURN 4
#endif
#ifndef FS2_WF_COOP_OBJ
#define FS2_WF_COOP_OBJ 0  // 1: scene-tree leaves tested by 4 ... 32 lanes per owner (ph_object_coop).  Measured on the
                           // 303-object scene: 146 vs 196 Mpaths/s per lane — the object loop dominates that scene, so
                           // most lanes of a warp already sit in PH_OBJ and the shuffles only add work
#endif
#ifndef FS2_WF_REPEAT_MIN_OBJ
#define FS2_WF_REPEAT_MIN_OBJ 2
#endif
#ifndef FS2_WF_COOP_NODE
#define FS2_WF_COOP_NODE 0  // 1: leaf-BVH nodes with the four child boxes on four lanes (ph_lb_node_coop).  Measured:
                            // config 2 623 vs 723 Mpaths/s per lane, grid4m 467 vs 613 — masked-off lanes cost nothing,
                            // the 12 shuffles, 3 ballots and two warp syncs of a pass cost more than the divergence
#endif
#if FS2_WF_COOP_NODE && FS2_PH_LB_STACK < 18
#error "ph_lb_node_coop has no overflow fallback: it needs the worst-case stack (FS2_PH_LB_STACK >= 18)"
#endif
#ifndef FS2_WF_REPEAT_MIN_NODE
#define FS2_WF_REPEAT_MIN_NODE 3
#endif
#ifndef FS2_WF_COOP_TRIS
#define FS2_WF_COOP_TRIS 1  // triangle groups tested by 4 lanes each (ph_lb_tris_coop)
#endif
#ifndef FS2_WF_MIN_BLOCKS
#define FS2_WF_MIN_BLOCKS 8  // measured (no scene tree, Mpaths/s): r1_h 5 -> 400, 6 -> 433, 7 -> 434, 8 -> 398; r1_k (152 B/lane stacks) 6 -> 472, 7 -> 479; r1_l (leaf best kept in the record) 6 -> 491, 7 -> 504, 8 -> 521 (64 regs)
#endif
// TREE = false: the scene has no scene octree (< 182 objects), which lets the
// compiler drop the scene-leaf bookkeeping (second hit record, leaf cursor) from the loop.
template <bool STATS, bool TREE, int OBJB = 1>
// (with a scene tree the per-lane stacks are 252 B: shared memory holds 7 CTAs per SM, so that
// specialisation may use the registers of 7)
__global__ void __launch_bounds__(FS2_PH_THREADS, TREE ? (FS2_WF_MIN_BLOCKS < 7 ? FS2_WF_MIN_BLOCKS : 7) : FS2_WF_MIN_BLOCKS)
k_wf_trace(DScene S, WfBuffers B, TraceParams tp, unsigned long long* __restrict__ stats) {
  extern __shared__ uint32_t ph_smem[];
  __shared__ uint8_t s_slot_owner[FS2_PH_THREADS / 32][8];
  PhSmem sm;
  sm.col = ph_smem + threadIdx.x;
#if FS2_WF_STAGE_TOP
  {
    // experiment (north_star: "shared-memory or TMA staging of hot octree nodes"): the root and
    // first level of the first eight objects' mesh trees, 288 contiguous bytes each
    __shared__ float4 s_top[FS2_STAGE_OBJECTS * 2 * FS2_STAGE_NODES];
    __shared__ uint32_t s_top_root[FS2_STAGE_OBJECTS];
    if (threadIdx.x < FS2_STAGE_OBJECTS) {
      uint32_t root = 0xFFFFFFFFu;
      if (threadIdx.x < S.n_objects) {
        const float4 w0 = obj_word(S, threadIdx.x, 0);
        if (__float_as_uint(w0.x) == FS_OBJ_MESH) {
          const uint32_t r = __float_as_uint(w0.z);
          if (r != 0xFFFFFFFFu) {
            const NodeRec n = load_node(S, r);
            if (n.count == FS_NODE_INTERNAL && n.first == r + 1u) root = r;
          }
        }
      }
      s_top_root[threadIdx.x] = root;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < FS2_STAGE_OBJECTS * 2 * FS2_STAGE_NODES; i += FS2_PH_THREADS) {
      const uint32_t r = s_top_root[i / (2 * FS2_STAGE_NODES)];
      if (r != 0xFFFFFFFFu) s_top[i] = __ldg(S.nodes + 2 * (size_t)r + i % (2 * FS2_STAGE_NODES));
    }
    __syncthreads();
    sm.top = s_top;
    sm.top_root = s_top_root;
  }
#endif
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const uint32_t n_sec = B.counters[0];
  const uint32_t n_in = n_sec + B.counters[8];
  const float4* st = B.state[B.counters[3] & 1u];
  const uint32_t n_warps = gridDim.x * (FS2_PH_THREADS / 32);

  Counters c = {0, 0, 0, 0, 0, 0};
  Ray ray = make_ray(v3(0.f, 0.f, 0.f), v3(0.f, 0.f, 0.f));
  LbRay q = make_lbray(ray);
  uint32_t idx = 0;
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
  uint32_t chunk_next = 0, chunk_end = 0;
  bool exhausted = false;
  const uint32_t fetch_min = max(1u, tp.tune & 31u);
#if FS2_EXP_ITERS
  uint32_t iters = 0;  // diagnostic build: loop iterations of this lane's current ray
#endif
  for (;;) {
#if FS2_EXP_ITERS
    iters++;
#endif
    // ---- fetch: idle lanes take the next rays of the warp's reserved chunk
    unsigned idle = __ballot_sync(FULL, w.st == PH_IDLE);
    if (idle && ((uint32_t)__popc(idle) >= fetch_min || idle == FULL || exhausted)) {
      if (chunk_next == chunk_end && !exhausted) {
        // guided self-scheduling: chunks shrink with the rays that are left, so the wave does
        // not end with a few warps still holding 128 rays each
        uint32_t b = 0, chunk = 0;
        if (lane == 0) {
          uint32_t seen = *(volatile uint32_t*)(B.counters + 2);
          uint32_t left = seen < n_in ? n_in - seen : 0u;
          chunk = min((uint32_t)FS2_WF_CHUNK,
                      max((uint32_t)FS2_WF_CHUNK_MIN, (left / (2u * n_warps)) & ~((uint32_t)FS2_WF_CHUNK_MIN - 1u)));
          b = atomicAdd(B.counters + 2, chunk);
        }
        b = __shfl_sync(FULL, b, 0);
        chunk = __shfl_sync(FULL, chunk, 0);
        if (b >= n_in) {
          exhausted = true;
        } else {
          chunk_next = b;
          chunk_end = min(b + chunk, n_in);
        }
      }
      uint32_t avail = chunk_end - chunk_next;
      if (avail) {
        uint32_t rank = (uint32_t)__popc(idle & ((1u << lane) - 1u));
        if (w.st == PH_IDLE && rank < avail) {
          idx = wf_phys(chunk_next + rank, n_sec, B.cap);
          ray = wf_load_ray(st, B.cap, idx);
          q = ph_segment_start<TREE>(S, w, ray);
#if FS2_EXP_ITERS
          iters = 0;
#endif
        }
        chunk_next += min((uint32_t)__popc(idle), avail);
      } else if (exhausted && idle == FULL) {
        break;
      }
    }
    // scene-tree leaves hold tens of objects (the reference's tree is shallow): stay in the phase
    if (TREE && FS2_WF_COOP_OBJ) {  // (TREE kernels are only launched for scenes with a scene tree)
#pragma unroll 1
      for (int rep_ = 0; rep_ < FS2_WF_REP_OBJ_TREE; ++rep_) {
        unsigned m_ = __ballot_sync(FULL, w.st == PH_OBJ);
        if (!m_ || (rep_ && __popc(m_) < FS2_WF_REPEAT_MIN_OBJ)) break;
        ph_object_coop<STATS>(S, w, ray, c, m_, s_slot_owner[threadIdx.x >> 5]);
      }
    } else {
      FS2_PHASE_N(PH_OBJ, (TREE ? FS2_WF_REP_OBJ_TREE : FS2_WF_REP_OBJ), FS2_WF_REPEAT_MIN,
                  (ph_object<STATS, TREE, OBJB>(S, w, ray, c)))
    }
    FS2_PHASE_N(PH_ENTER, FS2_WF_REP_ENTER, FS2_WF_REPEAT_MIN, (ph_enter<STATS, TREE>(S, w, ray, sm, c)))
#if FS2_WF_COOP_NODE
#pragma unroll 1
    for (int rep_ = 0; rep_ < FS2_WF_REP_NODE; ++rep_) {
      unsigned m_ = __ballot_sync(FULL, w.st == PH_LB_NODE);
      if (!m_ || (rep_ && __popc(m_) < FS2_WF_REPEAT_MIN_NODE)) break;
      ph_lb_node_coop<STATS>(S, w, ray, q, sm, c, m_, s_slot_owner[threadIdx.x >> 5]);
    }
#else
    FS2_PHASE_N(PH_LB_NODE, FS2_WF_REP_NODE, FS2_WF_REPEAT_MIN, ph_lb_node<STATS>(S, w, ray, q, sm, c))
#endif
#if FS2_WF_COOP_TRIS
#pragma unroll 1
    for (int rep_ = 0; rep_ < FS2_WF_REP_TRIS; ++rep_) {
      unsigned m_ = __ballot_sync(FULL, w.st == PH_LB_TRIS);
      if (!m_ || (rep_ && __popc(m_) < FS2_WF_REPEAT_MIN)) break;
      ph_lb_tris_coop<STATS>(S, w, ray, sm, c, m_, s_slot_owner[threadIdx.x >> 5]);
    }
#else
    FS2_PHASE_N(PH_LB_TRIS, FS2_WF_REP_TRIS, FS2_WF_REPEAT_MIN, ph_lb_tris<STATS>(S, w, ray, sm, c))
#endif
    FS2_PHASE_N(PH_RETURN, FS2_WF_REP_RETURN, FS2_WF_REPEAT_MIN, (ph_return<STATS, TREE>(S, w, ray, sm, c)))
    if (w.st == PH_SHADE) {  // nearest hit of the segment is final
#if FS2_EXP_ITERS
      if (STATS) {  // the work counters are borrowed: (sum of iterations, rays) for depth 0 / 1 / >= 2
        const uint32_t depth = (__float_as_uint(__ldcs(st + idx + 2 * (size_t)B.cap).y) >> 8) & 255u;
        const int k = depth == 0u ? 0 : depth == 1u ? 1 : 2;
        atomicAdd(stats + (k == 0 ? ST_NODE : k == 1 ? ST_BARY : ST_LBNODE), (unsigned long long)iters);
        atomicAdd(stats + (k == 0 ? ST_TRI : k == 1 ? ST_PRIM : ST_TRITEST), 1ull);
      }
#endif
      wf_store_hit(B, idx, w.rec);
      w.st = PH_IDLE;
    }
  }
  if (STATS && !FS2_EXP_ITERS) {
    warp_add(stats + ST_NODE, c.n_node);
    warp_add(stats + ST_TRI, c.n_tri);
    warp_add(stats + ST_BARY, c.n_bary);
    warp_add(stats + ST_PRIM, c.n_prim);
    warp_add(stats + ST_LBNODE, c.n_lbnode);
    warp_add(stats + ST_TRITEST, c.n_tritest);
  }
}

// FRESH: fill the pool (every thread starts a path).  Otherwise: finish the traced segment of
// live path `idx`, then continue it, or replace it by the next path id.
#ifndef FS2_WF_SHADE_MIN_BLOCKS
#define FS2_WF_SHADE_MIN_BLOCKS 8  // measured (config 3 scene / config 2 / config 1, Mpaths/s): 84 regs 3219 / 715 / 4040, 6 CTAs 3402 / 721 / 4234, 7 CTAs 3459 / 720 / 4281, 8 CTAs (64 regs) 3512 / 717 / 4247
#endif
template <bool FRESH>
__global__ void __launch_bounds__(128, FS2_WF_SHADE_MIN_BLOCKS)
k_wf_shade(DScene S, CamBasis cam, Film film, TraceParams tp, WfBuffers B,
           unsigned long long* __restrict__ stats) {
  const unsigned FULL = 0xFFFFFFFFu;
  const uint32_t n_sec = FRESH ? 0u : B.counters[0];
  const uint32_t n_in = FRESH ? B.cap : n_sec + B.counters[8];
  if (blockIdx.x * 128u >= n_in) return;  // whole block past the live paths
  const uint32_t cur = B.counters[3] & 1u;
  const uint32_t idx = wf_phys(blockIdx.x * 128u + threadIdx.x, n_sec, B.cap);
  const int lane = threadIdx.x & 31;
  const bool have = blockIdx.x * 128u + threadIdx.x < n_in;
  bool alive = false, need = FRESH && have, fresh = false;
  uint32_t n_shade = 0, finished = 0, pixslot = 0;
  PathState ps;
  ps.sum = v3(0.f, 0.f, 0.f);
  ps.s = 0;
  ps.fog_mat = 0;
  ps.fog_pos = v3(0.f, 0.f, 0.f);
  if (!FRESH && have) {
    Hit h;
    wf_load_path(B, B.state[cur], idx, ps, pixslot, h);
    int px, py;
    wf_slot_pixel(B, pixslot, px, py);
    const uint32_t pixel = (uint32_t)(py * film.width + px);
    if (shade_segment(S, cam, film, tp, pixel, ps.s + 1 == tp.n_samples, ps, h, n_shade)) {
      atomicAdd(film.sum + pixel, make_float4(ps.L.x, ps.L.y, ps.L.z, 1.0f));
      finished = 1;
      need = true;
    } else {
      alive = true;
    }
  }
  // Counter traffic is aggregated per CTA: with one atomic per warp on three shared addresses
  // the kernel was bound by same-address atomic latency (ncu r1_h: 82 % long-scoreboard stalls
  // on the atomics and the shuffles that wait for them).  Barriers are the next cost (ncu r2_c:
  // 20 % of the stall samples): the common case — no padding pixels — takes five of them.
  __shared__ uint32_t s_cnt[4][2];   // regeneration round: requests per warp
  __shared__ uint32_t s_cmp[4][2];   // compaction: continuing / fresh paths per warp
  __shared__ unsigned long long s_base64;
  __shared__ uint32_t s_base[2];
  __shared__ uint32_t s_stat[3];
  if (threadIdx.x < 3) s_stat[threadIdx.x] = 0;  // (first used after the loop's first barrier)
  const int warp = threadIdx.x >> 5;
  const unsigned below = (1u << lane) - 1u;
  // ---- regeneration: the threads that need a path take consecutive ids
  for (;;) {
    unsigned m = __ballot_sync(FULL, need);
    if (lane == 0) s_cnt[warp][0] = (uint32_t)__popc(m);
    __syncthreads();
    const uint32_t c0 = s_cnt[0][0], c1 = s_cnt[1][0], c2 = s_cnt[2][0], c3 = s_cnt[3][0];
    const uint32_t tot = c0 + c1 + c2 + c3;
    if (!tot) break;  // CTA-uniform
    if (threadIdx.x == 0) s_base64 = atomicAdd(B.next_path, (unsigned long long)tot);
    __syncthreads();
    if (need) {
      uint32_t before = (warp > 0 ? c0 : 0u) + (warp > 1 ? c1 : 0u) + (warp > 2 ? c2 : 0u);
      unsigned long long id = s_base64 + before + (unsigned)__popc(m & below);
      if (id >= B.total) {
        need = false;  // no camera paths left
      } else {
        ps.s = (uint32_t)(id / B.pix_cap);
        pixslot = (uint32_t)(id % B.pix_cap);
        int px, py;
        wf_slot_pixel(B, pixslot, px, py);
        if (px < film.width && py < film.height) {  // else: padding pixel, take another id
          begin_path(cam, tp, (uint32_t)(py * film.width + px), px, py, ps);
          alive = fresh = true;
          need = false;
        }
      }
    }
    // another round only when somebody drew a padding pixel (the barrier also orders the reads
    // of s_cnt / s_base64 above before the next round rewrites them)
    if (!__syncthreads_or(need ? 1 : 0)) break;
  }
  // ---- compaction: live threads write their state densely into the other array, bounce
  // paths from the front, new camera paths from the back; statistics ride on the same barriers
  unsigned ms = __ballot_sync(FULL, alive && !fresh), mp = __ballot_sync(FULL, alive && fresh);
  uint32_t f = 0, hv = 0, sh = 0;
  if (!FRESH) {
    f = __reduce_add_sync(FULL, finished);
    hv = __reduce_add_sync(FULL, have ? 1u : 0u);
    sh = __reduce_add_sync(FULL, n_shade);
  }
  if (lane == 0) {
    s_cmp[warp][0] = (uint32_t)__popc(ms);
    s_cmp[warp][1] = (uint32_t)__popc(mp);
    if (!FRESH) {
      atomicAdd(&s_stat[0], f);
      atomicAdd(&s_stat[1], hv);
      atomicAdd(&s_stat[2], sh);
    }
  }
  __syncthreads();
  if (threadIdx.x < 2) {
    uint32_t t = s_cmp[0][threadIdx.x] + s_cmp[1][threadIdx.x] + s_cmp[2][threadIdx.x] + s_cmp[3][threadIdx.x];
    s_base[threadIdx.x] = t ? atomicAdd(B.counters + (threadIdx.x ? 7 : 1), t) : 0u;
  } else if (!FRESH && threadIdx.x == 32) {
    // statistics: one atomic per CTA and counter
    atomicAdd(stats + ST_PATHS, (unsigned long long)s_stat[0]);
    atomicAdd(stats + ST_SEGMENTS, (unsigned long long)s_stat[1]);
    atomicAdd(stats + ST_SHADE, (unsigned long long)s_stat[2]);
  }
  __syncthreads();
  if (alive) {
    const int k = fresh ? 1 : 0;
    uint32_t off = s_base[k] + (warp > 0 ? s_cmp[0][k] : 0u) + (warp > 1 ? s_cmp[1][k] : 0u) +
                   (warp > 2 ? s_cmp[2][k] : 0u) + (uint32_t)__popc((fresh ? mp : ms) & below);
    wf_store_path(B.state[cur ^ 1u], B.cap, fresh ? B.cap - 1u - off : off, ps, pixslot);
  }
}

}  // namespace fs2
