// fs2_cabi.cu — libfs2cuda.so: the C ABI of include/fs2_cabi.h over the sm_100a kernels.
//
// Host side only does plumbing: validation, uploads, the per-frame camera basis
// (reference source/engine.cpp:170-197, evaluated with the host libm exactly like the
// reference does), launches, read-back.  There is no CPU tracing path in this library:
// without a CUDA device every compute entry point fails with FS_E_CUDA.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false
//        -Xcompiler -fPIC,-ffp-contract=off -shared   (see ../build.py)
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cub/device/device_scan.cuh>

#include "fs2_kernels.cuh"
#include "fs2_build.cuh"
#include "fs2_post.cuh"
#include "fs2_selftest.cuh"
#include "fs2_lbvh_build.h"

using namespace fs2;

namespace {

thread_local char g_error[512] = "";

struct DeviceBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

}  // namespace

#define FS2_WF_AHEAD 4  // waves kept enqueued per pipe
#define FS2_WF_PIPES 8  // most independent parts of the path pool (one stream each); ctx->pipes are used

struct fs_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;      // stream in use
  cudaStream_t own_stream = nullptr;  // created by fs_create
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evk0 = nullptr, evk1 = nullptr;
  bool timing_pending = false;
  char error[512] = "";
  // scene
  bool have_scene = false;
  DeviceBuf sref_sphere;  // per scene ref: centre + radius of a sphere, NaN radius otherwise
  DeviceBuf nodes, scene_refs, tri_plane, tri_aux, faces, objects, materials, textures, texels,
      lut, lb_nodes, lb_leaf, lb_plane, lb_aux, lb_order;
  DScene ds;
  uint32_t n_nodes = 0, n_tris = 0, n_faces = 0, n_materials = 0, n_objects = 0;
  float lb_m_floor = 1.0f;  // coordinate scale the leaf-BVH pads were sized for
  // film
  int width = 0, height = 0;
  DeviceBuf sum, aux, mat_id, mean, count, rgb8, scratch, auxpack, post;
  DeviceBuf wf_planes, wf_hit, wf_counters;  // wavefront path state
  uint32_t* wf_host_live = nullptr;   // pinned: live paths of the newest wave, per pipe
  cudaEvent_t wf_events[FS2_WF_AHEAD * FS2_WF_PIPES] = {};
  cudaStream_t wf_streams[FS2_WF_PIPES] = {};  // [0] unused (the context's stream), [p] pipe p
  cudaStream_t wf_stream = nullptr;            // == wf_streams[1]: "the pipeline has been set up"
  cudaEvent_t wf_fork = nullptr, wf_join[FS2_WF_PIPES] = {};
  bool prims_only = false;  // the resident scene lists no mesh (k_wf_trace's batched object loop)
  int pipes = 2;  // FS2_PIPES: parts the pool is split into
  // Sparse waves (the drain of a frame): a persistent k_wf_trace that fills every SM slot keeps the
  // other pipe's k_wf_shade — and with it that pipe's next wave — waiting until its own rays are
  // done, so the two pipes alternate instead of overlapping.  Below drain_live live paths the
  // traversal grid takes drain_ctas CTAs per SM only.
  int drain_ctas = 0;            // FS2_DRAIN_CTAS (0: always the full grid)
  uint32_t drain_live = 2u << 20;  // FS2_DRAIN_LIVE (paths)
  int sm_count = 148;
  // stats
  unsigned long long* d_stats = nullptr;
  uint64_t launches = 0;
  float last_trace_ms = 0.f, last_kernel_ms = 0.f;
  bool stats_build = false;
  const char* timeline = nullptr;  // FS2_TIMELINE=file: per-launch device time stamps of the wavefront pipeline
  uint32_t tune = 8;  // FS2_TUNE: bits 0-4 idle lanes that trigger a refill in k_wf_trace, bits 8.. pool size (1024 paths)
  // interleaved tile sharding of this context (fs_set_tile_shard); tile_n <= 1: every tile
  uint32_t tile_n = 1, tile_r = 0;
  // multi-device lead (fs_create_multi): the contexts on the other devices, in device order
  std::vector<fs_ctx*> peers;
  std::vector<int> peer_direct;  // 1: the lead reads that peer's film through a peer mapping
  DeviceBuf stage;               // staged copies of peer films when there is no peer mapping
  fs_ctx* lead = nullptr;        // set on peers
};

namespace {

int fail(fs_ctx* ctx, int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof g_error, fmt, ap);
  va_end(ap);
  if (ctx) memcpy(ctx->error, g_error, sizeof g_error);
  return code;
}

#define CUDA_TRY(ctx, call)                                                               \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail(ctx, FS_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                  __LINE__);                                                              \
  } while (0)

int dev_alloc(fs_ctx* ctx, DeviceBuf* b, size_t bytes) {
  if (b->p && b->bytes >= bytes && bytes) return FS_OK;
  if (b->p) {
    cudaFree(b->p);
    b->p = nullptr;
    b->bytes = 0;
  }
  if (!bytes) bytes = 16;  // keep pointers valid for empty arrays
  CUDA_TRY(ctx, cudaMalloc(&b->p, bytes));
  b->bytes = bytes;
  return FS_OK;
}

int dev_upload(fs_ctx* ctx, DeviceBuf* b, const void* src, size_t bytes) {
  int rc = dev_alloc(ctx, b, bytes);
  if (rc) return rc;
  if (bytes) CUDA_TRY(ctx, cudaMemcpyAsync(b->p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  return FS_OK;
}

void dev_free(DeviceBuf* b) {
  if (b->p) cudaFree(b->p);
  b->p = nullptr;
  b->bytes = 0;
}

// ---- host restatement of the reference's approximate vector helpers, used only for
// the per-frame camera basis (reference source/math/scalar.h:233-245, vector3.h:166-187)
struct H3 {
  float x, y, z;
};
float h_inv_sqrt(float f) {
  float half = 0.5f * f;
  int32_t i;
  memcpy(&i, &f, 4);
  i = 0x5f3759df - (i >> 1);
  memcpy(&f, &i, 4);
  return f * (1.5f - half * f * f);
}
H3 h_normalize(H3 a) {
  float l = 1.0f / h_inv_sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
  if (l == 0.0f) l = 1.0f;
  return H3{a.x / l, a.y / l, a.z / l};
}
H3 h_cross(H3 a, H3 b) {
  return H3{(a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x)};
}

// reference source/engine.cpp:170-197 (and :300-313 for TraceRange)
CamBasis camera_basis(const fs_camera* cam, int w, int h) {
  CamBasis b;
  memset(&b, 0, sizeof b);
  float width = (float)w, height = (float)h;
  float aspect = width / height;
  float fovy = cam->fov_y * FS2_PI / 180.0;
  float fovx = 2.0 * atan(tan(fovy * 0.5) * aspect);
  H3 origin{cam->origin[0], cam->origin[1], cam->origin[2]};
  H3 target{cam->target[0], cam->target[1], cam->target[2]};
  H3 fwd = h_normalize(H3{target.x - origin.x, target.y - origin.y, target.z - origin.z});
  H3 right = h_normalize(h_cross(H3{0.f, 1.f, 0.f}, fwd));
  H3 up = h_normalize(h_cross(fwd, right));
  b.half_h = tanf(fovy * 0.5f) * cam->z_far;
  b.half_w = tanf(fovx * 0.5f) * cam->z_far;
  H3 proj{origin.x + fwd.x * cam->z_far, origin.y + fwd.y * cam->z_far,
          origin.z + fwd.z * cam->z_far};
  // focal plane: calculate_plane(-forward, origin + forward * focal_depth) (engine.cpp:195-197)
  H3 n{fwd.x * -1.0f, fwd.y * -1.0f, fwd.z * -1.0f};
  H3 fp{origin.x + fwd.x * cam->focal_depth, origin.y + fwd.y * cam->focal_depth,
        origin.z + fwd.z * cam->focal_depth};
  float d = -(n.x * fp.x + n.y * fp.y + n.z * fp.z);
  float* dst[5] = {b.origin, b.right, b.up, b.forward, b.proj_origin};
  H3 src[5] = {origin, right, up, fwd, proj};
  for (int k = 0; k < 5; k++) {
    dst[k][0] = src[k].x;
    dst[k][1] = src[k].y;
    dst[k][2] = src[k].z;
  }
  b.focal_plane[0] = n.x;
  b.focal_plane[1] = n.y;
  b.focal_plane[2] = n.z;
  b.focal_plane[3] = d;
  b.wm1 = width - 1;
  b.hm1 = height - 1;
  b.aperture = cam->aperture_size;
  b.z_far = cam->z_far;
  b.fast_render = cam->fast_render_enabled;
  return b;
}

int validate_scene(fs_ctx* ctx, const fs_scene_desc* s) {
  if (s->abi_version != FS2_ABI_VERSION)
    return fail(ctx, FS_E_INVALID, "scene abi_version %u != %d", s->abi_version, FS2_ABI_VERSION);
  if (s->n_materials == 0 || s->sky_material >= s->n_materials)
    return fail(ctx, FS_E_INVALID, "sky_material %u out of range (%u materials)", s->sky_material,
                s->n_materials);
  if (s->n_lut == 0 || !s->lut) return fail(ctx, FS_E_INVALID, "normal LUT missing");
  if ((s->n_nodes && !s->nodes) || (s->n_tris && (!s->tri_plane || !s->tri_aux)) ||
      (s->n_faces && !s->faces) || (s->n_objects && !s->objects) || !s->materials ||
      (s->n_textures && (!s->textures || !s->texels)) || (s->n_scene_refs && !s->scene_refs))
    return fail(ctx, FS_E_INVALID, "null array with non-zero count");
  if (s->n_nodes >= (1u << FS2_NODE_INDEX_BITS))
    return fail(ctx, FS_E_INVALID, "more than 2^%d octree nodes", FS2_NODE_INDEX_BITS);
  if (s->scene_root >= 0 && (uint32_t)s->scene_root >= s->n_nodes)
    return fail(ctx, FS_E_INVALID, "scene_root %d out of range", s->scene_root);
  for (uint32_t i = 0; i < s->n_nodes; i++) {
    const fs_node& n = s->nodes[i];
    if (n.count == FS_NODE_INTERNAL) {
      if ((uint64_t)n.first + 8 > s->n_nodes)
        return fail(ctx, FS_E_INVALID, "node %u: children out of range", i);
    }
  }
  // Walk every tree once: children must follow their parent (no cycles), no node may be reached
  // twice, leaf ranges must lie inside the arrays they index, and the depth must fit the
  // kernels' frame stacks (a deeper subtree would silently read as a miss).
  {
    std::vector<uint8_t> seen(s->n_nodes, 0);
    struct Item {
      uint32_t node, depth;
    };
    std::vector<Item> stack;
    auto walk = [&](uint32_t root, bool mesh, uint32_t owner) -> int {
      const uint32_t max_internal_depth = mesh ? FS2_PH_MESH_FRAMES : FS2_PH_SCENE_FRAMES;
      const uint64_t ref_limit = mesh ? s->n_tris : s->n_scene_refs;
      stack.clear();
      stack.push_back(Item{root, 0});
      while (!stack.empty()) {
        Item it = stack.back();
        stack.pop_back();
        if (seen[it.node])
          return fail(ctx, FS_E_INVALID, "node %u is reachable twice (%s tree of %u)", it.node,
                      mesh ? "mesh" : "scene", owner);
        seen[it.node] = 1;
        const fs_node& n = s->nodes[it.node];
        if (n.count == FS_NODE_INTERNAL) {
          if (it.depth >= max_internal_depth)
            return fail(ctx, FS_E_INVALID, "%s tree of %u is deeper than %u levels at node %u",
                        mesh ? "mesh" : "scene", owner, max_internal_depth, it.node);
          if (n.first <= it.node)
            return fail(ctx, FS_E_INVALID, "node %u: children (%u) do not follow their parent", it.node,
                        n.first);
          for (uint32_t c = 0; c < 8; c++) stack.push_back(Item{n.first + c, it.depth + 1});
        } else if ((uint64_t)n.first + n.count > ref_limit) {
          return fail(ctx, FS_E_INVALID, "node %u: leaf range %u+%u exceeds the %llu %s", it.node, n.first,
                      n.count, (unsigned long long)ref_limit, mesh ? "triangle records" : "scene refs");
        }
      }
      return FS_OK;
    };
    int rc;
    if (s->scene_root >= 0 && (rc = walk((uint32_t)s->scene_root, false, 0))) return rc;
    for (uint32_t i = 0; i < s->n_objects; i++) {
      const fs_object& o = s->objects[i];
      if (o.type == FS_OBJ_MESH && o.u[0] != 0xFFFFFFFFu && o.u[0] < s->n_nodes)
        if ((rc = walk(o.u[0], true, i))) return rc;
    }
  }
  for (uint32_t i = 0; i < s->n_objects; i++) {
    const fs_object& o = s->objects[i];
    if (o.type > FS_OBJ_MESH) return fail(ctx, FS_E_INVALID, "object %u: bad type %u", i, o.type);
    if (o.material >= s->n_materials)
      return fail(ctx, FS_E_INVALID, "object %u: material %u out of range", i, o.material);
    if (o.type == FS_OBJ_MESH && o.u[0] != 0xFFFFFFFFu) {
      if (o.u[0] >= s->n_nodes) return fail(ctx, FS_E_INVALID, "object %u: mesh root out of range", i);
      if ((uint64_t)o.u[1] + o.u[3] > s->n_faces)
        return fail(ctx, FS_E_INVALID, "object %u: face range out of bounds", i);
    }
  }
  for (uint32_t i = 0; i < s->n_scene_refs; i++)
    if (s->scene_refs[i] >= s->n_objects)
      return fail(ctx, FS_E_INVALID, "scene ref %u out of range", i);
  for (uint32_t i = 0; i < s->n_tris; i++)
    if ((s->tri_aux[i].face & 0x3FFFFFFFu) >= s->n_faces)
      return fail(ctx, FS_E_INVALID, "triangle %u: face index out of range", i);
  for (uint32_t i = 0; i < s->n_materials; i++) {
    const fs_material& m = s->materials[i];
    if (m.type > FS_MAT_FOG) return fail(ctx, FS_E_INVALID, "material %u: bad type", i);
    if (m.texture >= (int32_t)s->n_textures)
      return fail(ctx, FS_E_INVALID, "material %u: texture out of range", i);
  }
  for (uint32_t i = 0; i < s->n_textures; i++) {
    const fs_texture& t = s->textures[i];
    if (!t.width || !t.height || t.offset + (uint64_t)t.width * t.height * 3 > s->n_texels)
      return fail(ctx, FS_E_INVALID, "texture %u: out of range", i);
  }
  return FS_OK;
}

int set_device(fs_ctx* ctx) {
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return FS_OK;
}

dim3 pixel_grid(int w, int h) { return dim3((w + 15) / 16, (h + 7) / 8); }

// Reads the event pairs of the last fs_trace; the stream must have been synchronised.
int collect_timing(fs_ctx* ctx) {
  if (!ctx->timing_pending) return FS_OK;
  CUDA_TRY(ctx, cudaEventElapsedTime(&ctx->last_trace_ms, ctx->ev0, ctx->ev1));
  CUDA_TRY(ctx, cudaEventElapsedTime(&ctx->last_kernel_ms, ctx->evk0, ctx->evk1));
  ctx->timing_pending = false;
  return FS_OK;
}

// Wavefront pipeline (fs2_wavefront.cuh): fill the pool, then per wave k_wf_flip / k_wf_trace /
// k_wf_shade until no path is live.
//  * The pool is split into FS2_WF_PIPES independent halves on two streams that draw path ids
//    from one counter: while the persistent traversal kernel of one half drains (its last
//    warps finish their rays), the CTAs of the other half's kernel take over the SMs that
//    become free, so the tail of every wave is overlapped instead of idling the machine.
//  * The host never drains a stream: k_wf_flip publishes the live count of the wave it starts
//    in a pinned host word, the host keeps FS2_WF_AHEAD waves enqueued per pipe (event ring)
//    and stops enqueueing once a wave reported zero live paths; the waves already in the
//    queue then exit at once.
int launch_wavefront_waves(fs_ctx* ctx, const CamBasis& cb, const Film& film, const TraceParams& tp) {
  uint32_t tiles_x = (uint32_t)(film.width + 15) / 16;
  const uint32_t tiles_y = (uint32_t)(film.height + 7) / 8;
  const uint32_t tile_n = ctx->tile_n > 1 ? ctx->tile_n : 1u;
  if (tile_n > 1) tiles_x = (tiles_x + tile_n - 1) / tile_n;  // this shard's tiles per row
  const uint32_t pix_cap = tiles_x * tiles_y * 128u;
  const unsigned long long total = (unsigned long long)pix_cap * tp.n_samples;
  unsigned long long pool = FS2_WF_MAX_PATHS;
  if (ctx->tune >> 8) pool = (unsigned long long)(ctx->tune >> 8) << 10;  // FS2_TUNE bits 8..
  if (pool > total) pool = total;
  bool grow = false;
  unsigned long long fit = FS2_WF_MAX_PATHS;
  if (!(ctx->tune >> 8) && (size_t)pool * 212 > ctx->wf_planes.bytes + ctx->wf_hit.bytes) {
    // the buffers have to grow: never take more than an eighth of what is free for the path
    // state.  Only then — cudaMemGetInfo costs milliseconds when several processes share the
    // driver, which showed as 35 -> 40 ms per frame at 8 GPUs when it ran on every call.
    grow = true;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
      fit = (free_b / 8) / 212;
      if (fit < (1u << 20)) fit = 1u << 20;
      if (pool > fit) pool = fit;
    }
  }
  int pipes = (ctx->tune & 32u) || pool < (1u << 18) ? 1 : ctx->pipes;  // FS2_TUNE bit 5: one pipe
  const uint32_t cap = (uint32_t)(((pool + pipes - 1) / pipes + 127u) & ~127ull);
  // Paths the buffers are sized for.  When they have to grow, the allocation (not the pool in use)
  // is rounded up to a power of two within the same bounds: a context whose jobs get larger step
  // by step reallocates a handful of times, not at every step.  (A reallocation frees and
  // allocates gigabytes; in a multi-device context with peer mappings that is ~100 ms, seen as
  // outliers in the resolution x spp sweep at 4 and 8 GPUs — even for a growth from 33.2 M to
  // 33.5 M paths.)
  size_t alloc_paths = (size_t)cap * pipes;
  if (grow) {
    size_t p2 = 1;
    while (p2 < alloc_paths) p2 <<= 1;
    const size_t bound = (size_t)(fit < FS2_WF_MAX_PATHS ? fit : FS2_WF_MAX_PATHS);
    if (p2 > bound) p2 = bound;
    if (p2 > alloc_paths) alloc_paths = p2;
  }
  int rc;
  if ((rc = dev_alloc(ctx, &ctx->wf_planes, alloc_paths * FS2_WF_PLANES * 16 * 2))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->wf_hit, alloc_paths * 20))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->wf_counters, 64 * (FS2_WF_PIPES + 1)))) return rc;
  if (!ctx->wf_host_live) {
    CUDA_TRY(ctx, cudaHostAlloc((void**)&ctx->wf_host_live, 64, cudaHostAllocMapped));
    for (int i = 0; i < FS2_WF_AHEAD * FS2_WF_PIPES; i++)
      CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->wf_events[i], cudaEventDisableTiming));
    for (int p = 1; p < FS2_WF_PIPES; p++) {
      CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->wf_streams[p], cudaStreamNonBlocking));
      CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->wf_join[p], cudaEventDisableTiming));
    }
    ctx->wf_stream = ctx->wf_streams[1];
    CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->wf_fork, cudaEventDisableTiming));
  }
  WfBuffers B[FS2_WF_PIPES];
  cudaStream_t streams[FS2_WF_PIPES];
  for (int p = 0; p < pipes; p++) {
    B[p].state[0] = (float4*)ctx->wf_planes.p + (size_t)p * 2 * cap * FS2_WF_PLANES;
    B[p].state[1] = B[p].state[0] + (size_t)cap * FS2_WF_PLANES;
    B[p].hit = (float4*)((char*)ctx->wf_hit.p + (size_t)p * cap * 20);
    B[p].hit_v = (float*)(B[p].hit + cap);
    B[p].counters = (uint32_t*)ctx->wf_counters.p + 16 * (p + 1);
    B[p].next_path = (unsigned long long*)ctx->wf_counters.p;
    B[p].host_live = ctx->wf_host_live + p;
    B[p].cap = cap;
    B[p].pix_cap = pix_cap;
    B[p].tiles_x = tiles_x;
    B[p].total = total;
    B[p].tile_n = tile_n;
    B[p].tile_r = ctx->tile_r;
    streams[p] = p == 0 ? ctx->stream : ctx->wf_streams[p];
    *(volatile uint32_t*)(ctx->wf_host_live + p) = 1u;
  }
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->wf_counters.p, 0, 64 * (FS2_WF_PIPES + 1), ctx->stream));
  if (pipes > 1) {
    CUDA_TRY(ctx, cudaEventRecord(ctx->wf_fork, ctx->stream));
    for (int p = 1; p < pipes; p++) CUDA_TRY(ctx, cudaStreamWaitEvent(streams[p], ctx->wf_fork, 0));
  }
  // FS2_TIMELINE (diagnostics only): a timing event after every launch; written out as
  // "pipe wave kernel end_ms live" rows relative to the start of the fill when the frame is done
  struct Stamp {
    int pipe;
    unsigned long long wave;
    const char* what;
    uint32_t live;
    cudaEvent_t ev;
  };
  std::vector<Stamp> stamps;
  cudaEvent_t tl_start = nullptr;
  auto stamp = [&](int p, unsigned long long wave, const char* what, uint32_t live) {
    if (!ctx->timeline) return;
    Stamp s{p, wave, what, live, nullptr};
    cudaEventCreate(&s.ev);
    cudaEventRecord(s.ev, streams[p]);
    stamps.push_back(s);
  };
  if (ctx->timeline) {
    cudaEventCreate(&tl_start);
    cudaEventRecord(tl_start, ctx->stream);
    for (int p = 1; p < pipes; p++) cudaStreamWaitEvent(streams[p], tl_start, 0);
  }
  for (int p = 0; p < pipes; p++) {
    k_wf_shade<true><<<cap / 128u, 128, 0, streams[p]>>>(ctx->ds, cb, film, tp, B[p], ctx->d_stats);
    stamp(p, 0, "fill", cap);
  }
  ctx->launches += pipes;
  // every path has at most 8 segments and the pool is refilled every wave
  const unsigned long long max_waves = 8ull * ((total + pool - 1) / pool) + 8ull;
  const unsigned smem = ph_smem_bytes(ctx->ds.scene_root >= 0);
  int per_sm = 0;
  // specialisations: scene tree or not; without one, brute-force lists that hold no mesh are
  // walked FS2_PH_OBJ_BATCH_PRIMS entries per step (fs2_phase.cuh ph_object)
  void (*trace_kernel)(DScene, WfBuffers, TraceParams, unsigned long long*);
  if (ctx->ds.scene_root >= 0)
    trace_kernel = ctx->stats_build ? k_wf_trace<true, true> : k_wf_trace<false, true>;
  else if (ctx->prims_only)
    trace_kernel = ctx->stats_build ? k_wf_trace<true, false, FS2_PH_OBJ_BATCH_PRIMS>
                                    : k_wf_trace<false, false, FS2_PH_OBJ_BATCH_PRIMS>;
  else
    trace_kernel = ctx->stats_build ? k_wf_trace<true, false> : k_wf_trace<false, false>;
  CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_kernel, FS2_PH_THREADS, smem));
  if (per_sm < 1) per_sm = 1;
  const int persistent = ctx->sm_count * per_sm;  // one resident generation of CTAs
  // k_wf_shade's grid covers the live paths of the wave.  The live count never grows from one
  // wave to the next (a finished path is replaced by at most one new one), so the count a
  // k_wf_flip published for an earlier wave bounds it; once the camera paths run out the tail
  // waves shrink and so do their grids (an empty 131 072-CTA grid alone costs 0.15 ms).
  uint32_t live_bound[FS2_WF_PIPES];
  for (int p = 0; p < FS2_WF_PIPES; p++) live_bound[p] = cap;
  bool done[FS2_WF_PIPES] = {};
  for (unsigned long long wave = 0; wave < max_waves; wave++) {
    int running = 0;
    for (int p = 0; p < pipes; p++) {
      if (done[p]) continue;
      cudaEvent_t ev = ctx->wf_events[p * FS2_WF_AHEAD + wave % FS2_WF_AHEAD];
      if (wave >= FS2_WF_AHEAD) {
        CUDA_TRY(ctx, cudaEventSynchronize(ev));
        const uint32_t live = *(volatile uint32_t*)(ctx->wf_host_live + p);
        if (ctx->timeline) stamps.push_back(Stamp{p, wave - FS2_WF_AHEAD, "live", live, nullptr});
        if (live == 0u) {
          done[p] = true;
          continue;
        }
        if (live < live_bound[p]) live_bound[p] = live;
      }
      const uint32_t shade_blocks = (live_bound[p] + 127u) / 128u;
      running++;
      k_wf_flip<<<1, 1, 0, streams[p]>>>(B[p].counters, B[p].host_live);
      CUDA_TRY(ctx, cudaEventRecord(ev, streams[p]));
      stamp(p, wave, "flip", live_bound[p]);
      int grid = persistent;
      if (ctx->drain_ctas > 0 && ctx->drain_ctas < per_sm && pipes > 1 && live_bound[p] < ctx->drain_live)
        grid = ctx->sm_count * ctx->drain_ctas;
      trace_kernel<<<grid, FS2_PH_THREADS, smem, streams[p]>>>(ctx->ds, B[p], tp, ctx->d_stats);
      stamp(p, wave, "trace", live_bound[p]);
      k_wf_shade<false><<<shade_blocks, 128, 0, streams[p]>>>(ctx->ds, cb, film, tp, B[p], ctx->d_stats);
      stamp(p, wave, "shade", live_bound[p]);
      ctx->launches += 3;
    }
    if (!running) break;
  }
  // The loop ends when every pipe has reported an empty wave.  Leaving it on the wave bound
  // instead would mean paths were dropped (pixel counts below n_samples): report it, never
  // return FS_OK over an incomplete film.
  for (int p = 0; p < pipes; p++)
    if (!done[p]) {
      for (int q = 0; q < pipes; q++) cudaStreamSynchronize(streams[q]);
      if (*(volatile uint32_t*)(ctx->wf_host_live + p) != 0u)
        return fail(ctx, FS_E_CUDA, "fs_trace: pipe %d still had %u live paths after %llu waves", p,
                    *(volatile uint32_t*)(ctx->wf_host_live + p), max_waves);
    }
  for (int p = 1; p < pipes; p++) {
    CUDA_TRY(ctx, cudaEventRecord(ctx->wf_join[p], streams[p]));
    CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ctx->wf_join[p], 0));
  }
  if (ctx->timeline) {
    cudaStreamSynchronize(ctx->stream);
    if (FILE* f = fopen(ctx->timeline, "a")) {
      fprintf(f, "# frame %dx%d, %u samples, pool %llu in %d pipe(s)\n", film.width, film.height,
              tp.n_samples, (unsigned long long)cap * pipes, pipes);
      for (const Stamp& s : stamps) {
        float ms = 0.f;
        if (s.ev) cudaEventElapsedTime(&ms, tl_start, s.ev);
        fprintf(f, "%d %llu %s %.4f %u\n", s.pipe, s.wave, s.what, ms, s.live);
      }
      fclose(f);
    }
    for (const Stamp& s : stamps)
      if (s.ev) cudaEventDestroy(s.ev);
    cudaEventDestroy(tl_start);
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return FS_OK;
}

// A failure part-way leaves waves enqueued on both streams: drain them before reporting, so the
// caller may free or reuse the film and the context's buffers.
int launch_wavefront(fs_ctx* ctx, const CamBasis& cb, const Film& film, const TraceParams& tp) {
  int rc = launch_wavefront_waves(ctx, cb, film, tp);
  if (rc) {
    for (int p = 1; p < FS2_WF_PIPES; p++)
      if (ctx->wf_streams[p]) cudaStreamSynchronize(ctx->wf_streams[p]);
    cudaStreamSynchronize(ctx->stream);
    cudaGetLastError();
  }
  return rc;
}

// Uploads prebuilt leaf BVHs; the host vectors must outlive the asynchronous copies, hence the
// synchronisation.
int upload_lbvh(fs_ctx* ctx, const LbvhHost& lb) {
  int rc;
  if ((rc = dev_upload(ctx, &ctx->lb_nodes, lb.nodes.data(), lb.nodes.size() * sizeof(LbNode)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->lb_leaf, lb.leaf.data(), lb.leaf.size() * sizeof(LbLeaf)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->lb_plane, lb.plane.data(), lb.plane.size() * 4))) return rc;
  if ((rc = dev_upload(ctx, &ctx->lb_aux, lb.aux.data(), lb.aux.size() * sizeof(fs_tri_aux)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->lb_order, lb.order.data(), lb.order.size() * 4))) return rc;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  DScene& d = ctx->ds;
  d.lb_nodes = (const float4*)ctx->lb_nodes.p;
  d.lb_leaf = (const uint2*)ctx->lb_leaf.p;
  d.lb_plane = (const float4*)ctx->lb_plane.p;
  d.lb_aux = (const float4*)ctx->lb_aux.p;
  d.lb_order = (const uint32_t*)ctx->lb_order.p;
  d.lb_origin_limit = lb.origin_limit();
  ctx->lb_m_floor = lb.m_floor;
  return FS_OK;
}

// The lead context followed by its peers (a plain context: itself only).
std::vector<fs_ctx*> all_contexts(fs_ctx* ctx) {
  std::vector<fs_ctx*> all{ctx};
  all.insert(all.end(), ctx->peers.begin(), ctx->peers.end());
  return all;
}

// Device copy of the node array: the `first` word of every internal node gets the occupancy of
// its eight children in its top byte (bit c set = child c is internal or a non-empty leaf), so
// the traversal kernel answers visits to empty leaves from the parent (fs2_phase.cuh).  Node
// indices stay below 2^FS2_NODE_INDEX_BITS (validate_scene).  The ABI's fs_node is unchanged.
std::vector<fs_node> tag_nodes(const fs_scene_desc* s) {
  std::vector<fs_node> t(s->nodes, s->nodes + s->n_nodes);
  for (uint32_t i = 0; i < s->n_nodes; i++) {
    if (t[i].count != FS_NODE_INTERNAL) continue;
    uint32_t mask = 0;
    for (uint32_t c = 0; c < 8; c++)
      if (s->nodes[t[i].first + c].count != 0u) mask |= 1u << c;
    t[i].first |= mask << FS2_NODE_INDEX_BITS;
  }
  return t;
}
void untag_nodes(std::vector<fs_node>& t) {
  for (fs_node& n : t)
    if (n.count == FS_NODE_INTERNAL) n.first &= (1u << FS2_NODE_INDEX_BITS) - 1u;
}

// One device's copy of the (validated) scene.
int scene_upload_one(fs_ctx* ctx, const fs_scene_desc* s, const std::vector<float>& lut4,
                     const LbvhHost& lb) {
  int rc;
  if ((rc = set_device(ctx))) return rc;
  ctx->have_scene = false;
  {
    std::vector<fs_node> tagged = tag_nodes(s);
    if ((rc = dev_upload(ctx, &ctx->nodes, tagged.data(), tagged.size() * sizeof(fs_node)))) return rc;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // `tagged` is pageable and dies here
  }
  if ((rc = dev_upload(ctx, &ctx->scene_refs, s->scene_refs, (size_t)s->n_scene_refs * 4))) return rc;
  {
    // Scene-tree leaves list their objects by index; most of a many-object scene's are spheres.
    // A copy of centre and radius beside every reference lets the leaf loop test a sphere from one
    // 16-byte load instead of the reference -> object header -> payload chain (same float values,
    // same sphere test).
    std::vector<float> sph(4 * (size_t)s->n_scene_refs + 4);
    for (uint32_t k = 0; k < s->n_scene_refs; k++) {
      const fs_object& o = s->objects[s->scene_refs[k]];
      float* q = &sph[4 * (size_t)k];
      if (o.type == FS_OBJ_SPHERE) {
        q[0] = o.f[0], q[1] = o.f[1], q[2] = o.f[2], q[3] = o.f[3];
        if (!(q[3] == q[3])) q[3] = NAN;  // (a NaN radius takes the generic path: same outcome)
      } else {
        q[0] = q[1] = q[2] = 0.f;
        q[3] = NAN;
      }
    }
    if ((rc = dev_upload(ctx, &ctx->sref_sphere, sph.data(), sph.size() * 4))) return rc;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  if ((rc = dev_upload(ctx, &ctx->tri_plane, s->tri_plane, (size_t)s->n_tris * 16))) return rc;
  if ((rc = dev_upload(ctx, &ctx->tri_aux, s->tri_aux, (size_t)s->n_tris * sizeof(fs_tri_aux)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->faces, s->faces, (size_t)s->n_faces * sizeof(fs_face_attr)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->objects, s->objects, (size_t)s->n_objects * sizeof(fs_object)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->materials, s->materials, (size_t)s->n_materials * sizeof(fs_material)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->textures, s->textures, (size_t)s->n_textures * sizeof(fs_texture)))) return rc;
  if ((rc = dev_upload(ctx, &ctx->texels, s->texels, (size_t)s->n_texels * 4))) return rc;
  if ((rc = dev_upload(ctx, &ctx->lut, lut4.data(), lut4.size() * 4))) return rc;
  if ((rc = upload_lbvh(ctx, lb))) return rc;  // synchronises the stream
  DScene& d = ctx->ds;
  d.nodes = (const float4*)ctx->nodes.p;
  d.scene_refs = (const uint32_t*)ctx->scene_refs.p;
  d.sref_sphere = (const float4*)ctx->sref_sphere.p;
  d.tri_plane = (const float4*)ctx->tri_plane.p;
  d.tri_aux = (const float4*)ctx->tri_aux.p;
  d.faces = (const float4*)ctx->faces.p;
  d.objects = (const float4*)ctx->objects.p;
  d.materials = (const float4*)ctx->materials.p;
  d.textures = (const fs_texture*)ctx->textures.p;
  d.texels = (const float*)ctx->texels.p;
  d.lut = (const float4*)ctx->lut.p;
  d.n_objects = s->n_objects;
  ctx->prims_only = true;
  for (uint32_t i = 0; i < s->n_objects; i++)
    if (s->objects[i].type == FS_OBJ_MESH) ctx->prims_only = false;
  d.n_lut = s->n_lut;
  d.sky_material = s->sky_material;
  d.scene_root = s->scene_root;
  ctx->n_nodes = s->n_nodes;
  ctx->n_tris = s->n_tris;
  ctx->n_faces = s->n_faces;
  ctx->n_materials = s->n_materials;
  ctx->n_objects = s->n_objects;
  ctx->have_scene = true;
  return FS_OK;
}

// The leaf-BVH box pads bound hit-point rounding for segments that start within
// DScene::lb_origin_limit of the origin (fs2_lbvh_build.h record_box).  A camera beyond it would
// send every camera ray down the linear leaf scans, so the leaf BVHs are rebuilt once with pads
// sized for that camera (rare: the arrays the builder reads are fetched back from the device).
int ensure_lbvh_for_camera(fs_ctx* ctx, const fs_camera* cam) {
  float m = 0.f;
  for (int k = 0; k < 3; k++) m = fmaxf(m, fabsf(cam->origin[k]));
  m += fabsf(cam->aperture_size);
  if (!(m > ctx->ds.lb_origin_limit) || !(m <= 1.0e30f) || !ctx->n_tris) return FS_OK;
  std::vector<fs_node> nodes(ctx->n_nodes);
  std::vector<float> plane(4 * (size_t)ctx->n_tris);
  std::vector<fs_tri_aux> aux(ctx->n_tris);
  std::vector<fs_object> objects(ctx->n_objects);
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CUDA_TRY(ctx, cudaMemcpy(nodes.data(), ctx->nodes.p, nodes.size() * sizeof(fs_node), cudaMemcpyDeviceToHost));
  untag_nodes(nodes);
  CUDA_TRY(ctx, cudaMemcpy(plane.data(), ctx->tri_plane.p, plane.size() * 4, cudaMemcpyDeviceToHost));
  CUDA_TRY(ctx, cudaMemcpy(aux.data(), ctx->tri_aux.p, aux.size() * sizeof(fs_tri_aux), cudaMemcpyDeviceToHost));
  CUDA_TRY(ctx, cudaMemcpy(objects.data(), ctx->objects.p, objects.size() * sizeof(fs_object), cudaMemcpyDeviceToHost));
  fs_scene_desc d;
  memset(&d, 0, sizeof d);
  d.n_nodes = ctx->n_nodes;
  d.n_tris = ctx->n_tris;
  d.n_objects = ctx->n_objects;
  d.nodes = nodes.data();
  d.tri_plane = plane.data();
  d.tri_aux = aux.data();
  d.objects = objects.data();
  for (int p = 1; p < FS2_WF_PIPES; p++)
    if (ctx->wf_streams[p]) CUDA_TRY(ctx, cudaStreamSynchronize(ctx->wf_streams[p]));
  LbvhHost lb;
  build_lbvh(&d, &lb, 9, 0, m / 2.0f);  // limit = 3 * m_floor >= 1.5 * |camera|
  return upload_lbvh(ctx, lb);
}

#include "fs2_build_host.inc"

}  // namespace

extern "C" {

int fs_abi_version(void) { return FS2_ABI_VERSION; }

const char* fs_last_error(const fs_ctx* ctx) { return ctx ? ctx->error : g_error; }

int fs_create(fs_ctx** out, int device, int flags) {
  if (!out) return fail(nullptr, FS_E_INVALID, "fs_create: null out pointer");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(nullptr, FS_E_CUDA, "no CUDA device: %s", cudaGetErrorString(e));
  if (device < 0 || device >= n)
    return fail(nullptr, FS_E_INVALID, "device %d out of range (%d devices)", device, n);
  fs_ctx* ctx = new (std::nothrow) fs_ctx();
  if (!ctx) return fail(nullptr, FS_E_NOMEM, "out of host memory");
  ctx->device = device;
  ctx->stats_build = (flags & 1) != 0;
  if (const char* e = getenv("FS2_TUNE")) ctx->tune = (uint32_t)strtoul(e, nullptr, 0);
  ctx->timeline = getenv("FS2_TIMELINE");
  if (const char* e = getenv("FS2_DRAIN_CTAS")) ctx->drain_ctas = atoi(e);
  if (const char* e = getenv("FS2_DRAIN_LIVE")) ctx->drain_live = (uint32_t)strtoul(e, nullptr, 0);
  if (const char* e = getenv("FS2_PIPES")) {
    int n = atoi(e);
    ctx->pipes = n < 1 ? 1 : n > FS2_WF_PIPES ? FS2_WF_PIPES : n;
  }
  int rc = set_device(ctx);
  if (rc == FS_OK) {
    cudaError_t e1 = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
    ctx->stream = ctx->own_stream;
    if (e1 == cudaSuccess) e1 = cudaEventCreate(&ctx->ev0);
    if (e1 == cudaSuccess) e1 = cudaEventCreate(&ctx->ev1);
    if (e1 == cudaSuccess) e1 = cudaEventCreate(&ctx->evk0);
    if (e1 == cudaSuccess) e1 = cudaEventCreate(&ctx->evk1);
    if (e1 == cudaSuccess)
      e1 = cudaFuncSetAttribute(k_trace_phase<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                FS2_PH_SMEM_BYTES);
    if (e1 == cudaSuccess)
      e1 = cudaFuncSetAttribute(k_trace_phase<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                FS2_PH_SMEM_BYTES);
    if (e1 == cudaSuccess)
      e1 = cudaFuncSetAttribute(k_wf_trace<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                FS2_PH_SMEM_BYTES);
    if (e1 == cudaSuccess)
      e1 = cudaFuncSetAttribute(k_wf_trace<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                FS2_PH_SMEM_BYTES);
    if (e1 == cudaSuccess)
      e1 = cudaFuncSetAttribute(k_wf_trace<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                FS2_PH_SMEM_BYTES);
    if (e1 == cudaSuccess)
      e1 = cudaFuncSetAttribute(k_wf_trace<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                FS2_PH_SMEM_BYTES);
    if (e1 == cudaSuccess)
      e1 = cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (e1 == cudaSuccess) e1 = cudaMalloc(&ctx->d_stats, ST_COUNT * sizeof(unsigned long long));
    if (e1 == cudaSuccess)
      e1 = cudaMemset(ctx->d_stats, 0, ST_COUNT * sizeof(unsigned long long));
    if (e1 != cudaSuccess) rc = fail(nullptr, FS_E_CUDA, "fs_create: %s", cudaGetErrorString(e1));
  }
  if (rc != FS_OK) {
    delete ctx;
    return rc;
  }
  *out = ctx;
  return FS_OK;
}

int fs_create_multi(fs_ctx** out, const int* device_ids, int n_devices, int flags) {
  if (!out) return fail(nullptr, FS_E_INVALID, "fs_create_multi: null out pointer");
  *out = nullptr;
  if (!device_ids || n_devices < 1 || n_devices > FS2_MAX_DEVICES)
    return fail(nullptr, FS_E_INVALID, "fs_create_multi: 1..%d devices", FS2_MAX_DEVICES);
  for (int i = 0; i < n_devices; i++)
    for (int j = 0; j < i; j++)
      if (device_ids[i] == device_ids[j])
        return fail(nullptr, FS_E_INVALID, "fs_create_multi: device %d listed twice", device_ids[i]);
  fs_ctx* lead = nullptr;
  int rc = fs_create(&lead, device_ids[0], flags);
  if (rc) return rc;
  for (int i = 1; i < n_devices; i++) {
    fs_ctx* peer = nullptr;
    if ((rc = fs_create(&peer, device_ids[i], flags))) {
      fs_destroy(lead);
      return rc;
    }
    peer->lead = lead;
    lead->peers.push_back(peer);
    // the combine kernel on the lead reads the peer's film over NVLink when a peer mapping
    // exists; otherwise the film is staged through cudaMemcpyPeerAsync
    int can = 0;
    cudaSetDevice(lead->device);
    if (cudaDeviceCanAccessPeer(&can, lead->device, peer->device) == cudaSuccess && can) {
      cudaError_t e = cudaDeviceEnablePeerAccess(peer->device, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        e = cudaSuccess;
      }
      can = e == cudaSuccess;
      if (!can) cudaGetLastError();
    }
    lead->peer_direct.push_back(can ? 1 : 0);
  }
  *out = lead;
  return FS_OK;
}

int fs_device_count(const fs_ctx* ctx) { return ctx ? 1 + (int)ctx->peers.size() : 0; }

int fs_set_tile_shard(fs_ctx* ctx, int rank, int world) {
  if (!ctx || world < 1 || rank < 0 || rank >= world || world > 65535)
    return fail(ctx, FS_E_INVALID, "fs_set_tile_shard: bad rank %d of %d", rank, world);
  if (!ctx->peers.empty())
    return fail(ctx, FS_E_INVALID, "fs_set_tile_shard: a multi-device context shards by itself");
  ctx->tile_n = (uint32_t)world;
  ctx->tile_r = (uint32_t)rank;
  return FS_OK;
}

int fs_destroy(fs_ctx* ctx) {
  if (!ctx) return FS_OK;
  for (fs_ctx* p : ctx->peers) {
    p->lead = nullptr;
    fs_destroy(p);
  }
  ctx->peers.clear();
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  DeviceBuf* all[] = {&ctx->sref_sphere, &ctx->nodes, &ctx->scene_refs, &ctx->tri_plane, &ctx->tri_aux, &ctx->faces,
                      &ctx->objects, &ctx->materials, &ctx->textures, &ctx->texels, &ctx->lut,
                      &ctx->lb_nodes, &ctx->lb_leaf, &ctx->lb_plane, &ctx->lb_aux, &ctx->lb_order,
                      &ctx->sum, &ctx->aux, &ctx->mat_id, &ctx->mean, &ctx->count, &ctx->rgb8,
                      &ctx->scratch, &ctx->auxpack, &ctx->post, &ctx->wf_planes, &ctx->wf_hit, &ctx->wf_counters,
                      &ctx->stage};
  for (DeviceBuf* b : all) dev_free(b);
  if (ctx->d_stats) cudaFree(ctx->d_stats);
  if (ctx->wf_host_live) {
    cudaFreeHost(ctx->wf_host_live);
    for (cudaEvent_t e : ctx->wf_events) cudaEventDestroy(e);
    for (int p = 1; p < FS2_WF_PIPES; p++) {
      if (ctx->wf_streams[p]) cudaStreamDestroy(ctx->wf_streams[p]);
      if (ctx->wf_join[p]) cudaEventDestroy(ctx->wf_join[p]);
    }
    if (ctx->wf_fork) cudaEventDestroy(ctx->wf_fork);
  }
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->evk0) cudaEventDestroy(ctx->evk0);
  if (ctx->evk1) cudaEventDestroy(ctx->evk1);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
  return FS_OK;
}

int fs_scene_upload(fs_ctx* ctx, const fs_scene_desc* s) {
  if (!ctx || !s) return fail(ctx, FS_E_INVALID, "fs_scene_upload: null argument");
  int rc = validate_scene(ctx, s);
  if (rc) return rc;
  // LUT repacked to 16 bytes per entry so one ld.global.v4 fetches a normal
  std::vector<float> lut4((size_t)s->n_lut * 4);
  for (uint32_t i = 0; i < s->n_lut; i++) {
    lut4[4 * (size_t)i + 0] = s->lut[3 * (size_t)i + 0];
    lut4[4 * (size_t)i + 1] = s->lut[3 * (size_t)i + 1];
    lut4[4 * (size_t)i + 2] = s->lut[3 * (size_t)i + 2];
    lut4[4 * (size_t)i + 3] = 0.f;
  }
  // leaf BVH over every mesh-tree leaf (fs2_lbvh_build.h): built once, uploaded to every device
  LbvhHost lb;
  build_lbvh(s, &lb, 9, 0, 0.0f);
  if (ctx->peers.empty()) return scene_upload_one(ctx, s, lut4, lb);
  // multi-device lead: the scene is replicated, one host thread per device
  std::vector<fs_ctx*> all = all_contexts(ctx);
  std::vector<int> rcs(all.size(), FS_OK);
  std::vector<std::thread> workers;
  for (size_t i = 1; i < all.size(); i++)
    workers.emplace_back([&, i]() { rcs[i] = scene_upload_one(all[i], s, lut4, lb); });
  rcs[0] = scene_upload_one(ctx, s, lut4, lb);
  for (std::thread& t : workers) t.join();
  for (size_t i = 0; i < all.size(); i++)
    if (rcs[i]) return i ? fail(ctx, rcs[i], "device %d: %s", all[i]->device, all[i]->error) : rcs[i];
  return FS_OK;
}

int fs_film_resize(fs_ctx* ctx, int width, int height) {
  if (!ctx || width <= 0 || height <= 0 || (int64_t)width * height > (1ll << 30))
    return fail(ctx, FS_E_INVALID, "fs_film_resize: bad size %dx%d", width, height);
  for (fs_ctx* p : ctx->peers) {
    int prc = fs_film_resize(p, width, height);
    if (prc) return fail(ctx, prc, "device %d: %s", p->device, p->error);
  }
  int rc = set_device(ctx);
  if (rc) return rc;
  size_t n = (size_t)width * height;
  if ((rc = dev_alloc(ctx, &ctx->sum, n * 16))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->aux, n * 16))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->mat_id, n * 4))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->mean, n * 12))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->count, n * 4))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->rgb8, n * 3))) return rc;
  ctx->width = width;
  ctx->height = height;
  return fs_film_reset(ctx);
}

int fs_film_reset(fs_ctx* ctx) {
  if (!ctx) return fail(ctx, FS_E_INVALID, "null context");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_reset before fs_film_resize");
  for (fs_ctx* p : ctx->peers) {
    int prc = fs_film_reset(p);
    if (prc) return fail(ctx, prc, "device %d: %s", p->device, p->error);
  }
  int rc = set_device(ctx);
  if (rc) return rc;
  size_t n = (size_t)ctx->width * ctx->height;
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->sum.p, 0, n * 16, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->aux.p, 0, n * 16, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->mat_id.p, 0, n * 4, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->mean.p, 0, n * 12, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->count.p, 0, n * 4, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->rgb8.p, 0, n * 3, ctx->stream));
  return FS_OK;  // stream-ordered with the launches that follow
}

int fs_first_hit(fs_ctx* ctx, const fs_camera* cam, int width, int height, float* t,
                 uint32_t* object, uint32_t* face, uint32_t* mat_id, float* range) {
  if (!ctx || !cam || width <= 1 || height <= 1)
    return fail(ctx, FS_E_INVALID, "fs_first_hit: bad argument");
  if (!ctx->have_scene) return fail(ctx, FS_E_NOSCENE, "fs_first_hit before fs_scene_upload");
  int rc = set_device(ctx);
  if (rc) return rc;
  if ((rc = ensure_lbvh_for_camera(ctx, cam))) return rc;
  size_t n = (size_t)width * height;
  if ((rc = dev_alloc(ctx, &ctx->scratch, n * 20))) return rc;
  char* base = (char*)ctx->scratch.p;
  float* d_t = (float*)base;
  uint32_t* d_obj = (uint32_t*)(base + n * 4);
  uint32_t* d_face = (uint32_t*)(base + n * 8);
  uint32_t* d_mat = (uint32_t*)(base + n * 12);
  float* d_range = (float*)(base + n * 16);
  CamBasis cb = camera_basis(cam, width, height);
  k_first_hit<true><<<pixel_grid(width, height), 128, 0, ctx->stream>>>(ctx->ds, cb, width, height, d_t,
                                                                  d_obj, d_face, d_mat, d_range,
                                                                  ctx->d_stats);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  if (t) CUDA_TRY(ctx, cudaMemcpyAsync(t, d_t, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (object) CUDA_TRY(ctx, cudaMemcpyAsync(object, d_obj, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (face) CUDA_TRY(ctx, cudaMemcpyAsync(face, d_face, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (mat_id) CUDA_TRY(ctx, cudaMemcpyAsync(mat_id, d_mat, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (range) CUDA_TRY(ctx, cudaMemcpyAsync(range, d_range, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

int fs_trace_range(fs_ctx* ctx, const fs_camera* cam, int width, int height, float x, float y,
                   float* range) {
  // TraceRange(viewer, scene, frame, x, y): fractional pixel coordinates are expressed as a
  // 2x2 query whose pixel (0,0) carries the requested offset: (x/(W-1))*2-1 only depends on
  // x and W, so a single-pixel launch with the real width/height is used instead.
  if (!ctx || !cam || !range || width <= 1 || height <= 1)
    return fail(ctx, FS_E_INVALID, "fs_trace_range: bad argument");
  if (!ctx->have_scene) return fail(ctx, FS_E_NOSCENE, "fs_trace_range before fs_scene_upload");
  if (!(x >= 0.0f && y >= 0.0f && x <= (float)(width - 1) && y <= (float)(height - 1)))
    return fail(ctx, FS_E_INVALID, "fs_trace_range: (%g,%g) outside the image", x, y);
  int rc = set_device(ctx);
  if (rc) return rc;
  if ((rc = ensure_lbvh_for_camera(ctx, cam))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->scratch, 64))) return rc;
  CamBasis cb = camera_basis(cam, width, height);
  // shift the projection origin so that pixel (0,0) of a 1x1 launch is (x, y)
  float xd = (float)((double)cb.half_w * ((double)(x / cb.wm1) * 2.0 - 1.0));
  float yd = (float)((double)cb.half_h * ((double)(y / cb.hm1) * 2.0 - 1.0));
  CamBasis one = cb;
  // pixel (0,0) maps to xd' = half_w * (0*2 - 1) = -half_w; choose half extents so that
  // -half_w' == xd, -half_h' == yd (exact: negation only).
  one.half_w = -xd;
  one.half_h = -yd;
  float* d_range = (float*)ctx->scratch.p;
  k_first_hit<true><<<dim3(1, 1), 128, 0, ctx->stream>>>(ctx->ds, one, 1, 1, nullptr, nullptr, nullptr,
                                                  nullptr, d_range, ctx->d_stats);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaMemcpyAsync(range, d_range, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

int fs_trace_rays(fs_ctx* ctx, const float* segments, int n_rays, fs_ray_hit* out) {
  if (!ctx || n_rays < 0 || (n_rays && (!segments || !out)))
    return fail(ctx, FS_E_INVALID, "fs_trace_rays: bad argument");
  if (!ctx->have_scene) return fail(ctx, FS_E_NOSCENE, "fs_trace_rays before fs_scene_upload");
  if (!n_rays) return FS_OK;
  int rc = set_device(ctx);
  if (rc) return rc;
  // origins beyond the leaf BVHs' pad range take the linear leaf scans on the device
  // (DScene::lb_origin_limit), so no rebuild is needed here
  const size_t in_b = (size_t)n_rays * 24, out_b = (size_t)n_rays * sizeof(fs_ray_hit);
  if ((rc = dev_alloc(ctx, &ctx->scratch, in_b + out_b + 64))) return rc;
  float* d_seg = (float*)ctx->scratch.p;
  fs_ray_hit* d_out = (fs_ray_hit*)((char*)ctx->scratch.p + ((in_b + 63) & ~(size_t)63));
  CUDA_TRY(ctx, cudaMemcpyAsync(d_seg, segments, in_b, cudaMemcpyHostToDevice, ctx->stream));
  k_probe_rays<<<(n_rays + FS2_LB_THREADS - 1) / FS2_LB_THREADS, FS2_LB_THREADS, 0, ctx->stream>>>(
      ctx->ds, d_seg, n_rays, d_out);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaMemcpyAsync(out, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

// ---- fs_trace_rays_static: the same query with its scratch in the library image (no allocation).
// Layout of the array: FS_TRACE_RAYS_STATIC_MAX segments (24 B each), then as many fs_ray_hit
// records (52 B each); both parts start on a 256-byte boundary.
#define FS2_STATIC_SEG_BYTES ((size_t)FS_TRACE_RAYS_STATIC_MAX * 24)
#define FS2_STATIC_HIT_BYTES ((size_t)FS_TRACE_RAYS_STATIC_MAX * sizeof(fs_ray_hit))
static_assert(FS2_STATIC_SEG_BYTES % 256 == 0, "the hit records must start on a 256-byte boundary");
static __device__ __align__(256) unsigned char g_static_ray_scratch[FS2_STATIC_SEG_BYTES + FS2_STATIC_HIT_BYTES];

namespace {
// In host code `g_static_ray_scratch` is a host-side stand-in, not a device address: a kernel given
// it faults (or, where the GPU can reach host memory, silently uses the wrong memory).  Every
// device has its own copy of the array, so its address is looked up once per device, with that
// device current, and kept here.
#define FS2_STATIC_DEVICES 64
std::mutex g_static_lock;
void* g_static_ptr[FS2_STATIC_DEVICES] = {};

int static_ray_scratch(fs_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= FS2_STATIC_DEVICES)
    return fail(ctx, FS_E_INVALID, "fs_trace_rays_static: device ordinal %d beyond %d", ctx->device,
                FS2_STATIC_DEVICES - 1);
  std::lock_guard<std::mutex> hold(g_static_lock);
  if (!g_static_ptr[ctx->device]) {
    int rc = set_device(ctx);  // the lookup answers for the current device
    if (rc) return rc;
    void* p = nullptr;
    CUDA_TRY(ctx, cudaGetSymbolAddress(&p, g_static_ray_scratch));
    if (!p) return fail(ctx, FS_E_CUDA, "fs_trace_rays_static: cudaGetSymbolAddress returned a null address");
    g_static_ptr[ctx->device] = p;
  }
  *out = (unsigned char*)g_static_ptr[ctx->device];
  return FS_OK;
}
}  // namespace

int fs_trace_rays_static(fs_ctx* ctx, const float* segments, int n_rays, fs_ray_hit* out) {
  if (!ctx || n_rays < 0 || (n_rays && (!segments || !out)))
    return fail(ctx, FS_E_INVALID, "fs_trace_rays_static: bad argument");
  if (n_rays > FS_TRACE_RAYS_STATIC_MAX)
    return fail(ctx, FS_E_INVALID,
                "fs_trace_rays_static: %d rays exceed the fixed scratch of FS_TRACE_RAYS_STATIC_MAX = %d "
                "(split the batch or call fs_trace_rays)", n_rays, FS_TRACE_RAYS_STATIC_MAX);
  if (!ctx->have_scene) return fail(ctx, FS_E_NOSCENE, "fs_trace_rays_static before fs_scene_upload");
  if (!n_rays) return FS_OK;
  int rc = set_device(ctx);
  if (rc) return rc;
  unsigned char* base = nullptr;
  if ((rc = static_ray_scratch(ctx, &base))) return rc;
  // Nothing in the array has to be cleared: the n_rays segments the kernel reads are overwritten
  // here, and it writes every one of the n_rays records that are copied back.
  float* d_seg = (float*)base;
  fs_ray_hit* d_out = (fs_ray_hit*)(base + FS2_STATIC_SEG_BYTES);
  const size_t in_b = (size_t)n_rays * 24, out_b = (size_t)n_rays * sizeof(fs_ray_hit);
  CUDA_TRY(ctx, cudaMemcpyAsync(d_seg, segments, in_b, cudaMemcpyHostToDevice, ctx->stream));
  k_probe_rays<<<(n_rays + FS2_LB_THREADS - 1) / FS2_LB_THREADS, FS2_LB_THREADS, 0, ctx->stream>>>(
      ctx->ds, d_seg, n_rays, d_out);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaMemcpyAsync(out, d_out, out_b, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

int fs_sample_sky(fs_ctx* ctx, const float* views, int n, float* rgb) {
  if (!ctx || n < 0 || (n && (!views || !rgb))) return fail(ctx, FS_E_INVALID, "fs_sample_sky: bad argument");
  if (!ctx->have_scene) return fail(ctx, FS_E_NOSCENE, "fs_sample_sky before fs_scene_upload");
  if (!n) return FS_OK;
  int rc = set_device(ctx);
  if (rc) return rc;
  const size_t b = (size_t)n * 12;
  if ((rc = dev_alloc(ctx, &ctx->scratch, 2 * b + 64))) return rc;
  float* d_in = (float*)ctx->scratch.p;
  float* d_out = (float*)((char*)ctx->scratch.p + ((b + 63) & ~(size_t)63));
  CUDA_TRY(ctx, cudaMemcpyAsync(d_in, views, b, cudaMemcpyHostToDevice, ctx->stream));
  k_probe_sky<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->ds, d_in, n, d_out);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaMemcpyAsync(rgb, d_out, b, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

int fs_film_resolve(fs_ctx* ctx) {
  if (!ctx) return fail(ctx, FS_E_INVALID, "null context");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_resolve before fs_film_resize");
  int rc = set_device(ctx);
  if (rc) return rc;
  int n = ctx->width * ctx->height;
  k_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>((const float4*)ctx->sum.p, n,
                                                      (float*)ctx->mean.p, (uint32_t*)ctx->count.p,
                                                      (uint8_t*)ctx->rgb8.p);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

// First half of fs_trace on one device: the path-tracing kernels of this context's shard.
static int trace_begin(fs_ctx* ctx, const fs_camera* cam, uint32_t first_sample, uint32_t n_samples,
                       uint32_t sample_stride, uint64_t seed, int flags) {
  if (!ctx || !cam) return fail(ctx, FS_E_INVALID, "fs_trace: null argument");
  if (!ctx->have_scene) return fail(ctx, FS_E_NOSCENE, "fs_trace before fs_scene_upload");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_trace before fs_film_resize");
  if (ctx->width < 2 || ctx->height < 2)
    return fail(ctx, FS_E_INVALID, "fs_trace: film must be at least 2x2");
  if (sample_stride == 0) sample_stride = 1;
  int rc = set_device(ctx);
  if (rc) return rc;
  if ((rc = ensure_lbvh_for_camera(ctx, cam))) return rc;
  CamBasis cb = camera_basis(cam, ctx->width, ctx->height);
  Film film{(float4*)ctx->sum.p, (float4*)ctx->aux.p, (uint32_t*)ctx->mat_id.p, ctx->width,
            ctx->height};
  TraceParams tp{first_sample, n_samples, sample_stride, seed, flags, ctx->tune};
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  CUDA_TRY(ctx, cudaEventRecord(ctx->evk0, ctx->stream));
  if (n_samples) {
    dim3 grid = pixel_grid(ctx->width, ctx->height);
    const bool per_thread = (flags & FS_TRACE_PER_THREAD) != 0;
    const bool coop = (flags & FS_TRACE_COOP) != 0;
    const bool st = ctx->stats_build;
#define FS2_LAUNCH(K) K<<<grid, 128, 0, ctx->stream>>>(ctx->ds, cb, film, tp, ctx->d_stats)
    const int mega = FS_TRACE_PER_THREAD | FS_TRACE_COOP | FS_TRACE_LBVH_LANE | FS_TRACE_MEGAKERNEL;
    if (!(flags & mega)) {
      if ((rc = launch_wavefront(ctx, cb, film, tp))) return rc;
      ctx->launches--;  // the common increment below counts one launch
    } else if (per_thread) {
      if (st) FS2_LAUNCH((k_trace_mega<true, false>)); else FS2_LAUNCH((k_trace_mega<false, false>));
    } else if (coop) {
      if (st) FS2_LAUNCH(k_trace_coop<true>); else FS2_LAUNCH(k_trace_coop<false>);
    } else if (flags & FS_TRACE_LBVH_LANE) {
      if (st) FS2_LAUNCH((k_trace_mega<true, true>)); else FS2_LAUNCH((k_trace_mega<false, true>));
    } else {
      if (st)
        k_trace_phase<true><<<grid, FS2_PH_THREADS, ph_smem_bytes(ctx->ds.scene_root >= 0), ctx->stream>>>(
            ctx->ds, cb, film, tp, ctx->d_stats);
      else
        k_trace_phase<false><<<grid, FS2_PH_THREADS, ph_smem_bytes(ctx->ds.scene_root >= 0), ctx->stream>>>(
            ctx->ds, cb, film, tp, ctx->d_stats);
    }
#undef FS2_LAUNCH
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
  }
  CUDA_TRY(ctx, cudaEventRecord(ctx->evk1, ctx->stream));
  return FS_OK;
}

// Second half: resolve (mean, gamma, RGB8), timing events, synchronisation.
static int trace_finish(fs_ctx* ctx, int flags) {
  if (!(flags & FS_TRACE_NO_RESOLVE)) {
    int n = ctx->width * ctx->height;
    k_resolve<<<(n + 255) / 256, 256, 0, ctx->stream>>>(
        (const float4*)ctx->sum.p, n, (float*)ctx->mean.p, (uint32_t*)ctx->count.p,
        (uint8_t*)ctx->rgb8.p);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
  }
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  ctx->timing_pending = true;
  if (flags & FS_TRACE_ASYNC) return FS_OK;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return collect_timing(ctx);
}

// Multi-device lead: shard, trace on every device (one host thread each, like the reference fans
// TraceScene out over its worker threads, engine.cpp:261-274), combine the films on the lead.
//  * sample sharding: device r takes the requested samples j with j == (n - 1 - r) mod N, so the
//    lead owns the last sample (whose depth-0 hit fills the aux planes, frame.cpp:78-85);
//  * tile sharding: every device traces all samples of the 16x8-pixel tiles (tx, ty) with
//    (tx + ty) mod N == r (diagonal interleave).  Chosen when there are fewer samples than
//    devices, or on request; the combine then moves one film in total instead of N - 1.
static int trace_multi(fs_ctx* lead, const fs_camera* cam, uint32_t first_sample, uint32_t n_samples,
                       uint32_t sample_stride, uint64_t seed, int flags) {
  if (!lead->have_scene) return fail(lead, FS_E_NOSCENE, "fs_trace before fs_scene_upload");
  if (!lead->width) return fail(lead, FS_E_NOFILM, "fs_trace before fs_film_resize");
  std::vector<fs_ctx*> all = all_contexts(lead);
  const uint32_t N = (uint32_t)all.size();
  const int mega = FS_TRACE_PER_THREAD | FS_TRACE_COOP | FS_TRACE_LBVH_LANE | FS_TRACE_MEGAKERNEL;
  bool tiles = (flags & FS_TRACE_SHARD_TILES) || (!(flags & FS_TRACE_SHARD_SAMPLES) && n_samples < N);
  if (tiles && (flags & mega)) {
    if (flags & FS_TRACE_SHARD_TILES)
      return fail(lead, FS_E_INVALID, "fs_trace: tile sharding needs the default (wavefront) pipeline");
    tiles = false;
  }
  if (sample_stride == 0) sample_stride = 1;
  const size_t n_px = (size_t)lead->width * lead->height;
  const int inner = (flags | FS_TRACE_NO_RESOLVE) & ~(FS_TRACE_SHARD_TILES | FS_TRACE_SHARD_SAMPLES);
  std::vector<int> rcs(N, FS_OK);
  auto run = [&](uint32_t r) {
    fs_ctx* c = all[r];
    uint32_t first = first_sample, count = n_samples, stride = sample_stride;
    if (tiles) {
      c->tile_n = N;
      c->tile_r = r;
    } else {
      c->tile_n = 1;
      c->tile_r = 0;
      // samples jr, jr + N, ... below n_samples; devices beyond the sample count get none
      const uint32_t jr = (uint32_t)(((uint64_t)n_samples + N - 1 - r) % N);  // (n - 1 - r) mod N
      count = n_samples > jr ? (n_samples - jr + N - 1) / N : 0;
      first = first_sample + jr * sample_stride;
      stride = sample_stride * N;
    }
    int rc = set_device(c);
    if (!rc && r) {
      // a peer's film holds this call's contribution only
      if (cudaMemsetAsync(c->sum.p, 0, n_px * 16, c->stream) != cudaSuccess)
        rc = fail(c, FS_E_CUDA, "cudaMemsetAsync(peer film)");
    }
    if (!rc) rc = trace_begin(c, cam, first, count, stride, seed, r ? (inner & ~FS_TRACE_ASYNC) : inner);
    if (!rc && r && cudaStreamSynchronize(c->stream) != cudaSuccess)
      rc = fail(c, FS_E_CUDA, "cudaStreamSynchronize(peer): %s", cudaGetErrorString(cudaGetLastError()));
    c->tile_n = 1;
    c->tile_r = 0;
    rcs[r] = rc;
  };
  std::vector<std::thread> workers;
  for (uint32_t r = 1; r < N; r++) workers.emplace_back(run, r);
  run(0);
  for (std::thread& t : workers) t.join();
  int rc = set_device(lead);
  for (uint32_t r = 0; r < N && !rc; r++)
    if (rcs[r]) rc = r ? fail(lead, rcs[r], "device %d: %s", all[r]->device, all[r]->error) : rcs[r];
  if (rc) {
    cudaStreamSynchronize(lead->stream);
    cudaGetLastError();
    return rc;
  }
  // combine on the lead, in device order
  PeerFilms pf;
  memset(&pf, 0, sizeof pf);
  pf.n_peers = (int)N - 1;
  pf.tile_n = tiles ? (int)N : 0;
  const size_t per_peer = n_px * (tiles ? 36 : 16);
  bool need_stage = false;
  for (uint32_t r = 1; r < N; r++) need_stage |= !lead->peer_direct[r - 1];
  if (need_stage && (rc = dev_alloc(lead, &lead->stage, per_peer * (N - 1)))) return rc;
  for (uint32_t r = 1; r < N; r++) {
    fs_ctx* c = all[r];
    if (lead->peer_direct[r - 1]) {
      pf.sum[r - 1] = (const float4*)c->sum.p;
      pf.aux[r - 1] = (const float4*)c->aux.p;
      pf.mat_id[r - 1] = (const uint32_t*)c->mat_id.p;
    } else {
      char* base = (char*)lead->stage.p + per_peer * (r - 1);
      CUDA_TRY(lead, cudaMemcpyPeerAsync(base, lead->device, c->sum.p, c->device, n_px * 16, lead->stream));
      pf.sum[r - 1] = (const float4*)base;
      if (tiles) {
        CUDA_TRY(lead, cudaMemcpyPeerAsync(base + n_px * 16, lead->device, c->aux.p, c->device, n_px * 16, lead->stream));
        CUDA_TRY(lead, cudaMemcpyPeerAsync(base + n_px * 32, lead->device, c->mat_id.p, c->device, n_px * 4, lead->stream));
        pf.aux[r - 1] = (const float4*)(base + n_px * 16);
        pf.mat_id[r - 1] = (const uint32_t*)(base + n_px * 32);
      }
    }
  }
  k_film_combine<<<(unsigned)((n_px + 255) / 256), 256, 0, lead->stream>>>(
      (float4*)lead->sum.p, (float4*)lead->aux.p, (uint32_t*)lead->mat_id.p, pf, lead->width, (int)n_px);
  lead->launches++;
  CUDA_TRY(lead, cudaGetLastError());
  // the peers' films may be cleared by the next call only after the combine has read them
  if ((rc = trace_finish(lead, flags & ~FS_TRACE_ASYNC))) return rc;
  return FS_OK;
}

int fs_trace(fs_ctx* ctx, const fs_camera* cam, uint32_t first_sample, uint32_t n_samples,
             uint32_t sample_stride, uint64_t seed, int flags) {
  if (!ctx || !cam) return fail(ctx, FS_E_INVALID, "fs_trace: null argument");
  if (!ctx->peers.empty()) return trace_multi(ctx, cam, first_sample, n_samples, sample_stride, seed, flags);
  int rc = trace_begin(ctx, cam, first_sample, n_samples, sample_stride, seed, flags);
  if (rc) return rc;
  return trace_finish(ctx, flags);
}

int fs_film_read(fs_ctx* ctx, float* mean_rgb, uint32_t* count, uint8_t* rgb8, float* normal,
                 float* depth, uint64_t* mat_id) {
  if (!ctx) return fail(ctx, FS_E_INVALID, "null context");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_read before fs_film_resize");
  int rc = set_device(ctx);
  if (rc) return rc;
  size_t n = (size_t)ctx->width * ctx->height;
  if (mean_rgb) CUDA_TRY(ctx, cudaMemcpyAsync(mean_rgb, ctx->mean.p, n * 12, cudaMemcpyDeviceToHost, ctx->stream));
  if (count) CUDA_TRY(ctx, cudaMemcpyAsync(count, ctx->count.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (rgb8) CUDA_TRY(ctx, cudaMemcpyAsync(rgb8, ctx->rgb8.p, n * 3, cudaMemcpyDeviceToHost, ctx->stream));
  if (normal || depth || mat_id) {
    // split the aux planes on the device into the caller's layouts, then copy straight into
    // the caller's buffers (full-speed DMA when they are page-locked, see fs_host_register)
    if ((rc = dev_alloc(ctx, &ctx->auxpack, n * 24))) return rc;
    unsigned long long* d_ids = (unsigned long long*)ctx->auxpack.p;
    float* d_normal = (float*)(d_ids + n);
    float* d_depth = d_normal + 3 * n;
    k_unpack_aux<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(
        (const float4*)ctx->aux.p, (const uint32_t*)ctx->mat_id.p, (int)n, d_normal, d_depth, d_ids);
    ctx->launches++;
    if (normal) CUDA_TRY(ctx, cudaMemcpyAsync(normal, d_normal, n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    if (depth) CUDA_TRY(ctx, cudaMemcpyAsync(depth, d_depth, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (mat_id) CUDA_TRY(ctx, cudaMemcpyAsync(mat_id, d_ids, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

int fs_film_post(fs_ctx* ctx, const fs_post_params* pr, float* rgb, uint8_t* rgb8) {
  if (!ctx || !pr) return fail(ctx, FS_E_INVALID, "fs_film_post: null argument");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_post before fs_film_resize");
  if (pr->denoise_passes < 0 || pr->denoise_passes > 8 || pr->flare_ghosts < 0 ||
      pr->flare_ghosts > 16)
    return fail(ctx, FS_E_INVALID, "fs_film_post: denoise_passes in 0..8, flare_ghosts in 0..16");
  const bool glow = pr->bloom_strength > 0.0f || (pr->flare_strength > 0.0f && pr->flare_ghosts > 0);
  if (glow && (pr->bloom_radius < 1 || pr->bloom_radius > 128))
    return fail(ctx, FS_E_INVALID, "fs_film_post: bloom_radius in 1..128");
  if (pr->denoise_passes && !(pr->sigma_depth > 0.0f && pr->sigma_colour >= 0.0f && pr->sigma_normal >= 0.0f))
    return fail(ctx, FS_E_INVALID, "fs_film_post: sigma_depth > 0, sigma_colour >= 0, sigma_normal >= 0");
  int rc = set_device(ctx);
  if (rc) return rc;
  const int w = ctx->width, h = ctx->height, n = w * h;
  // planes: A, B, C (float4 each), sigma (float), out rgb (3 floats), out rgb8 (3 bytes)
  if ((rc = dev_alloc(ctx, &ctx->post, (size_t)n * (48 + 4 + 12 + 3) + 64))) return rc;
  float4* A = (float4*)ctx->post.p;
  float4 *B = A + n, *C = B + n;
  float* sigma = (float*)(C + n);
  float* d_rgb = sigma + n;
  uint8_t* d_rgb8 = (uint8_t*)(d_rgb + 3 * (size_t)n);
  const float4* aux = (const float4*)ctx->aux.p;
  const uint32_t* ids = (const uint32_t*)ctx->mat_id.p;
  PostParams pp{pr->sigma_normal, pr->sigma_depth, pr->sigma_colour, pr->bloom_threshold,
                pr->bloom_strength, pr->bloom_radius, pr->flare_strength, pr->flare_ghosts};
  const dim3 g2((w + 31) / 32, (h + 7) / 8);
  const int g1 = (n + 255) / 256;
  cudaStream_t st = ctx->stream;
  k_post_prep<<<g1, 256, 0, st>>>((const float*)ctx->mean.p, n, A);
  ctx->launches++;
  float4 *cur = A, *other = B;
  if (pr->denoise_passes) {
    k_post_sigma<<<g2, 256, 0, st>>>(cur, ids, w, h, sigma);
    ctx->launches++;
    float scale = 1.0f;
    for (int i = 0; i < pr->denoise_passes; i++, scale *= 0.5f) {
      k_post_atrous<<<g2, 256, 0, st>>>(cur, aux, ids, sigma, w, h, 1 << i, scale, pp, other);
      ctx->launches++;
      float4* t = cur;
      cur = other;
      other = t;
    }
  }
  const float4* glow_plane = nullptr;
  if (glow) {
    k_post_bright<<<g1, 256, 0, st>>>(cur, n, pr->bloom_threshold, other);
    k_post_blur<<<g2, 256, 0, st>>>(other, w, h, pr->bloom_radius, 0, C);
    k_post_blur<<<g2, 256, 0, st>>>(C, w, h, pr->bloom_radius, 1, other);
    ctx->launches += 3;
    glow_plane = other;
  }
  k_post_compose<<<g2, 256, 0, st>>>(cur, glow_plane, w, h, pp, d_rgb, d_rgb8);
  ctx->launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  if (rgb) CUDA_TRY(ctx, cudaMemcpyAsync(rgb, d_rgb, (size_t)n * 12, cudaMemcpyDeviceToHost, st));
  if (rgb8) CUDA_TRY(ctx, cudaMemcpyAsync(rgb8, d_rgb8, (size_t)n * 3, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(ctx, cudaStreamSynchronize(st));
  return FS_OK;
}

int fs_device_selftest(int device, uint32_t* words, uint32_t n_words) {
  if (!words || n_words < FS2_SELFTEST_WORDS)
    return fail(nullptr, FS_E_INVALID, "fs_device_selftest: need room for %d words", FS2_SELFTEST_WORDS);
  cudaError_t e = cudaSetDevice(device);
  void* buf = nullptr;
  if (e == cudaSuccess) e = cudaMalloc(&buf, 1024);
  if (e == cudaSuccess) {
    float4* plane = (float4*)buf;
    fs_tri_aux* aux = (fs_tri_aux*)((char*)buf + 64);
    uint32_t* out = (uint32_t*)((char*)buf + 256);
    k_selftest_pack<<<1, 1>>>(plane, aux);
    k_selftest<<<1, 1>>>(plane, aux, out);
    e = cudaGetLastError();
    if (e == cudaSuccess)
      e = cudaMemcpy(words, out, FS2_SELFTEST_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost);
  }
  if (buf) cudaFree(buf);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(nullptr, FS_E_CUDA, "fs_device_selftest: %s", cudaGetErrorString(e));
  }
  return FS_OK;
}

int fs_host_register(void* p, size_t bytes) {
  if (!p || !bytes) return fail(nullptr, FS_E_INVALID, "fs_host_register: null buffer");
  cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(nullptr, FS_E_CUDA, "cudaHostRegister: %s", cudaGetErrorString(e));
  }
  return FS_OK;
}

int fs_host_unregister(void* p) {
  if (!p) return FS_OK;
  cudaError_t e = cudaHostUnregister(p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(nullptr, FS_E_CUDA, "cudaHostUnregister: %s", cudaGetErrorString(e));
  }
  return FS_OK;
}

int fs_film_device(fs_ctx* ctx, void** sum_rgba, size_t* n_floats) {
  if (!ctx || !sum_rgba || !n_floats) return fail(ctx, FS_E_INVALID, "null argument");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_device before fs_film_resize");
  *sum_rgba = ctx->sum.p;
  *n_floats = (size_t)ctx->width * ctx->height * 4;
  return FS_OK;
}

int fs_film_device_aux(fs_ctx* ctx, void** aux_xyzd, void** mat_id_u32, size_t* n_pixels) {
  if (!ctx || !aux_xyzd || !mat_id_u32 || !n_pixels) return fail(ctx, FS_E_INVALID, "null argument");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_device_aux before fs_film_resize");
  *aux_xyzd = ctx->aux.p;
  *mat_id_u32 = ctx->mat_id.p;
  *n_pixels = (size_t)ctx->width * ctx->height;
  return FS_OK;
}

int fs_film_save(fs_ctx* ctx, float* sum_rgba) {
  if (!ctx || !sum_rgba) return fail(ctx, FS_E_INVALID, "null argument");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_save before fs_film_resize");
  int rc = set_device(ctx);
  if (rc) return rc;
  size_t n = (size_t)ctx->width * ctx->height;
  CUDA_TRY(ctx, cudaMemcpyAsync(sum_rgba, ctx->sum.p, n * 16, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return FS_OK;
}

int fs_film_load(fs_ctx* ctx, const float* sum_rgba) {
  if (!ctx || !sum_rgba) return fail(ctx, FS_E_INVALID, "null argument");
  if (!ctx->width) return fail(ctx, FS_E_NOFILM, "fs_film_load before fs_film_resize");
  int rc = set_device(ctx);
  if (rc) return rc;
  size_t n = (size_t)ctx->width * ctx->height;
  // the sample count rides in the float w channel: exact below 2^24 samples per pixel only
  for (size_t i = 0; i < n; i++) {
    float c = sum_rgba[4 * i + 3];
    if (!(c >= 0.0f && c < 16777216.0f) || c != floorf(c))
      return fail(ctx, FS_E_INVALID, "fs_film_load: pixel %zu has sample count %g (must be an integer below 2^24)", i, (double)c);
  }
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->sum.p, sum_rgba, n * 16, cudaMemcpyHostToDevice, ctx->stream));
  return fs_film_resolve(ctx);
}

int fs_get_stats(fs_ctx* ctx, fs_stats* out) {
  if (!ctx || !out) return fail(ctx, FS_E_INVALID, "null argument");
  int rc = set_device(ctx);
  if (rc) return rc;
  unsigned long long h[ST_COUNT];
  CUDA_TRY(ctx, cudaMemcpyAsync(h, ctx->d_stats, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if ((rc = collect_timing(ctx))) return rc;
  memset(out, 0, sizeof *out);
  out->n_paths = h[ST_PATHS];
  out->n_segments = h[ST_SEGMENTS];
  out->n_node = h[ST_NODE];
  out->n_tri = h[ST_TRI];
  out->n_bary = h[ST_BARY];
  out->n_prim = h[ST_PRIM];
  out->n_shade = h[ST_SHADE];
  out->n_lbnode = h[ST_LBNODE];
  out->n_tritest = h[ST_TRITEST];
  out->launches = ctx->launches;
  out->last_trace_ms = ctx->last_trace_ms;
  out->last_kernel_ms = ctx->last_kernel_ms;
  for (fs_ctx* p : ctx->peers) {  // multi-device lead: work counters of the whole job
    fs_stats ps;
    if ((rc = fs_get_stats(p, &ps))) return fail(ctx, rc, "device %d: %s", p->device, p->error);
    out->n_paths += ps.n_paths;
    out->n_segments += ps.n_segments;
    out->n_node += ps.n_node;
    out->n_tri += ps.n_tri;
    out->n_bary += ps.n_bary;
    out->n_prim += ps.n_prim;
    out->n_shade += ps.n_shade;
    out->n_lbnode += ps.n_lbnode;
    out->n_tritest += ps.n_tritest;
    out->launches += ps.launches;
  }
  return set_device(ctx);
}

int fs_reset_stats(fs_ctx* ctx) {
  if (!ctx) return fail(ctx, FS_E_INVALID, "null context");
  for (fs_ctx* p : ctx->peers) fs_reset_stats(p);
  int rc = set_device(ctx);
  if (rc) return rc;
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_stats, 0, ST_COUNT * sizeof(unsigned long long), ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->launches = 0;
  return FS_OK;
}

int fs_set_stream(fs_ctx* ctx, void* stream) {
  if (!ctx) return fail(ctx, FS_E_INVALID, "null context");
  int rc = set_device(ctx);
  if (rc) return rc;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if ((rc = collect_timing(ctx))) return rc;
  ctx->stream = stream ? (cudaStream_t)stream : ctx->own_stream;
  return FS_OK;
}

int fs_get_stream(fs_ctx* ctx, void** stream) {
  if (!ctx || !stream) return fail(ctx, FS_E_INVALID, "null argument");
  *stream = (void*)ctx->stream;
  return FS_OK;
}

int fs_mesh_tree_build(int device, const float* verts, uint32_t n_verts, const uint32_t* face_vertices,
                       const float* face_planes, uint32_t n_faces, uint32_t face_base,
                       uint32_t node_base, uint32_t tri_base, fs_mesh_tree* out) {
  if (!out) return fail(nullptr, FS_E_INVALID, "fs_mesh_tree_build: null out");
  memset(out, 0, sizeof *out);
  if (!verts || !face_vertices || !face_planes || !n_verts || !n_faces)
    return fail(nullptr, FS_E_INVALID, "fs_mesh_tree_build: empty mesh");
  for (size_t i = 0; i < 3 * (size_t)n_faces; i++)
    if (face_vertices[i] >= n_verts)
      return fail(nullptr, FS_E_INVALID, "fs_mesh_tree_build: vertex index out of range");
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail(nullptr, FS_E_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  int rc = bd_build(verts, n_verts, face_vertices, face_planes, n_faces, face_base, node_base, tri_base, out);
  if (rc != FS_OK) fs_mesh_tree_free(out);
  return rc;
}

void fs_mesh_tree_free(fs_mesh_tree* t) {
  if (!t) return;
  free(t->nodes);
  free(t->tri_plane);
  free(t->tri_aux);
  memset(t, 0, sizeof *t);
}

}  // extern "C"
