/*
 * fs2_cabi.h — C ABI of the B200-native per-sample path-tracing hot path of
 * Final Stage 2.0 (libfs2cuda.so).
 *
 * Plain pointers and sizes only; no C++/torch types cross this boundary.  The
 * reference has no FFI layer: its seam is the pair of free functions
 *   void    base::TraceScene(const Camera&, Scene*, DisplayFrame*, ImagePlaneCache*)
 *   float32 base::TraceRange(const Camera&, Scene*, DisplayFrame*, float32 x, float32 y)
 * (reference source/engine.h:68-75).  Everything under those two calls —
 * camera ray generation (source/engine.cpp:166-251), the TraceStep recursion
 * (source/engine.cpp:59-159), Scene::Trace and the two octrees
 * (source/scene.cpp:76-103,227-249, source/bvh.h:154-231, source/mesh.cpp:78-115,
 * 277-322), the primitive intersectors (source/object.cpp:23-237), materials
 * (source/material.cpp:28-339) and DisplayFrame::WritePixel (source/frame.cpp:53-85)
 * — runs behind the entry points declared here.
 *
 * The host (C++ mirror of the reference classes, or any other language through
 * FFI) flattens its scene into the `fs_scene_desc` arrays below, uploads them
 * once, and then calls fs_trace()/fs_first_hit().  All inputs are copied during
 * the call; the caller may free them afterwards.  Every function returns 0 on
 * success or a negative FS_E_* code; fs_last_error() gives the message.
 * Nothing throws across the boundary.  A context is not re-entrant.
 */
#ifndef FS2_CABI_H_
#define FS2_CABI_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FS2_ABI_VERSION 1

/* ---- error codes -------------------------------------------------------- */
#define FS_OK 0
#define FS_E_INVALID (-1)   /* bad argument / malformed scene description   */
#define FS_E_CUDA (-2)      /* a CUDA runtime call failed                   */
#define FS_E_NOSCENE (-3)   /* trace requested before fs_scene_upload       */
#define FS_E_NOFILM (-4)    /* film requested before fs_film_resize         */
#define FS_E_NOMEM (-5)
#define FS_E_IO (-6)

/* ---- object / material tags -------------------------------------------- */
/* object types (reference source/object.h:63-152, source/mesh.h:124-149) */
#define FS_OBJ_SPHERE 0
#define FS_OBJ_PLANE 1
#define FS_OBJ_DISC 2
#define FS_OBJ_CUBOID 3
#define FS_OBJ_QUAD 4
#define FS_OBJ_MESH 5

/* material types (reference source/material.h:88-328) */
#define FS_MAT_DIFFUSE 0
#define FS_MAT_LIGHT 1
#define FS_MAT_METAL 2
#define FS_MAT_MIRROR 3
#define FS_MAT_GLASS 4
#define FS_MAT_LIQUID 5
#define FS_MAT_CERAMIC 6
#define FS_MAT_GLOW 7
#define FS_MAT_FOG 8

#define FS_NODE_INTERNAL 0xFFFFFFFFu /* fs_node.count value marking an internal node */
#define FS_NO_HIT_PARAM 2.0f         /* reference source/math/trace.cpp:18-21 */
#define FS_CREATE_STATS 1            /* fs_create flag: count node/triangle/primitive tests */
#define FS_LUT_SIZE 32768            /* reference source/material.cpp:17 */

/*
 * Octree node, 32 bytes, used by both the scene tree (leaf refs = object
 * indices) and the per-mesh trees (leaf refs = packed triangle records).
 * Internal nodes always own 8 consecutive children starting at `first`
 * (reference source/bvh.h:108-137: child i = x | z<<1 | y<<2).  The three
 * split planes are the axis planes through (bmin+bmax)*0.5f and are derived
 * on the fly.
 */
typedef struct fs_node {
  float bmin[3];
  uint32_t first; /* internal: index of child 0; leaf: offset of first ref */
  float bmax[3];
  uint32_t count; /* FS_NODE_INTERNAL, or number of refs in the leaf      */
} fs_node;

/*
 * Packed triangle record, one per (leaf, face) reference, in leaf order.
 * Split in two arrays so that the plane stage of the reference's triangle test
 * (source/math/intersect.cpp:382-388) streams 16 B per candidate:
 *   tri_plane[i] = face plane (source/mesh.cpp:137-146)
 *   tri_aux[i]   = everything the barycentric stage needs, with the
 *                  per-triangle constants of source/math/intersect.cpp:96-129,
 *                  168-173 precomputed by the same float operations.
 */
typedef struct fs_tri_aux {
  float p1[3];       /* first vertex                                              */
  uint32_t face;     /* bits 0..29 global face index, bits 30..31 dropped axis    */
  float v0a, v0b;    /* the two kept components of (p2 - p1)                      */
  float v1a, v1b;    /* the two kept components of (p3 - p1)                      */
  float dot00, dot01, dot11, inv_denom;
} fs_tri_aux;        /* 48 bytes */

/* Per-face shading attributes, gathered only for the winning face
 * (reference source/mesh.cpp:289-311). */
typedef struct fs_face_attr {
  float n[3][3]; /* vertex normals (already approx-normalised at load)  */
  float uv[3][2];
  float pad;
} fs_face_attr; /* 64 bytes */

/*
 * Object record, 128 bytes.  f[] by type:
 *  SPHERE  f[0..2] centre, f[3] radius                       (object.cpp:23-36)
 *  PLANE   f[0..3] plane                                     (object.cpp:59-73)
 *  DISC    f[0..3] plane, f[4..6] origin, f[7] radius        (object.cpp:97-113)
 *  CUBOID  f[4*i..4*i+3] face plane i, i<6                   (object.cpp:126-165)
 *  QUAD    f[0..3] plane, f[4..6] origin, f[7..9] tangent,
 *          f[10..12] bitangent, f[13] half_width, f[14] half_height (object.cpp:213-237)
 *  MESH    no floats; u[0] root node, u[1] first face_attr, u[2] flags
 *          (bit0 has normals, bit1 has texcoords), u[3] face count (mesh.cpp:277-322)
 */
typedef struct fs_object {
  uint32_t type;
  uint32_t material; /* index into materials[] */
  uint32_t u[4];
  float f[24];
  float bmin[3]; /* Object::GetBounds(): tree build input; for a cuboid it must contain the solid
                    (its eight vertices) — the device skips the six-plane test for segments that
                    stay outside the sphere through the box's corners */
  float bmax[3];
} fs_object; /* 24 + 96 + 24 = 144 bytes */

/* Material record (reference source/material.h:54-328), 64 bytes. */
typedef struct fs_material {
  uint32_t type;
  uint32_t id;        /* Material::GetID() truncated to 32 bit (material.cpp:26) */
  float diffuse[3];   /* diffuse_                                           */
  float emissive[3];  /* LightMaterial::emissive_ or GlowMaterial::glow_    */
  float roughness;    /* Metal roughness_ / Ceramic+Glow shininess_         */
  float index;        /* Glass/Liquid index_                                */
  float reflectivity; /* Glass/Liquid reflectivity_                         */
  float frost;        /* Glass frost_                                       */
  float density;      /* Fog density_ (already ×1000, material.cpp:306)     */
  float tex_scale;    /* texture_scale_                                     */
  int32_t texture;    /* index into textures[] or -1                        */
  uint32_t pad;
} fs_material;

typedef struct fs_texture {
  uint32_t width, height;
  uint64_t offset; /* float offset of texel (0,0) in texels[] (RGB float)  */
} fs_texture;

typedef struct fs_scene_desc {
  uint32_t abi_version; /* FS2_ABI_VERSION */
  uint32_t n_nodes;
  uint32_t n_scene_refs; /* object indices referenced by scene-tree leaves */
  uint32_t n_tris;       /* packed triangle records (leaf references)      */
  uint32_t n_faces;      /* fs_face_attr records                           */
  uint32_t n_objects;
  uint32_t n_materials;
  uint32_t n_textures;
  uint64_t n_texels;  /* number of floats in texels[]                       */
  int32_t scene_root; /* node index of the scene tree root, or -1: brute force
                         over objects[] in order (scene.cpp:230-233)        */
  uint32_t sky_material;
  uint32_t n_lut; /* FS_LUT_SIZE */
  uint32_t pad;
  const fs_node* nodes;
  const uint32_t* scene_refs;
  const float* tri_plane; /* 4 floats per record */
  const fs_tri_aux* tri_aux;
  const fs_face_attr* faces;
  const fs_object* objects;
  const fs_material* materials;
  const fs_texture* textures;
  const float* texels;
  const float* lut; /* 3 floats per entry (normal.cpp:9-27) */
} fs_scene_desc;

/* Camera POD (reference source/camera.h:43-63). */
typedef struct fs_camera {
  float origin[3];
  float target[3];
  float fov_y;
  float aperture_size;
  float focal_depth;
  float z_near;
  float z_far;
  int32_t fast_render_enabled;
} fs_camera;

/* fs_trace flags */
#define FS_TRACE_DEFAULT 0     /* wavefront pipeline: path pool in HBM, traversal and shading kernels */
#define FS_TRACE_NO_JITTER 1  /* pixel-centre rays (TraceRange ray construction) */
#define FS_TRACE_NO_DOF 2     /* ignore aperture_size                            */
#define FS_TRACE_WAVEFRONT 4  /* the default, named (kept for callers that set it) */
#define FS_TRACE_NO_RESOLVE 8 /* skip the mean/gamma resolve (multi-GPU: resolve after the reduce) */
#define FS_TRACE_ASYNC 16     /* skip the final synchronisation: the film is complete in stream order.
                                 (The default pipeline's host loop follows the waves through an event
                                 ring, so the call still returns only when the last wave has been
                                 enqueued; the megakernel variants return at once.) */
#define FS_TRACE_PER_THREAD 32 /* megakernel; every thread scans its own mesh leaves linearly, in reference order */
#define FS_TRACE_COOP 64       /* megakernel; linear leaf scans done by the whole warp, 32 triangles per step */
#define FS_TRACE_LBVH_LANE 128 /* megakernel; per-leaf 4-wide BVH walked by free-running lanes (nested loops) */
#define FS_TRACE_MEGAKERNEL 256 /* megakernel; warp-synchronous phase machine over the leaf BVH (one launch,
                                 * a pixel's samples are summed in order: bit-reproducible films) */
#define FS_TRACE_SHARD_TILES 512    /* multi-device context: interleaved 16x8-pixel tiles per device */
#define FS_TRACE_SHARD_SAMPLES 1024 /* multi-device context: sample indices per device (the default
                                     * whenever there are at least as many samples as devices) */
/* default (none of the four megakernel flags): the wavefront pipeline.  Its traversal kernel is
 * the same phase machine over the per-leaf 4-wide BVH built by fs_scene_upload; every path has
 * the RNG stream and arithmetic of the megakernels, but a pixel's samples reach the film through
 * float atomics, so films agree with the megakernels to summation order, not bit for bit. */

typedef struct fs_stats {
  uint64_t n_paths;    /* camera paths traced since the last reset           */
  uint64_t n_segments; /* ray segments = Σ TraceResult::ray_count (engine.cpp:72) */
  uint64_t n_node;     /* ray_intersect_bounds calls  (stats build only)     */
  uint64_t n_tri;      /* ray_intersect_triangle calls (stats build only)    */
  uint64_t n_bary;     /* triangle tests reaching the barycentric stage      */
  uint64_t n_prim;     /* primitive (non-mesh object) tests                  */
  uint64_t n_shade;    /* material evaluations + sky lookups                 */
  uint64_t launches;   /* kernels launched by this context since the last reset */
  float last_trace_ms;  /* device time of the last fs_trace, all its kernels (CUDA events) */
  float last_kernel_ms; /* device time of the path-tracing kernel(s) alone in that call   */
  /* n_node/n_tri/n_bary above count reference-equivalent work (what the reference's loops
   * execute for the same rays).  The default kernel's own work (stats build only): */
  uint64_t n_lbnode;  /* leaf-BVH nodes visited (128 B each)                               */
  uint64_t n_tritest; /* triangle plane tests actually executed                            */
} fs_stats;

typedef struct fs_ctx fs_ctx;

/* Lifetime.  `device` is a CUDA ordinal; the context uses that device's
 * primary context, so pointers interoperate with other runtime users. */
int fs_create(fs_ctx** out, int device, int flags);
/* One context over several devices of the box: what the reference's thread fan-out over every
 * core (source/engine.cpp:261-274) becomes on an 8-GPU machine.  The scene is replicated by
 * fs_scene_upload, fs_trace shards the requested samples (or, with fewer samples than devices,
 * interleaved pixel tiles) across the devices — one host thread each — and combines the films on
 * device_ids[0] with a kernel that reads the other devices' films over NVLink peer mappings, in
 * device order (deterministic); every other call acts on the film of device_ids[0].  Per-sample
 * random streams depend only on (seed, pixel, sample index), so the result is the single-device
 * sample set.  Every entry point below accepts such a context. */
int fs_create_multi(fs_ctx** out, const int* device_ids, int n_devices, int flags);
int fs_device_count(const fs_ctx* ctx); /* 1 for fs_create contexts */
int fs_destroy(fs_ctx* ctx);
const char* fs_last_error(const fs_ctx* ctx); /* ctx may be NULL: global last error */
int fs_abi_version(void);

/* Scene: replaces walking Scene::object_list_/object_tree_ (scene.h:150-156). */
int fs_scene_upload(fs_ctx* ctx, const fs_scene_desc* scene);

/* Film: replaces DisplayFrame ctor / Reset (frame.cpp:7-51). */
int fs_film_resize(fs_ctx* ctx, int width, int height);
int fs_film_reset(fs_ctx* ctx);

/* Replaces TraceRange (engine.cpp:298-329) for every pixel at once: pixel
 * centre rays, no jitter, no DOF.  Outputs are host arrays of width*height
 * elements, row 0 = bottom scanline; any may be NULL.
 *   t      : hit parameter along the camera segment, FS_NO_HIT_PARAM on miss
 *   object : index into objects[], 0xFFFFFFFF on miss
 *   face   : global face index for mesh hits, cuboid plane for cuboids, else 0xFFFFFFFF
 *   mat_id : material id of the hit (sky material id on miss)
 *   range  : TraceRange's return value                                        */
int fs_first_hit(fs_ctx* ctx, const fs_camera* cam, int width, int height,
                 float* t, uint32_t* object, uint32_t* face, uint32_t* mat_id,
                 float* range);

/* Single-ray form of the above with fractional pixel coordinates, exactly
 * TraceRange(viewer, scene, frame, x, y). */
int fs_trace_range(fs_ctx* ctx, const fs_camera* cam, int width, int height,
                   float x, float y, float* range);

/* One ray query as Scene::Trace leaves its ObjectCollision (reference source/object.h:47-61,
 * source/scene.cpp:227-249: nearest hit, then the back-face flip of the normal). */
typedef struct fs_ray_hit {
  float param;          /* portion of the segment, FS_NO_HIT_PARAM when nothing was hit */
  float point[3];
  float normal[3];      /* flipped towards the ray origin when struck from behind */
  float uv[2];
  uint32_t object;      /* index into objects[], 0xFFFFFFFF = none */
  uint32_t face;        /* mesh face / cuboid plane, 0xFFFFFFFF otherwise */
  uint32_t material;    /* index into materials[] (sky_material when nothing was hit) */
  uint32_t is_internal; /* 1 when the origin lies behind the surface (scene.cpp:244-247) */
} fs_ray_hit;

/* Replaces Scene::Trace(const ray&, ObjectCollision*) (reference source/scene.h:128,
 * source/scene.cpp:227-249) for n_rays segments at once: segments[6*i..] = start xyz, stop xyz.
 * Host pointers. */
int fs_trace_rays(fs_ctx* ctx, const float* segments, int n_rays, fs_ray_hit* out);

/* fs_trace_rays for callers that must not allocate device memory (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as fs_trace_rays, bit for bit,
 * but the staged segments and hit records live in a fixed array in device global memory that is
 * part of the library image (FS_TRACE_RAYS_STATIC_MAX rays = 2 490 368 bytes, held by every GPU
 * that loads the library until the process exits); the call makes no cudaMalloc / cudaFree and
 * uses no allocating library routine.
 *   - n_rays > FS_TRACE_RAYS_STATIC_MAX is refused with FS_E_INVALID and a message naming the limit
 *     (split the batch, or use fs_trace_rays).
 *   - Every call on a device shares that one array; it keeps the previous call's contents.  Nothing
 *     in it has to start at zero: a call overwrites the n_rays segments it reads and the kernel
 *     writes all n_rays records it returns.
 *   - Calls that overlap on the GPU race (two contexts on one device, or two host threads): callers
 *     serialise them.  The call itself is synchronous like fs_trace_rays. */
#define FS_TRACE_RAYS_STATIC_MAX 32768
int fs_trace_rays_static(fs_ctx* ctx, const float* segments, int n_rays, fs_ray_hit* out);

/* Replaces Scene::SampleSky(depth, view) (reference source/scene.h:131, source/scene.cpp:157-162)
 * for n view directions (3 floats each); rgb gets 3 floats per query.  The reference's `depth`
 * argument does not reach LightMaterial::Sample's result and is not part of the call. */
int fs_sample_sky(fs_ctx* ctx, const float* views, int n, float* rgb);

/* Replaces n_samples calls of TraceScene (engine.cpp:253-296): accumulates
 * samples first_sample, first_sample+stride, ... (n_samples of them) of every
 * pixel into the device film.  Per-sample random streams depend only on
 * (seed, pixel, sample index), so sharding sample indices across GPUs
 * reproduces the single-GPU sample set. */
int fs_trace(fs_ctx* ctx, const fs_camera* cam, uint32_t first_sample,
             uint32_t n_samples, uint32_t sample_stride, uint64_t seed, int flags);

/* Film read-back into host buffers (any may be NULL), replaces reading
 * DisplayFrame's buffers (frame.h:79-91).  mean_rgb: 3 floats/pixel,
 * rgb8: 3 bytes/pixel (frame.cpp:60-69), normal: 3 floats, mat_id: u64. */
int fs_film_read(fs_ctx* ctx, float* mean_rgb, uint32_t* count, uint8_t* rgb8,
                 float* normal, float* depth, uint64_t* mat_id);

/* Device views for a multi-process reduce (one process per GPU, e.g. NCCL): sum = float4 per
 * pixel (rgb sum, w = sample count as float).  After an external reduce call fs_film_resolve
 * to refresh mean/rgb8. */
int fs_film_device(fs_ctx* ctx, void** sum_rgba, size_t* n_floats);
/* ... and the aux planes (float4 normal xyz + depth, u32 material id per pixel): with tile
 * sharding every pixel is non-zero on exactly one rank, so summing them gathers them. */
int fs_film_device_aux(fs_ctx* ctx, void** aux_xyzd, void** mat_id_u32, size_t* n_pixels);
/* Multi-process tile sharding: this context renders only the 16x8-pixel tiles (tx, ty) with
 * (tx + ty) mod world == rank (all samples of them); world = 1 restores the whole image.
 * Default pipeline only. */
int fs_set_tile_shard(fs_ctx* ctx, int rank, int world);

/* Film checkpoint / resume: copies the accumulation film (4 floats per pixel: rgb sum, sample
 * count) to / from a host buffer of width*height*4 floats.  After fs_film_load the resolved
 * buffers are current and the next fs_trace continues at first_sample = the loaded count. */
int fs_film_save(fs_ctx* ctx, float* sum_rgba);
int fs_film_load(fs_ctx* ctx, const float* sum_rgba);
int fs_film_resolve(fs_ctx* ctx);

/*
 * Film post-processing on the device.  The reference's README describes
 * variance-adaptive denoising, bloom and flare; its code implements none of them but keeps the
 * guide buffers for it (normal / depth / material id, reference source/frame.h:86-91).  This
 * call filters the resolved mean image with those guides; the film itself is not modified, so
 * accumulation can continue.  Stages (each off when its count / strength is 0):
 *   denoise : `denoise_passes` edge-avoiding a-trous passes (stride 1, 2, 4, ...), weights from
 *             normal^sigma_normal, relative depth / sigma_depth, luminance difference in units of
 *             sigma_colour local standard deviations; taps of another material id are rejected
 *   bloom   : luminance above bloom_threshold, Gaussian of bloom_radius pixels, x bloom_strength
 *   flare   : flare_ghosts mirrored copies of the blurred bright pass through the image centre
 * Multi-ray anti-aliasing (k jittered rays per pixel per call) is fs_trace's n_samples.
 * Outputs are host arrays (either may be NULL): rgb = 3 floats / pixel, rgb8 = display bytes
 * with DisplayFrame's gamma (frame.cpp:60-69).
 */
typedef struct fs_post_params {
  int32_t denoise_passes; /* 0 = off, at most 8 */
  float sigma_normal;     /* e.g. 64   */
  float sigma_depth;      /* e.g. 0.02 */
  float sigma_colour;     /* e.g. 4    */
  float bloom_threshold;  /* e.g. 1.0  */
  float bloom_strength;   /* 0 = off   */
  int32_t bloom_radius;   /* pixels, 1..128 */
  float flare_strength;   /* 0 = off   */
  int32_t flare_ghosts;   /* 0..16     */
} fs_post_params;
int fs_film_post(fs_ctx* ctx, const fs_post_params* params, float* rgb, uint8_t* rgb8);

/* Diagnostic: one device thread evaluates the known-answer inputs of the reference (fast
 * inverse square root, plane / triangle / sphere / box intersections of two probe rays,
 * normalize / reflect / refract / rotate, sphere-map coordinates, the xorshift64* stream from the
 * reference's default seed, WritePixel's quantiser) with the device functions the kernels are
 * built from and returns the raw 32-bit words (48 of them; layout in csrc/fs2_selftest.cuh).
 * tests/test_gpu_kat.py compares them with the bits the unmodified reference prints. */
int fs_device_selftest(int device, uint32_t* words, uint32_t n_words);

int fs_get_stats(fs_ctx* ctx, fs_stats* out);
int fs_reset_stats(fs_ctx* ctx);

/* Page-locks / releases a caller-owned host buffer (cudaHostRegister) so that fs_film_read and
 * fs_first_hit copy into it by DMA at full PCIe speed; optional, any host pointer works. */
int fs_host_register(void* buffer, size_t bytes);
int fs_host_unregister(void* buffer);

/* CUDA stream the context launches on (cudaStream_t as void*).  fs_set_stream makes the
 * context enqueue on a caller-owned stream (e.g. the one an NCCL reduce of the film runs
 * on); NULL restores the context's own stream. */
int fs_get_stream(fs_ctx* ctx, void** stream);
int fs_set_stream(fs_ctx* ctx, void* stream);

/*
 * GPU build of one mesh octree: replaces MeshBvh::BuildBvh / MeshBvhNode::Subdivide
 * (reference source/mesh.cpp:39-76,117-152; SAT test source/math/intersect.cpp:193-311) and
 * the packing of the leaf records.  Same leaf membership, leaf order, node numbering and
 * record bytes as the host builder (host/fs2_octree.cpp build_mesh_tree).
 *   verts          3 floats per vertex
 *   face_vertices  3 vertex indices per face
 *   face_planes    4 floats per face (MeshFace::plane, never all zero)
 *   face_base      global index of this mesh's face 0 (fs_tri_aux::face)
 *   node_base      index the returned nodes[0] will have in fs_scene_desc::nodes
 *   tri_base       index the returned record 0 will have in tri_plane / tri_aux
 * The arrays of `out` are malloc'ed; release them with fs_mesh_tree_free.
 */
typedef struct fs_mesh_tree {
  fs_node* nodes;
  float* tri_plane;
  fs_tri_aux* tri_aux;
  uint32_t n_nodes, n_tris;
  float build_ms; /* device time from the first to the last build kernel (CUDA events) */
} fs_mesh_tree;

int fs_mesh_tree_build(int device, const float* verts, uint32_t n_verts, const uint32_t* face_vertices,
                       const float* face_planes, uint32_t n_faces, uint32_t face_base,
                       uint32_t node_base, uint32_t tri_base, fs_mesh_tree* out);
void fs_mesh_tree_free(fs_mesh_tree* tree);

#ifdef __cplusplus
}
#endif
#endif /* FS2_CABI_H_ */
