// fs2_engine.cpp — DisplayFrame, TraceScene / TraceRange of the host mirror and the C
// surface used from Python.  This is where the reference's seam (source/engine.h:68-75)
// meets the C ABI: flatten-if-dirty -> fs_scene_upload -> fs_trace / fs_first_hit ->
// fs_film_read into the DisplayFrame's host buffers.  No tracing happens on the host.
#include <math.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <exception>
#include <memory>
#include <string>
#include <vector>

#include "fs2_impl.h"
#include "fs2_octree.h"

namespace {

struct TraceState {
  int device = 0;
  std::vector<int> devices;  // SetTraceDevices: more than one -> multi-device context
  bool devices_changed = false;
  int shard_flags = 0;       // FS_TRACE_SHARD_TILES / FS_TRACE_SHARD_SAMPLES / 0 = automatic
  uint64_t seed = 0x46533230ull;  // "FS20"
  fs_ctx* ctx = nullptr;
  int ctx_device = -1;
  const void* scene = nullptr;  // SceneImpl whose flattening is resident on the device
  uint64_t scene_version = 0;
  const void* frame = nullptr;  // FrameImpl that owns the device film
  uint32_t film_w = 0, film_h = 0;
  // identity of what the device film accumulates (checkpoint header): the last scene / camera
  // traced into it, and — after LoadFilmCheckpoint — what the next TraceSamples must match
  uint64_t scene_hash = 0, camera_hash = 0;
  bool expect_identity = false;
  uint64_t expect_seed = 0, expect_scene = 0, expect_camera = 0;
  char error[512] = "";
};

// FNV-1a over a byte range (identity check, not a security boundary)
uint64_t fnv(const void* p, size_t n, uint64_t h = 0xcbf29ce484222325ull) {
  const unsigned char* b = (const unsigned char*)p;
  for (size_t i = 0; i < n; i++) h = (h ^ b[i]) * 0x100000001b3ull;
  return h;
}
// Small arrays in full, the big ones by their size and both ends (a render is not resumed on
// "almost the same" mesh by accident; hashing a gigabyte on every upload is not worth more).
uint64_t hash_range(const void* p, size_t bytes, uint64_t h) {
  const size_t edge = 1 << 16;
  h = fnv(&bytes, sizeof bytes, h);
  if (!p || !bytes) return h;
  if (bytes <= 2 * edge) return fnv(p, bytes, h);
  h = fnv(p, edge, h);
  return fnv((const char*)p + bytes - edge, edge, h);
}
uint64_t hash_scene(const fs_scene_desc* d) {
  uint64_t h = fnv(d, offsetof(fs_scene_desc, nodes));
  h = hash_range(d->nodes, (size_t)d->n_nodes * sizeof(fs_node), h);
  h = hash_range(d->scene_refs, (size_t)d->n_scene_refs * 4, h);
  h = hash_range(d->tri_plane, (size_t)d->n_tris * 16, h);
  h = hash_range(d->faces, (size_t)d->n_faces * sizeof(fs_face_attr), h);
  h = hash_range(d->objects, (size_t)d->n_objects * sizeof(fs_object), h);
  h = hash_range(d->materials, (size_t)d->n_materials * sizeof(fs_material), h);
  h = hash_range(d->textures, (size_t)d->n_textures * sizeof(fs_texture), h);
  return hash_range(d->texels, (size_t)d->n_texels * 4, h);
}

TraceState& state() {
  static TraceState s;
  return s;
}

int fail(const char* what, fs_ctx* ctx) {
  TraceState& st = state();
  snprintf(st.error, sizeof st.error, "%s: %s", what, fs_last_error(ctx));
  fprintf(stderr, "fs2: %s\n", st.error);
  return -1;
}

void fail_msg(const char* msg) {
  TraceState& st = state();
  snprintf(st.error, sizeof st.error, "%s", msg);
  fprintf(stderr, "fs2: %s\n", st.error);
}

void fetch_aux(fs2h::FrameImpl* f);

int ensure_context() {
  TraceState& st = state();
  if (st.ctx && st.ctx_device == st.device && !st.devices_changed) return 0;
  if (st.ctx && st.frame) fetch_aux((fs2h::FrameImpl*)const_cast<void*>(st.frame));
  if (st.ctx) fs_destroy(st.ctx);
  st.ctx = nullptr;
  st.scene = nullptr;
  st.frame = nullptr;
  st.film_w = st.film_h = 0;
  st.devices_changed = false;
  if (st.devices.size() > 1) {
    // every GPU of the box behind the one TraceScene call, as the reference puts every core
    // behind it (engine.cpp:261-274)
    if (fs_create_multi(&st.ctx, st.devices.data(), (int)st.devices.size(), 0) != FS_OK)
      return fail("fs_create_multi", nullptr);
  } else if (fs_create(&st.ctx, st.device, 0) != FS_OK) {
    return fail("fs_create", nullptr);
  }
  st.ctx_device = st.device;
  return 0;
}

int ensure_scene(base::Scene* scene) {
  TraceState& st = state();
  fs2h::SceneImpl* si = scene->impl();
  const fs_scene_desc* d = scene->Flatten();
  if (st.scene == si && st.scene_version == si->flat_version) return 0;
  if (fs_scene_upload(st.ctx, d) != FS_OK) return fail("fs_scene_upload", st.ctx);
  st.scene = si;
  st.scene_version = si->flat_version;
  st.scene_hash = hash_scene(d);
  return 0;
}

// The guide planes (normal, depth, material id: private in the reference, frame.h:86-91) are
// fetched from the device when somebody reads them, not after every TraceScene: they are 50 of
// the 89 MB a 1080p frame would otherwise move per call.
void fetch_aux(fs2h::FrameImpl* f) {
  TraceState& st = state();
  if (!f || !f->aux_stale) return;
  f->aux_stale = false;
  if (!st.ctx || st.frame != f) return;
  if (fs_film_read(st.ctx, nullptr, nullptr, nullptr, f->normal.data(), f->depth.data(), f->mat_id.data()) != FS_OK)
    fail("fs_film_read", st.ctx);
}

fs_camera to_pod(const base::Camera& c) {
  fs_camera o;
  memset(&o, 0, sizeof o);
  o.origin[0] = c.origin.x;
  o.origin[1] = c.origin.y;
  o.origin[2] = c.origin.z;
  o.target[0] = c.target.x;
  o.target[1] = c.target.y;
  o.target[2] = c.target.z;
  o.fov_y = c.fov_y;
  o.aperture_size = c.aperture_size;
  o.focal_depth = c.focal_depth;
  o.z_near = c.z_near;
  o.z_far = c.z_far;
  o.fast_render_enabled = c.fast_render_enabled ? 1 : 0;
  return o;
}

}  // namespace

namespace base {

using namespace fs2h;

// ---------------------------------------------------------------- DisplayFrame
DisplayFrame::DisplayFrame(uint32 width, uint32 height) : impl_(new FrameImpl()) {
  FrameImpl& f = *impl_;
  f.width = width;
  f.height = height;
  size_t n = (size_t)width * height;
  f.mean.assign(n * 3, 0.f);
  f.normal.assign(n * 3, 0.f);
  f.depth.assign(n, 0.f);
  f.count.assign(n, 0);
  f.rgb8.assign(n * 3, 0);
  f.mat_id.assign(n, 0);
  // page-lock the film so fs_film_read lands in it by DMA; without a device this fails and the
  // buffers simply stay pageable
  f.pinned = n > 0 && fs_host_register(f.mean.data(), n * 12) == FS_OK;
  if (f.pinned) {
    fs_host_register(f.normal.data(), n * 12);
    fs_host_register(f.depth.data(), n * 4);
    fs_host_register(f.count.data(), n * 4);
    fs_host_register(f.rgb8.data(), n * 3);
    fs_host_register(f.mat_id.data(), n * 8);
  }
}
DisplayFrame::~DisplayFrame() {
  if (state().frame == impl_.get()) state().frame = nullptr;
  FrameImpl& f = *impl_;
  if (f.pinned) {
    fs_host_unregister(f.mean.data());
    fs_host_unregister(f.normal.data());
    fs_host_unregister(f.depth.data());
    fs_host_unregister(f.count.data());
    fs_host_unregister(f.rgb8.data());
    fs_host_unregister(f.mat_id.data());
  }
}
// Reset() clears 43 bytes per pixel on the host (frame.cpp:43-51) that the next TraceScene
// overwrites anyway (5 ms at 1080p, more than a sixth of an 8-GPU frame): the fill is deferred
// until somebody looks at the buffers before they have been rewritten.
static void zero_picture(FrameImpl& f) {
  std::fill(f.mean.begin(), f.mean.end(), 0.f);
  std::fill(f.count.begin(), f.count.end(), 0u);
  std::fill(f.rgb8.begin(), f.rgb8.end(), (uint8_t)0);
}
static void zero_guides(FrameImpl& f) {
  std::fill(f.normal.begin(), f.normal.end(), 0.f);
  std::fill(f.depth.begin(), f.depth.end(), 0.f);
  std::fill(f.mat_id.begin(), f.mat_id.end(), (uint64_t)0);
}
static void materialise(FrameImpl& f) {
  if (!f.zero_pending) return;
  f.zero_pending = false;
  zero_picture(f);
  zero_guides(f);
}
void DisplayFrame::Reset() {  // frame.cpp:43-51
  FrameImpl& f = *impl_;
  f.zero_pending = true;
  f.samples = 0;
  f.device_dirty = true;
  f.aux_stale = false;
  f.host_written = false;
  if (state().frame == impl_.get()) state().expect_identity = false;
}
// frame.cpp:53-76, same operation order (running mean in float, quantiser in double)
void DisplayFrame::WritePixel(const vector3& pixel, uint32 x, uint32 y) {
  FrameImpl& f = *impl_;
  if (x >= f.width || y >= f.height) return;
  materialise(f);
  const size_t i = (size_t)y * f.width + x;
  float* m = &f.mean[3 * i];
  const float n0 = (float)f.count[i];
  const float n1 = (float)(++f.count[i]);
  const float in[3] = {pixel.x, pixel.y, pixel.z};
  for (int k = 0; k < 3; k++) {
    m[k] = (m[k] * n0 + in[k]) / n1;
    float c = m[k] < 0.0f ? 0.0f : m[k] > 1.0f ? 1.0f : m[k];
    f.rgb8[3 * i + k] = (uint8_t)(255.0 * pow((double)c, 1.0 / 2.2) + 0.5);
  }
  f.host_written = true;
}
void DisplayFrame::WritePixel(const TraceResult& result, uint32 x, uint32 y) {  // frame.cpp:78-85
  FrameImpl& f = *impl_;
  if (x >= f.width || y >= f.height) return;
  WritePixel(result.color, x, y);
  const size_t i = (size_t)y * f.width + x;
  f.normal[3 * i] = result.normal.x;
  f.normal[3 * i + 1] = result.normal.y;
  f.normal[3 * i + 2] = result.normal.z;
  f.depth[i] = result.depth;
  f.mat_id[i] = result.material_id;
}
uint8* DisplayFrame::GetDisplayBuffer() {
  materialise(*impl_);
  return impl_->rgb8.data();
}
uint32 DisplayFrame::GetWidth() { return impl_->width; }
uint32 DisplayFrame::GetHeight() { return impl_->height; }
float32 DisplayFrame::GetAspectRatio() { return (float32)impl_->width / impl_->height; }
void DisplayFrame::SetFrameCount(uint32 count) { impl_->frame_count = count; }
const float32* DisplayFrame::GetRenderTarget() {
  materialise(*impl_);
  return impl_->mean.data();
}
const uint32* DisplayFrame::GetCountBuffer() {
  materialise(*impl_);
  return impl_->count.data();
}
const float32* DisplayFrame::GetNormalBuffer() {
  materialise(*impl_);
  fetch_aux(impl_.get());
  return impl_->normal.data();
}
const float32* DisplayFrame::GetDepthBuffer() {
  materialise(*impl_);
  fetch_aux(impl_.get());
  return impl_->depth.data();
}
const uint64* DisplayFrame::GetMaterialIdBuffer() {
  materialise(*impl_);
  fetch_aux(impl_.get());
  return impl_->mat_id.data();
}

// ---------------------------------------------------------------- tracing
void SetTraceDevice(int device) {
  TraceState& st = state();
  if (st.devices.size() > 1) st.devices_changed = true;
  st.devices.clear();
  st.device = device;
}
void SetTraceDevices(const int* devices, int n_devices) {
  TraceState& st = state();
  std::vector<int> ids(devices, devices + (n_devices > 0 ? n_devices : 0));
  if (ids.empty()) ids.push_back(0);
  if (ids != st.devices) st.devices_changed = true;
  st.devices = ids;
  st.device = ids[0];
}
int GetTraceDeviceCount() { return state().ctx ? fs_device_count(state().ctx) : (int)std::max<size_t>(1, state().devices.size()); }
void SetTraceSharding(int mode) {
  state().shard_flags = mode == 1 ? FS_TRACE_SHARD_SAMPLES : mode == 2 ? FS_TRACE_SHARD_TILES : 0;
}
void SetTraceSeed(uint64 seed) { state().seed = seed; }
const char* LastTraceError() { return state().error; }

int TraceSamples(const Camera& view, Scene* scene, DisplayFrame* output, uint32 n_samples) {
  if (!scene || !output) return -1;
  TraceState& st = state();
  if (ensure_context() || ensure_scene(scene)) return -1;
  FrameImpl* f = output->impl();
  if (st.frame && st.frame != f) fetch_aux((FrameImpl*)const_cast<void*>(st.frame));  // the device film changes hands
  if (st.film_w != f->width || st.film_h != f->height) {
    if (fs_film_resize(st.ctx, (int)f->width, (int)f->height) != FS_OK)
      return fail("fs_film_resize", st.ctx);
    st.film_w = f->width;
    st.film_h = f->height;
    st.frame = f;
    f->device_dirty = false;
    f->samples = 0;
  } else if (st.frame != f || f->device_dirty) {
    // The device film accumulates for one DisplayFrame at a time.
    if (fs_film_reset(st.ctx) != FS_OK) return fail("fs_film_reset", st.ctx);
    st.frame = f;
    f->device_dirty = false;
    f->samples = 0;
  }
  if (f->host_written) {
    // DisplayFrame::WritePixel changed the host buffers: the device film continues from them
    // (sum = mean * count per pixel)
    std::vector<float> sums((size_t)f->width * f->height * 4);
    for (size_t i = 0; i < (size_t)f->width * f->height; i++) {
      const float n = (float)f->count[i];
      sums[4 * i] = f->mean[3 * i] * n;
      sums[4 * i + 1] = f->mean[3 * i + 1] * n;
      sums[4 * i + 2] = f->mean[3 * i + 2] * n;
      sums[4 * i + 3] = n;
    }
    if (fs_film_load(st.ctx, sums.data()) != FS_OK) return fail("fs_film_load", st.ctx);
    f->host_written = false;
  }
  fs_camera cam = to_pod(view);
  st.camera_hash = fnv(&cam, sizeof cam);
  if (st.expect_identity) {
    // a resumed film continues one image: same scene, same camera, same seed (the sample
    // sequence restarts at the stored count, so another seed would merely be unreproducible,
    // but another scene or camera blends two pictures)
    if (st.expect_scene != st.scene_hash || st.expect_camera != st.camera_hash || st.expect_seed != st.seed) {
      fail_msg("TraceSamples: the resumed checkpoint was written for another scene, camera or seed "
               "(Reset() the frame to start a new image)");
      return -1;
    }
    st.expect_identity = false;
  }
  if (fs_trace(st.ctx, &cam, f->samples, n_samples, 1, st.seed, FS_TRACE_DEFAULT | st.shard_flags) != FS_OK)
    return fail("fs_trace", st.ctx);
  f->samples += n_samples;
  f->frame_count += n_samples;
  if (fs_film_read(st.ctx, f->mean.data(), f->count.data(), f->rgb8.data(), nullptr, nullptr, nullptr) != FS_OK)
    return fail("fs_film_read", st.ctx);
  f->aux_stale = true;
  f->zero_pending = false;  // the picture was just rewritten, the guide planes come from the device
  return 0;
}

namespace {
struct CheckpointHeader {
  char magic[8];  // "FS2FILM2"
  uint32_t width, height, samples, reserved;
  uint64_t seed, scene_hash, camera_hash;  // what the sums were accumulated from
};
}  // namespace

bool PostProcessFrame(DisplayFrame* frame, const PostProcess& post, float* rgb, uint8* rgb8) {
  TraceState& st = state();
  if (!frame || !st.ctx || st.frame != frame->impl() || frame->impl()->device_dirty) {
    fail_msg("PostProcessFrame: the frame has no samples on the device");
    return false;
  }
  fs_post_params p = {post.denoise_passes, post.sigma_normal, post.sigma_depth, post.sigma_colour,
                      post.bloom_threshold, post.bloom_strength, post.bloom_radius,
                      post.flare_strength, post.flare_ghosts};
  if (fs_film_post(st.ctx, &p, rgb, rgb8) != FS_OK) {
    fail("fs_film_post", st.ctx);
    return false;
  }
  return true;
}

bool SaveFrameEXR(DisplayFrame* frame, const char* path) {
  if (!frame || !path) return false;
  return fs2h::save_exr(path, frame->GetWidth(), frame->GetHeight(), frame->GetRenderTarget(), true);
}

bool SaveFilmCheckpoint(DisplayFrame* frame, const char* path) {
  TraceState& st = state();
  if (!frame || !path || !st.ctx || st.frame != frame->impl() || frame->impl()->device_dirty) {
    fail_msg("SaveFilmCheckpoint: the frame has no samples on the device");
    return false;
  }
  FrameImpl* f = frame->impl();
  std::vector<float> sums((size_t)f->width * f->height * 4);
  if (fs_film_save(st.ctx, sums.data()) != FS_OK) {
    fail("fs_film_save", st.ctx);
    return false;
  }
  FILE* fp = fopen(path, "wb");
  if (!fp) return false;
  CheckpointHeader h = {{'F', 'S', '2', 'F', 'I', 'L', 'M', '2'}, f->width, f->height, f->samples, 0,
                        st.seed, st.scene_hash, st.camera_hash};
  bool ok = fwrite(&h, sizeof h, 1, fp) == 1 && fwrite(sums.data(), 4, sums.size(), fp) == sums.size();
  return fclose(fp) == 0 && ok;
}

bool LoadFilmCheckpoint(DisplayFrame* frame, const char* path) {
  if (!frame || !path) return false;
  TraceState& st = state();
  FrameImpl* f = frame->impl();
  FILE* fp = fopen(path, "rb");
  if (!fp) return false;
  CheckpointHeader h;
  std::vector<float> sums((size_t)f->width * f->height * 4);
  bool ok = fread(&h, sizeof h, 1, fp) == 1 && !memcmp(h.magic, "FS2FILM2", 8) &&
            h.width == f->width && h.height == f->height &&
            fread(sums.data(), 4, sums.size(), fp) == sums.size();
  fclose(fp);
  if (!ok) {
    fail_msg("LoadFilmCheckpoint: not a checkpoint of this frame size");
    return false;
  }
  if (ensure_context()) return false;
  if (st.film_w != f->width || st.film_h != f->height) {
    if (fs_film_resize(st.ctx, (int)f->width, (int)f->height) != FS_OK) {
      fail("fs_film_resize", st.ctx);
      return false;
    }
    st.film_w = f->width;
    st.film_h = f->height;
  }
  if (f->zero_pending) {  // the picture is rewritten below; the guide planes stay cleared
    zero_guides(*f);
    f->zero_pending = false;
  }
  if (fs_film_load(st.ctx, sums.data()) != FS_OK ||
      fs_film_read(st.ctx, f->mean.data(), f->count.data(), f->rgb8.data(), nullptr, nullptr, nullptr) != FS_OK) {
    fail("fs_film_load", st.ctx);
    return false;
  }
  st.frame = f;
  f->device_dirty = false;
  f->samples = h.samples;
  f->frame_count = h.samples;
  st.expect_identity = h.samples > 0;
  st.expect_seed = h.seed;
  st.expect_scene = h.scene_hash;
  st.expect_camera = h.camera_hash;
  return true;
}

// ---------------------------------------------------------------- single-ray queries
int Scene::TraceRays(const ray* trajectories, int n, ObjectCollision* hit_info) {
  if (n <= 0 || !trajectories || !hit_info) return 0;
  TraceState& st = state();
  if (ensure_context() || ensure_scene(this)) return -1;
  std::vector<float> seg(6 * (size_t)n);
  for (int i = 0; i < n; i++) {
    const ray& r = trajectories[i];
    float* g = &seg[6 * (size_t)i];
    g[0] = r.start.x, g[1] = r.start.y, g[2] = r.start.z;
    g[3] = r.stop.x, g[4] = r.stop.y, g[5] = r.stop.z;
  }
  std::vector<fs_ray_hit> hits((size_t)n);
  if (fs_trace_rays(st.ctx, seg.data(), n, hits.data()) != FS_OK) return fail("fs_trace_rays", st.ctx);
  const std::vector<Material*>& owner = impl_->flat_material_owner;
  int n_hit = 0;
  for (int i = 0; i < n; i++) {
    const fs_ray_hit& h = hits[(size_t)i];
    ObjectCollision& o = hit_info[i];
    if (h.object == 0xFFFFFFFFu) continue;  // the record stays as the caller initialised it
    n_hit++;
    o.param = h.param;
    o.point = vector3(h.point[0], h.point[1], h.point[2]);
    o.surface_normal = vector3(h.normal[0], h.normal[1], h.normal[2]);
    o.surface_texcoords = vector2(h.uv[0], h.uv[1]);
    o.surface_material = h.material < owner.size() ? owner[h.material] : nullptr;
    o.is_internal = h.is_internal != 0;
  }
  return n_hit;
}

bool Scene::Trace(const ray& trajectory, ObjectCollision* hit_info) {
  return hit_info && TraceRays(&trajectory, 1, hit_info) == 1;
}

const vector3 Scene::SampleSky(uint32 /*depth*/, const vector3& view) {
  TraceState& st = state();
  if (ensure_context() || ensure_scene(this)) return vector3();
  const float v[3] = {view.x, view.y, view.z};
  float rgb[3] = {0.f, 0.f, 0.f};
  if (fs_sample_sky(st.ctx, v, 1, rgb) != FS_OK) {
    fail("fs_sample_sky", st.ctx);
    return vector3();
  }
  return vector3(rgb[0], rgb[1], rgb[2]);
}

void TraceScene(const Camera& view, Scene* scene, DisplayFrame* output, ImagePlaneCache*) {
  TraceSamples(view, scene, output, 1);
}

float32 TraceRange(const Camera& viewer, Scene* scene, DisplayFrame* frame, float32 x, float32 y) {
  if (!scene || !frame) return 0.0f;
  TraceState& st = state();
  if (ensure_context() || ensure_scene(scene)) return viewer.z_far;
  fs_camera cam = to_pod(viewer);
  float range = viewer.z_far;
  if (fs_trace_range(st.ctx, &cam, (int)frame->GetWidth(), (int)frame->GetHeight(), x, y,
                     &range) != FS_OK) {
    fail("fs_trace_range", st.ctx);
    return viewer.z_far;
  }
  return range;
}

}  // namespace base

// ================================================================= C surface (ctypes)
extern "C" {

void* fs2h_scene_new() { return new base::Scene(); }
void fs2h_scene_free(void* s) { delete (base::Scene*)s; }
// Nothing may unwind through the C surface: a damaged asset that drives an allocation over the
// edge (std::bad_alloc / length_error) is a load failure, not an abort.
int fs2h_scene_load(void* s, const char* path) {
  try {
    return ((base::Scene*)s)->LoadScene(path) ? 0 : -1;
  } catch (const std::exception& e) {
    fail_msg((std::string("fs2h_scene_load: ") + e.what()).c_str());
    return -1;
  }
}
void fs2h_initialize_materials() { base::InitializeMaterials(); }
// Restarts the process-wide random stream at the reference's default state, so that a
// scene loaded later in the same process gets the ids / normal table a fresh process would.
void fs2h_reset_rng() { fs2h::main_rng() = fs2h::Xorshift(); }
const fs_scene_desc* fs2h_scene_flatten(void* s) { return ((base::Scene*)s)->Flatten(); }
int fs2h_scene_save_blob(void* s, const char* path) { return ((base::Scene*)s)->SaveBlob(path) ? 0 : -1; }
int fs2h_scene_camera(void* s, int index, fs_camera* out) {
  base::Camera* c = ((base::Scene*)s)->GetCamera((uint32_t)index);
  base::Camera def(base::vector3(-5.80f, 7.05f, -47.06f), base::vector3(0.0f, 8.94f, 0.0f));
  *out = to_pod(c ? *c : def);  // default viewpoint of reference source/main.cpp:87-88
  return c ? 0 : 1;
}
// API-only content for generated scenes (FogMaterial cannot be written in a .scene file,
// reference source/scene.cpp:285-303): a rotated cuboid of fog, through the mirrored C++ API.
// v = position xyz, width, height, depth, rotation axis xyz, angle, colour rgb, density.
int fs2h_scene_add_fog_cuboid(void* s, const float* v) {
  base::Scene* scene = (base::Scene*)s;
  base::CuboidObject* box =
      scene->AddCuboidObject(base::vector3(v[0], v[1], v[2]), v[3], v[4], v[5]);
  if (!box) return -1;
  if (v[9] != 0.0f) box->Rotate(base::vector3(v[6], v[7], v[8]), v[9]);
  box->SetMaterial(std::make_shared<base::FogMaterial>(base::vector3(v[10], v[11], v[12]), v[13]));
  scene->Optimize();
  return 0;
}
void* fs2h_frame_new(uint32_t w, uint32_t h) { return new base::DisplayFrame(w, h); }
void fs2h_frame_free(void* f) { delete (base::DisplayFrame*)f; }
void fs2h_frame_reset(void* f) { ((base::DisplayFrame*)f)->Reset(); }
const uint8_t* fs2h_frame_display(void* f) { return ((base::DisplayFrame*)f)->GetDisplayBuffer(); }
const float* fs2h_frame_mean(void* f) { return ((base::DisplayFrame*)f)->GetRenderTarget(); }
const uint32_t* fs2h_frame_count(void* f) { return ((base::DisplayFrame*)f)->GetCountBuffer(); }

static base::Camera from_pod(const fs_camera* c) {
  base::Camera cam(base::vector3(c->origin[0], c->origin[1], c->origin[2]),
                   base::vector3(c->target[0], c->target[1], c->target[2]));
  cam.fov_y = c->fov_y;
  cam.aperture_size = c->aperture_size;
  cam.focal_depth = c->focal_depth;
  cam.z_near = c->z_near;
  cam.z_far = c->z_far;
  cam.fast_render_enabled = c->fast_render_enabled != 0;
  return cam;
}
int fs2h_trace_samples(void* scene, const fs_camera* cam, void* frame, uint32_t n) {
  return base::TraceSamples(from_pod(cam), (base::Scene*)scene, (base::DisplayFrame*)frame, n);
}
float fs2h_trace_range(void* scene, const fs_camera* cam, void* frame, float x, float y) {
  return base::TraceRange(from_pod(cam), (base::Scene*)scene, (base::DisplayFrame*)frame, x, y);
}
// Scene::Trace / SampleSky / DisplayFrame::WritePixel through the mirror (tests).
// out: param, point xyz, normal xyz, uv, material id, is_internal  (11 floats); returns hit.
int fs2h_scene_trace(void* scene, const float* start, const float* stop, float* out) {
  base::ray r(base::vector3(start[0], start[1], start[2]), base::vector3(stop[0], stop[1], stop[2]));
  base::ObjectCollision c;
  bool hit = ((base::Scene*)scene)->Trace(r, &c);
  out[0] = c.param;
  out[1] = c.point.x, out[2] = c.point.y, out[3] = c.point.z;
  out[4] = c.surface_normal.x, out[5] = c.surface_normal.y, out[6] = c.surface_normal.z;
  out[7] = c.surface_texcoords.x, out[8] = c.surface_texcoords.y;
  out[9] = c.surface_material ? (float)c.surface_material->GetID() : -1.0f;
  out[10] = c.is_internal ? 1.0f : 0.0f;
  return hit ? 1 : 0;
}
void fs2h_scene_sample_sky(void* scene, const float* view, float* rgb) {
  base::vector3 c = ((base::Scene*)scene)->SampleSky(0, base::vector3(view[0], view[1], view[2]));
  rgb[0] = c.x, rgb[1] = c.y, rgb[2] = c.z;
}
void fs2h_frame_write_pixel(void* f, const float* rgb, uint32_t x, uint32_t y) {
  ((base::DisplayFrame*)f)->WritePixel(base::vector3(rgb[0], rgb[1], rgb[2]), x, y);
}
void fs2h_set_device(int device) { base::SetTraceDevice(device); }
void fs2h_set_devices(const int* devices, int n) { base::SetTraceDevices(devices, n); }
void fs2h_set_sharding(int mode) { base::SetTraceSharding(mode); }
int fs2h_device_count() { return base::GetTraceDeviceCount(); }
void fs2h_set_seed(uint64_t seed) { base::SetTraceSeed(seed); }
// Mesh octrees are built by the CUDA library on `device` (-1: by the host builder).
void fs2h_set_gpu_build(int device) { fs2h::gpu_build_device() = device; }
float fs2h_gpu_build_ms() { return fs2h::gpu_build_ms(); }
const char* fs2h_last_error() { return base::LastTraceError(); }
int fs2h_save_exr(const char* path, uint32_t width, uint32_t height, const float* rgb, int bottom_up) {
  return path && fs2h::save_exr(path, width, height, rgb, bottom_up != 0) ? 0 : -1;
}
// Texture readers by file extension (.bmp / .hdr / .exr), for tests: writes width, height and up
// to cap_floats floats (3 per texel, rows in file order); returns 0, -1 unreadable, -2 too small.
int fs2h_load_image(const char* path, uint32_t* width, uint32_t* height, float* rgb, size_t cap_floats) {
  fs2h::Image img;
  std::string p = path ? path : "", err;
  auto ends = [&](const char* e) { return p.size() >= strlen(e) && p.compare(p.size() - strlen(e), strlen(e), e) == 0; };
  bool ok = false;
  try {
    ok = ends(".bmp") ? fs2h::load_bmp(p, &img)
       : ends(".hdr") ? fs2h::load_hdr(p, &img)
       : ends(".exr") ? fs2h::load_exr(p, &img, &err) : false;
  } catch (const std::exception& e) {
    err = e.what();
  }
  if (!ok) {
    fail_msg(("fs2h_load_image: " + p + ": " + (err.empty() ? "unreadable" : err)).c_str());
    return -1;
  }
  *width = img.width;
  *height = img.height;
  if (img.rgb.size() > cap_floats) return -2;
  if (rgb) memcpy(rgb, img.rgb.data(), img.rgb.size() * sizeof(float));
  return 0;
}

}  // extern "C"
