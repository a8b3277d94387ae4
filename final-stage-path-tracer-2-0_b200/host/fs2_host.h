// fs2_host.h — C++ host mirror of Final Stage 2.0's engine / scene / material / camera API
// for the path-tracing hot path (libfs2host.so).
//
// Same names, argument meaning and ownership rules as the reference's public classes in
// namespace base, so that code written against the reference (e.g. its main.cpp loop)
// compiles against this header:
//   Camera                         reference source/camera.h:43-63
//   Material family                reference source/material.h:54-330
//   Object family                  reference source/object.h:63-152, source/mesh.h:124-149
//   Scene                          reference source/scene.h:94-185
//   DisplayFrame                   reference source/frame.h:49-92
//   TraceScene / TraceRange        reference source/engine.h:68-75
// What differs is where the work happens: scene loading, octree construction and
// flattening stay on the host (this library); ray generation, traversal, intersection,
// shading and film accumulation run in libfs2cuda.so behind include/fs2_cabi.h.  There is
// no CPU tracing path here: TraceScene fails loudly (prints + returns) without a GPU.
#ifndef FS2_HOST_H_
#define FS2_HOST_H_

#include <stdint.h>

#include <memory>
#include <string>
#include <vector>

#include "fs2_cabi.h"

namespace fs2h {
struct SceneImpl;
struct MaterialImpl;
struct ObjectImpl;
struct FrameImpl;
}  // namespace fs2h

namespace base {

typedef float float32;
typedef uint8_t uint8;
typedef uint32_t uint32;
typedef int32_t int32;
typedef uint64_t uint64;

struct vector2 {
  float32 x, y;
  vector2() : x(0), y(0) {}
  vector2(float32 a, float32 b) : x(a), y(b) {}
};
struct vector3 {
  float32 x, y, z;
  vector3() : x(0), y(0), z(0) {}
  vector3(float32 a, float32 b, float32 c) : x(a), y(b), z(c) {}
};
struct vector4 {
  float32 x, y, z, w;
  vector4() : x(0), y(0), z(0), w(0) {}
  vector4(float32 a, float32 b, float32 c, float32 d) : x(a), y(b), z(c), w(d) {}
};
typedef vector4 plane;

// reference source/math/trace.h:40-54 (the part the tracing API needs)
struct ray {
  vector3 start, stop, dir;
  ray(const vector3& init_start, const vector3& init_stop)
      : start(init_start), stop(init_stop),
        dir(init_stop.x - init_start.x, init_stop.y - init_start.y, init_stop.z - init_start.z) {}
};

// ---------------------------------------------------------------- camera
class Camera {
 public:
  Camera();
  Camera(const vector3& new_origin, const vector3& new_target);
  vector3 origin;
  vector3 target;
  float32 fov_y;
  float32 aperture_size;
  float32 focal_depth;
  float32 z_near;
  float32 z_far;
  bool fast_render_enabled;
};

// ---------------------------------------------------------------- materials
// Host-side records only: Reflection()/Sample() are the CUDA kernels' job.
class Material {
 public:
  virtual ~Material();
  uint32 GetID() const;
  virtual bool IsLight() const { return false; }
  const fs_material& Record() const;
  fs2h::MaterialImpl* impl() const { return impl_.get(); }

 protected:
  explicit Material(uint32_t type);  // draws the id from the process-wide stream
  std::unique_ptr<fs2h::MaterialImpl> impl_;
};

class DiffuseMaterial : public Material {
 public:
  DiffuseMaterial();
  explicit DiffuseMaterial(const vector3& diffuse);
  explicit DiffuseMaterial(const std::string& filename, float32 tex_scale = 1.0f);
  // .bmp and .exr as in the reference (.exr: scan-line HALF / FLOAT files, host/fs2_exr.cpp);
  // .hdr (Radiance RGBE) is an extension.
  void LoadDiffuseTexture(const std::string& filename, float32 tex_scale = 1.0f);

 protected:
  explicit DiffuseMaterial(uint32_t type);
};

class LightMaterial : public DiffuseMaterial {
 public:
  LightMaterial();
  explicit LightMaterial(const vector3& emissive);
  bool IsLight() const override { return true; }
};
class MetalMaterial : public DiffuseMaterial {
 public:
  MetalMaterial(const vector3& diffuse, float32 roughness);
};
class MirrorMaterial : public DiffuseMaterial {
 public:
  explicit MirrorMaterial(const vector3& diffuse);
};
class GlassMaterial : public DiffuseMaterial {
 public:
  GlassMaterial(const vector3 diffuse, float32 index = 0.75f, float32 reflectivity = 0.1f,
                float32 frost = 0.0f);
};
class LiquidMaterial : public DiffuseMaterial {
 public:
  LiquidMaterial(const vector3 diffuse, float32 index = 0.75f, float32 reflectivity = 0.4f);
};
class CeramicMaterial : public DiffuseMaterial {
 public:
  CeramicMaterial(const vector3& diffuse, float32 shininess);

 protected:
  CeramicMaterial(uint32_t type, const vector3& diffuse, float32 shininess);
};
class GlowMaterial : public CeramicMaterial {
 public:
  GlowMaterial(const vector3& diffuse, const vector3& glow, float32 shininess);
};
class FogMaterial : public DiffuseMaterial {
 public:
  FogMaterial(const vector3 diffuse, float32 density);
};

// Fills the 32 K random-normal table from the process-wide stream (reference
// source/material.cpp:17, source/math/normal.cpp:9-27).  Call once after the scene's
// materials exist and before tracing, like the reference's main.cpp does.
void InitializeMaterials();

// ---------------------------------------------------------------- objects
class Object {
 public:
  virtual ~Object();
  void SetMaterial(std::shared_ptr<Material> material);
  Material* GetMaterial() { return material_.get(); }
  void GetBounds(vector3* bmin, vector3* bmax) const;
  fs2h::ObjectImpl* impl() const { return impl_.get(); }

 protected:
  Object();
  friend class Scene;
  std::shared_ptr<Material> material_;
  std::unique_ptr<fs2h::ObjectImpl> impl_;
};
class SphericalObject : public Object {
 public:
  SphericalObject(const vector3& origin, float32 radius);
};
class PlanarObject : public Object {
 public:
  explicit PlanarObject(const plane& data);
};
class DiscObject : public Object {
 public:
  DiscObject(const vector3& origin, const vector3& normal, float32 radius);
};
class CuboidObject : public Object {
 public:
  CuboidObject(const vector3& origin, float32 width, float32 height, float32 depth);
  void Rotate(const vector3& axis, float32 angle);
};
class QuadObject : public Object {
 public:
  QuadObject(const vector3& origin, const vector3& normal, float32 width, float32 height);
  QuadObject(const vector3& position, const vector3& u, const vector3& v);
};
class MeshObject : public Object {
 public:
  MeshObject(const std::string& filename, bool invert_normals = false,
             const vector3& translation = vector3(0, 0, 0),
             const vector3& scale = vector3(1, 1, 1),
             const vector4& rotation = vector4(0, 0, 0, 0));
};

// reference source/object.h:47-61
class Material;
struct ObjectCollision {
  float32 param;              // 2.0 (collision ctor, trace.cpp:18-21) when nothing was hit
  vector3 point;
  vector3 surface_normal;
  vector2 surface_texcoords;
  Material* surface_material;
  bool is_internal;
  ObjectCollision() : param(2.0f), surface_material(nullptr), is_internal(false) {}
};

// reference source/frame.h:40-47
struct TraceResult {
  vector3 color;
  vector3 normal;
  float32 depth;
  uint64 material_id;
  uint64 ray_count;
  TraceResult() : depth(0.0f), material_id(0), ray_count(0) {}
};

// ---------------------------------------------------------------- scene
class Scene {
 public:
  Scene();
  ~Scene();
  bool LoadScene(const std::string& filename);
  MeshObject* AddMeshObject(const std::string& filename, bool invert_normals = false,
                            const vector3& translation = vector3(0, 0, 0),
                            const vector3& scale = vector3(1, 1, 1),
                            const vector4& rotation = vector4(0, 0, 0, 0));
  SphericalObject* AddSphericalObject(const vector3& origin, float32 radius);
  DiscObject* AddDiscObject(const vector3& origin, const vector3& normal, float32 radius);
  PlanarObject* AddPlanarObject(const plane& data);
  CuboidObject* AddCuboidObject(const vector3& origin, float32 width, float32 height,
                                float32 depth);
  QuadObject* AddQuadObject(const vector3& origin, const vector3& normal, float32 width,
                            float32 height);
  QuadObject* AddQuadObject(const vector3& position, const vector3& u, const vector3& v);
  void SetSkyMaterial(std::shared_ptr<LightMaterial> material);
  LightMaterial* GetSkyMaterial();
  // One ray against the scene on the device (fs_trace_rays; scene.cpp:227-249).  A single ray
  // per call pays a launch and two copies: for many rays use TraceRays below.
  bool Trace(const ray& trajectory, ObjectCollision* hit_info);
  // Sky radiance seen along `view` (fs_sample_sky; scene.cpp:157-162).
  const vector3 SampleSky(uint32 depth, const vector3& view);
  void Optimize();
  uint32 GetCameraCount();
  Camera* GetCamera(uint32 index);

  // ---- additions for the GPU path
  // Batched Scene::Trace: n rays, hit_info[i] filled like Trace(); returns the number of hits.
  int TraceRays(const ray* trajectories, int n, ObjectCollision* hit_info);
  // Flattens objects, octrees, materials, textures and the normal table into the arrays
  // of include/fs2_cabi.h.  The returned description stays valid until the scene changes.
  const fs_scene_desc* Flatten();
  bool SaveBlob(const std::string& path);  // include/fs2_blob.h
  fs2h::SceneImpl* impl() { return impl_.get(); }

 private:
  std::unique_ptr<fs2h::SceneImpl> impl_;
};

// ---------------------------------------------------------------- film
class ImagePlaneCache;  // accepted and ignored (dead code in the reference, engine.cpp:33-57)

class DisplayFrame {
 public:
  DisplayFrame(uint32 width, uint32 height);
  ~DisplayFrame();
  void Reset();
  // frame.cpp:53-85: running mean, count and display byte of one pixel, on the host buffers;
  // the next TraceScene / TraceSamples call uploads the frame before it adds its samples.
  void WritePixel(const vector3& pixel, uint32 x, uint32 y);
  void WritePixel(const TraceResult& result, uint32 x, uint32 y);
  uint8* GetDisplayBuffer();
  uint32 GetWidth();
  uint32 GetHeight();
  float32 GetAspectRatio();
  void SetFrameCount(uint32 count);
  // ---- additions: read access to the buffers the reference keeps private (frame.h:79-91)
  const float32* GetRenderTarget();     // float mean, 3 per pixel
  const uint32* GetCountBuffer();
  const float32* GetNormalBuffer();
  const float32* GetDepthBuffer();
  const uint64* GetMaterialIdBuffer();
  fs2h::FrameImpl* impl() { return impl_.get(); }

 private:
  std::unique_ptr<fs2h::FrameImpl> impl_;
};

// One sample per pixel appended to `output` (never cleared here), like the reference.
void TraceScene(const Camera& view, Scene* scene, DisplayFrame* output,
                ImagePlaneCache* cache = nullptr);
// Distance to the object seen at fractional pixel (x, y); z_far on a miss.
float32 TraceRange(const Camera& viewer, Scene* scene, DisplayFrame* frame, float32 x, float32 y);

// ---- additions
// n samples per pixel in one launch (TraceScene == TraceSamples(…, 1)).
int TraceSamples(const Camera& view, Scene* scene, DisplayFrame* output, uint32 n_samples);
// Selects the CUDA device used by TraceScene/TraceRange (default 0) and the seed.
void SetTraceDevice(int device);
// Puts several GPUs of the box behind TraceScene / TraceSamples (fs_create_multi): the scene is
// replicated, every call's samples (or, with fewer samples than GPUs, interleaved pixel tiles)
// are sharded across them and the film is combined on devices[0] before it is copied into the
// DisplayFrame.  The GPU counterpart of the reference's fan-out over hardware_concurrency()
// threads (engine.cpp:261-274).
void SetTraceDevices(const int* devices, int n_devices);
int GetTraceDeviceCount();
// 0 automatic (sample indices; tiles when a call has fewer samples than GPUs), 1 samples, 2 tiles
void SetTraceSharding(int mode);
void SetTraceSeed(uint64 seed);
const char* LastTraceError();
// Film checkpoint / resume (headless replacement of keeping the window open): the accumulation
// film of `frame` goes to / comes from `path`; after a load the next TraceScene / TraceSamples
// call continues the sample sequence where the checkpoint stopped.
bool SaveFilmCheckpoint(DisplayFrame* frame, const char* path);
bool LoadFilmCheckpoint(DisplayFrame* frame, const char* path);

// Writes the frame's float mean image as an OpenEXR file (FLOAT R, G, B, ZIP; host/fs2_exr.cpp):
// the lossless output of the headless front end next to PPM / PFM.
bool SaveFrameEXR(DisplayFrame* frame, const char* path);

// Film post-processing on the device (fs_film_post): guided a-trous denoise, bloom and flare of
// the frame's resolved image, using its normal / depth / material-id planes (frame.h:86-91).
// The reference's README describes these stages, its code has none.  Outputs (either may be
// null): rgb = 3 floats per pixel, rgb8 = display bytes; the frame itself is not modified.
struct PostProcess {
  int denoise_passes = 0;
  float sigma_normal = 64.0f, sigma_depth = 0.02f, sigma_colour = 4.0f;
  float bloom_threshold = 1.0f, bloom_strength = 0.0f;
  int bloom_radius = 12;
  float flare_strength = 0.0f;
  int flare_ghosts = 0;
};
bool PostProcessFrame(DisplayFrame* frame, const PostProcess& post, float* rgb, uint8* rgb8);

}  // namespace base
#endif
