// fs2_impl.h — private records behind the public classes of fs2_host.h.
#ifndef FS2_IMPL_H_
#define FS2_IMPL_H_

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "fs2_assets.h"
#include "fs2_host.h"
#include "fs2_octree.h"

namespace fs2h {

// process-wide main-thread random stream (material ids, then the normal table)
Xorshift& main_rng();
const std::vector<float>& normal_table();  // 3 floats per entry; empty before InitializeMaterials

struct MaterialImpl {
  fs_material rec;
  Image texture;
};

struct ObjectImpl {
  fs_object rec;  // type + float payload; material / mesh indices are filled when flattening
  Box box;        // Object::GetBounds()
  // cuboid
  Box cube_box;
  Vec3 cube_verts[8];
  // mesh
  std::vector<Vec3> verts, normals;
  std::vector<float> texcoords;  // 2 per entry
  std::vector<MeshFaceRec> faces;
  bool mesh_loaded = false;
};

struct SceneImpl {
  std::vector<std::unique_ptr<base::Object>> objects;
  std::vector<base::Camera> cameras;
  std::shared_ptr<base::LightMaterial> sky;
  bool tree_valid = false;
  // flattened form
  bool flat_valid = false;
  uint64_t flat_version = 0;
  FlatArrays flat;
  fs_scene_desc desc;
  std::vector<base::Material*> flat_material_owner;  // flattened material index -> object
};

struct FrameImpl {
  uint32_t width = 0, height = 0, frame_count = 0;
  std::vector<float> mean, normal, depth;
  std::vector<uint32_t> count;
  std::vector<uint8_t> rgb8;
  std::vector<uint64_t> mat_id;
  uint32_t samples = 0;  // samples accumulated since the last Reset()
  bool device_dirty = true;  // device film must be cleared before the next launch
  bool pinned = false;       // buffers are page-locked (fs_host_register)
  bool zero_pending = false; // Reset() was called: the host buffers read as zero, the fill is deferred
  bool aux_stale = false;    // normal / depth / material-id planes are newer on the device
  bool host_written = false; // WritePixel changed the host buffers: upload before the next launch
};

inline Vec3 to3(const base::vector3& v) { return Vec3(v.x, v.y, v.z); }
inline void put3(float* d, Vec3 v) {
  d[0] = v.x;
  d[1] = v.y;
  d[2] = v.z;
}

}  // namespace fs2h
#endif
