// fs2_octree.h — octree construction into the flattened arrays (see fs2_octree.cpp).
#ifndef FS2_OCTREE_H_
#define FS2_OCTREE_H_

#include <thread>
#include <vector>

#include "fs2_cabi.h"
#include "fs2_hmath.h"
#include "fs2_pack.h"

namespace fs2h {

// reference source/mesh.h:63-74 without the unused per-face material
struct MeshFaceRec {
  uint32_t v[3], n[3], t[3];
  Plane plane;
};

// Growing arrays of an fs_scene_desc.
struct FlatArrays {
  std::vector<fs_node> nodes;
  std::vector<uint32_t> scene_refs;
  std::vector<float> tri_plane;
  std::vector<fs_tri_aux> tri_aux;
  std::vector<fs_face_attr> faces;
  std::vector<fs_object> objects;
  std::vector<fs_material> materials;
  std::vector<fs_texture> textures;
  std::vector<float> texels;
  std::vector<float> lut;
  int32_t scene_root = -1;
  uint32_t sky = 0;
};

// body(i) for every i in [0, n), split over `threads` host threads in contiguous slices; the calls
// must be independent of each other (every loop it is used for writes element i only).
template <class Body>
inline void parallel_for(size_t n, unsigned threads, Body body) {
  if (threads > 32) threads = 32;
  if (threads <= 1 || n < 65536) {
    for (size_t i = 0; i < n; i++) body(i);
    return;
  }
  std::vector<std::thread> pool;
  for (unsigned t = 0; t < threads; t++)
    pool.emplace_back([=]() {
      const size_t lo = n / threads * t, hi = t + 1 == threads ? n : n / threads * (t + 1);
      for (size_t i = lo; i < hi; i++) body(i);
    });
  for (std::thread& t : pool) t.join();
}

// Device the mesh octrees are built on (-1: on the host), and the device time spent there.
int& gpu_build_device();
float& gpu_build_ms();

bool triangle_touches_box(Vec3 p1, Vec3 p2, Vec3 p3, const Box& b);
bool box_touches_box(const Box& a, const Box& b);
void child_boxes(const Box& parent, Box out[8]);

// Builds the per-mesh octree; fills missing face planes.  Returns the root node index or
// 0xFFFFFFFF for an empty mesh.
uint32_t build_mesh_tree(const std::vector<Vec3>& verts, std::vector<MeshFaceRec>& faces,
                         uint32_t face_base, FlatArrays* out, unsigned threads);

// Builds the scene octree when the object count asks for one (>= 182 objects); returns
// the root node index or -1 for "brute force".
int32_t build_scene_tree(const std::vector<Box>& object_boxes, FlatArrays* out);

}  // namespace fs2h
#endif
