// harness.cpp — headless driver around the UNMODIFIED reference renderer.
//
// TEST INFRASTRUCTURE (oracle side).  Built by oracle/ref/Makefile together
// with the reference translation units compiled where they lie under
// /root/reference/source; the result is oracle/_ref/fs2_ref.  It
//   * loads a .scene with the reference's own Scene::LoadScene,
//   * flattens the reference-built structures (object list, scene octree,
//     per-mesh octrees, materials, normal LUT) into an .fs2 blob
//     (include/fs2_blob.h) so the CUDA path / CPU restatement can be run on
//     exactly the data the reference traces,
//   * dumps the reference's first hit per pixel (TraceRange ray construction),
//   * renders N frames with TraceScene and dumps the float mean film,
//   * times TraceScene on the host cores (the CPU baseline).
// This TU is compiled with -fno-access-control to read private members; no
// reference file is edited or copied.  Determinism comes from interposing
// gettimeofday() (per-frame seeds, reference engine.cpp:19-27,179) and
// std::thread::hardware_concurrency() (band count, engine.cpp:174,261-274).

#include <chrono>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "camera.h"
#include "engine.h"
#include "frame.h"
#include "material.h"
#include "math/intersect.h"
#include "math/normal.h"
#include "math/random.h"
#include "mesh.h"
#include "object.h"
#include "scene.h"

#include "fs2_blob.h"
#include "fs2_cabi.h"
#include "fs2_pack.h"
#ifdef FS2_REF_CTR
#include "ctr_rng.h"
#endif

// ---------------------------------------------------------------- interposers
static bool g_fake_clock = true;
static uint64_t g_fake_ms = 1700000000000ull;
static unsigned g_threads = 1;

extern "C" int gettimeofday(struct timeval* tv, void*) noexcept {
  if (g_fake_clock) {
    tv->tv_sec = (time_t)(g_fake_ms / 1000);
    tv->tv_usec = (suseconds_t)((g_fake_ms % 1000) * 1000);
  } else {
    struct timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    tv->tv_sec = ts.tv_sec;
    tv->tv_usec = ts.tv_nsec / 1000;
  }
  return 0;
}

unsigned int std::thread::hardware_concurrency() noexcept { return g_threads; }

namespace base {
extern normal_sphere normal_generator;  // reference material.cpp:15
}

using namespace base;

// ---------------------------------------------------------------- flattening
struct Flat {
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
  std::map<Material*, uint32_t> material_index;
};

static void put3(float* d, const vector3& v) {
  d[0] = v.x;
  d[1] = v.y;
  d[2] = v.z;
}

static uint32_t flatten_material(Flat* F, Material* m) {
  if (!m) return 0xFFFFFFFFu;
  auto it = F->material_index.find(m);
  if (it != F->material_index.end()) return it->second;
  fs_material r;
  memset(&r, 0, sizeof r);
  r.texture = -1;
  r.id = m->GetID();
  DiffuseMaterial* d = dynamic_cast<DiffuseMaterial*>(m);
  if (GlowMaterial* g = dynamic_cast<GlowMaterial*>(m)) {
    r.type = FS_MAT_GLOW;
    put3(r.emissive, g->glow_);
    r.roughness = g->shininess_;
  } else if (CeramicMaterial* c = dynamic_cast<CeramicMaterial*>(m)) {
    r.type = FS_MAT_CERAMIC;
    r.roughness = c->shininess_;
  } else if (FogMaterial* f = dynamic_cast<FogMaterial*>(m)) {
    r.type = FS_MAT_FOG;
    r.density = f->density_;
  } else if (LiquidMaterial* l = dynamic_cast<LiquidMaterial*>(m)) {
    r.type = FS_MAT_LIQUID;
    r.index = l->index_;
    r.reflectivity = l->reflectivity_;
  } else if (GlassMaterial* gl = dynamic_cast<GlassMaterial*>(m)) {
    r.type = FS_MAT_GLASS;
    r.index = gl->index_;
    r.reflectivity = gl->reflectivity_;
    r.frost = gl->frost_;
  } else if (dynamic_cast<MirrorMaterial*>(m)) {
    r.type = FS_MAT_MIRROR;
  } else if (MetalMaterial* me = dynamic_cast<MetalMaterial*>(m)) {
    r.type = FS_MAT_METAL;
    r.roughness = me->roughness_;
  } else if (LightMaterial* li = dynamic_cast<LightMaterial*>(m)) {
    r.type = FS_MAT_LIGHT;
    put3(r.emissive, li->emissive_);
  } else {
    r.type = FS_MAT_DIFFUSE;
  }
  if (d) {
    put3(r.diffuse, d->diffuse_);
    r.tex_scale = d->texture_scale_;
    if (d->diffuse_map_.buffer.size() && d->diffuse_map_.width && d->diffuse_map_.height) {
      fs_texture t;
      t.width = d->diffuse_map_.width;
      t.height = d->diffuse_map_.height;
      t.offset = F->texels.size();
      F->texels.insert(F->texels.end(), d->diffuse_map_.buffer.begin(),
                       d->diffuse_map_.buffer.end());
      r.texture = (int32_t)F->textures.size();
      F->textures.push_back(t);
    }
  }
  uint32_t idx = (uint32_t)F->materials.size();
  F->materials.push_back(r);
  F->material_index[m] = idx;
  return idx;
}

static void set_node_bounds(fs_node* n, const bounds& b) {
  put3(n->bmin, b.bounds_min);
  put3(n->bmax, b.bounds_max);
}

// Mesh tree: node `self` already allocated.
static void flatten_mesh_node(Flat* F, const MeshBvhNode* node, uint32_t self, uint32_t face_base) {
  set_node_bounds(&F->nodes[self], node->aabb_);
  if (node->is_leaf_node_) {
    F->nodes[self].first = (uint32_t)F->tri_aux.size();
    F->nodes[self].count = (uint32_t)node->face_indices_.size();
    for (uint32_t fi : node->face_indices_) {
      const MeshFace& face = node->tree_faces_->at(fi);
      const vector3& a = node->tree_vertices_->at(face.vertex_indices[0]);
      const vector3& b = node->tree_vertices_->at(face.vertex_indices[1]);
      const vector3& c = node->tree_vertices_->at(face.vertex_indices[2]);
      float plane[4] = {face.face_plane.x, face.face_plane.y, face.face_plane.z,
                        face.face_plane.w};
      float out_plane[4];
      fs_tri_aux aux;
      fs_pack_triangle(a.v, b.v, c.v, plane, face_base + fi, out_plane, &aux);
      F->tri_plane.insert(F->tri_plane.end(), out_plane, out_plane + 4);
      F->tri_aux.push_back(aux);
    }
    return;
  }
  uint32_t first = (uint32_t)F->nodes.size();
  F->nodes.resize(first + 8);
  F->nodes[self].first = first;
  F->nodes[self].count = FS_NODE_INTERNAL;
  for (int i = 0; i < 8; i++)
    flatten_mesh_node(F, static_cast<const MeshBvhNode*>(node->children_[i].get()), first + i,
                      face_base);
}

static void flatten_scene_node(Flat* F, const SceneBvhNode* node, uint32_t self) {
  set_node_bounds(&F->nodes[self], node->aabb_);
  if (node->is_leaf_node_) {
    F->nodes[self].first = (uint32_t)F->scene_refs.size();
    F->nodes[self].count = (uint32_t)node->object_indices_.size();
    for (uint32_t oi : node->object_indices_) F->scene_refs.push_back(oi);
    return;
  }
  uint32_t first = (uint32_t)F->nodes.size();
  F->nodes.resize(first + 8);
  F->nodes[self].first = first;
  F->nodes[self].count = FS_NODE_INTERNAL;
  for (int i = 0; i < 8; i++)
    flatten_scene_node(F, static_cast<const SceneBvhNode*>(node->children_[i].get()), first + i);
}

static void flatten_scene(Scene* scene, Flat* F) {
  for (auto& up : scene->object_list_) {
    Object* o = up.get();
    fs_object r;
    memset(&r, 0, sizeof r);
    r.material = flatten_material(F, o->GetMaterial());
    bounds bb = o->GetBounds();
    put3(r.bmin, bb.bounds_min);
    put3(r.bmax, bb.bounds_max);
    if (SphericalObject* s = dynamic_cast<SphericalObject*>(o)) {
      r.type = FS_OBJ_SPHERE;
      put3(r.f, s->origin_);
      r.f[3] = s->radius_;
    } else if (DiscObject* d = dynamic_cast<DiscObject*>(o)) {
      r.type = FS_OBJ_DISC;
      for (int k = 0; k < 4; k++) r.f[k] = d->plane_[k];
      put3(r.f + 4, d->origin_);
      r.f[7] = d->radius_;
    } else if (PlanarObject* p = dynamic_cast<PlanarObject*>(o)) {
      r.type = FS_OBJ_PLANE;
      for (int k = 0; k < 4; k++) r.f[k] = p->plane_[k];
    } else if (CuboidObject* c = dynamic_cast<CuboidObject*>(o)) {
      r.type = FS_OBJ_CUBOID;
      for (int i = 0; i < 6; i++) {
        const plane& pl = c->cube_data_.query_plane(i);
        for (int k = 0; k < 4; k++) r.f[4 * i + k] = pl[k];
      }
    } else if (QuadObject* q = dynamic_cast<QuadObject*>(o)) {
      r.type = FS_OBJ_QUAD;
      for (int k = 0; k < 4; k++) r.f[k] = q->plane_[k];
      put3(r.f + 4, q->origin_);
      put3(r.f + 7, q->tangent_);
      put3(r.f + 10, q->bitangent_);
      r.f[13] = q->half_width_;
      r.f[14] = q->half_height_;
    } else if (MeshObject* m = dynamic_cast<MeshObject*>(o)) {
      r.type = FS_OBJ_MESH;
      uint32_t face_base = (uint32_t)F->faces.size();
      r.u[1] = face_base;
      r.u[2] = (m->normals_.size() ? 1u : 0u) | (m->texcoords_.size() ? 2u : 0u);
      r.u[3] = (uint32_t)m->face_list.size();
      for (const MeshFace& face : m->face_list) {
        fs_face_attr fa;
        memset(&fa, 0, sizeof fa);
        for (int k = 0; k < 3; k++) {
          if (m->normals_.size() && face.normal_indices[k] < m->normals_.size())
            put3(fa.n[k], m->normals_[face.normal_indices[k]]);
          if (m->texcoords_.size() && face.texcoord_indices[k] < m->texcoords_.size()) {
            fa.uv[k][0] = m->texcoords_[face.texcoord_indices[k]].x;
            fa.uv[k][1] = m->texcoords_[face.texcoord_indices[k]].y;
          }
        }
        F->faces.push_back(fa);
      }
      if (m->shape_tree.root_node_.get()) {
        uint32_t root = (uint32_t)F->nodes.size();
        F->nodes.resize(root + 1);
        r.u[0] = root;
        flatten_mesh_node(F, m->shape_tree.root_node_.get(), root, face_base);
      } else {
        r.u[0] = 0xFFFFFFFFu;  // mesh failed to load: never hit (mesh.cpp:154-159)
      }
    }
    F->objects.push_back(r);
  }
  F->sky = flatten_material(F, scene->sky_material_.get());
  if (scene->is_tree_valid_ && scene->object_tree_.root_node_.get()) {
    uint32_t root = (uint32_t)F->nodes.size();
    F->nodes.resize(root + 1);
    F->scene_root = (int32_t)root;
    flatten_scene_node(F, scene->object_tree_.root_node_.get(), root);
  }
  for (const vector3& n : normal_generator.normal_list) {
    F->lut.push_back(n.x);
    F->lut.push_back(n.y);
    F->lut.push_back(n.z);
  }
}

static int save_flat(const Flat& F, const char* path) {
  fs_scene_desc s;
  memset(&s, 0, sizeof s);
  s.abi_version = FS2_ABI_VERSION;
  s.n_nodes = (uint32_t)F.nodes.size();
  s.n_scene_refs = (uint32_t)F.scene_refs.size();
  s.n_tris = (uint32_t)F.tri_aux.size();
  s.n_faces = (uint32_t)F.faces.size();
  s.n_objects = (uint32_t)F.objects.size();
  s.n_materials = (uint32_t)F.materials.size();
  s.n_textures = (uint32_t)F.textures.size();
  s.n_texels = F.texels.size();
  s.scene_root = F.scene_root;
  s.sky_material = F.sky;
  s.n_lut = (uint32_t)(F.lut.size() / 3);
  s.nodes = F.nodes.data();
  s.scene_refs = F.scene_refs.data();
  s.tri_plane = F.tri_plane.data();
  s.tri_aux = F.tri_aux.data();
  s.faces = F.faces.data();
  s.objects = F.objects.data();
  s.materials = F.materials.data();
  s.textures = F.textures.data();
  s.texels = F.texels.data();
  s.lut = F.lut.data();
  return fs_blob_save(&s, path);
}

// ---------------------------------------------------------------- first hit
#pragma pack(push, 1)
struct FirstHitRecord {  // 56 bytes; numpy dtype in tests/refdata.py
  float t;
  float point[3];
  float normal[3];
  float uv[2];
  uint32_t mat_id;
  uint32_t object;
  uint32_t face;
  float range;  // reference TraceRange return value
  uint32_t hit;
};
#pragma pack(pop)

static bool same_bits(float a, float b) { return memcmp(&a, &b, 4) == 0; }

// The camera-ray construction of TraceRange (reference engine.cpp:300-319),
// restated with the reference's own vector classes so that the complete
// ObjectCollision (not only the range) can be captured.  Its agreement with
// the real TraceRange is asserted per pixel below.
static ray range_ray(const Camera& cam, float width, float height, float x, float y) {
  float aspect = width / height;
  float fovy = cam.fov_y * BASE_PI / 180.0;
  float fovx = 2.0 * atan(tan(fovy * 0.5) * aspect);
  vector3 fwd = (cam.target - cam.origin).normalize();
  vector3 right = (vector3(0, 1, 0).cross(fwd)).normalize();
  vector3 up = (fwd.cross(right)).normalize();
  float hh = tanf(fovy * 0.5f) * cam.z_far;
  float hw = tanf(fovx * 0.5f) * cam.z_far;
  vector3 po = cam.origin + fwd * cam.z_far;
  float xd = hw * ((x / (width - 1)) * 2.0 - 1.0);
  float yd = hh * ((y / (height - 1)) * 2.0 - 1.0);
  vector3 stop = po + right * xd + up * yd;
  return ray(cam.origin, stop);
}

static int dump_first_hit(Scene* scene, const Camera& cam, uint32_t W, uint32_t H,
                          const char* path) {
  DisplayFrame frame(W, H);
  std::vector<FirstHitRecord> out((size_t)W * H);
  size_t range_mismatch = 0;
  for (uint32_t j = 0; j < H; j++)
    for (uint32_t i = 0; i < W; i++) {
      FirstHitRecord& r = out[(size_t)j * W + i];
      memset(&r, 0, sizeof r);
      r.range = TraceRange(cam, scene, &frame, (float)i, (float)j);
      ray tr = range_ray(cam, (float)W, (float)H, (float)i, (float)j);
      ObjectCollision hit;
      bool got = scene->Trace(tr, &hit);
      r.hit = got ? 1u : 0u;
      r.object = 0xFFFFFFFFu;
      r.face = 0xFFFFFFFFu;
      if (!got) {
        r.t = FS_NO_HIT_PARAM;
        r.mat_id = scene->GetSkyMaterial()->GetID();
        if (!same_bits(r.range, cam.z_far)) range_mismatch++;
        continue;
      }
      r.t = hit.param;
      put3(r.point, hit.point);
      put3(r.normal, hit.surface_normal);
      r.uv[0] = hit.surface_texcoords.x;
      r.uv[1] = hit.surface_texcoords.y;
      r.mat_id = hit.surface_material ? hit.surface_material->GetID() : 0;
      if (!same_bits(r.range, (hit.point - cam.origin).length())) range_mismatch++;
      // Object id: first object (list order) whose own Trace on this ray gives
      // the identical parameter and point.
      for (size_t k = 0; k < scene->object_list_.size(); k++) {
        ObjectCollision c;
        Object* o = scene->object_list_[k].get();
        if (o->Trace(tr, &c) && same_bits(c.param, hit.param) &&
            same_bits(c.point.x, hit.point.x) && same_bits(c.point.y, hit.point.y) &&
            same_bits(c.point.z, hit.point.z)) {
          r.object = (uint32_t)k;
          if (MeshObject* m = dynamic_cast<MeshObject*>(o)) {
            MeshCollision mc;
            if (m->shape_tree.Trace(tr, &mc)) r.face = mc.face_index;
          }
          break;
        }
      }
    }
  FILE* f = fopen(path, "wb");
  if (!f) return 1;
  uint32_t hdr[4] = {0x48463246u /* "F2FH" */, W, H, (uint32_t)sizeof(FirstHitRecord)};
  fwrite(hdr, sizeof hdr, 1, f);
  fwrite(out.data(), sizeof(FirstHitRecord), out.size(), f);
  fclose(f);
  printf("FS2REF firsthit %ux%u written to %s; TraceRange consistency mismatches: %zu\n", W, H,
         path, range_mismatch);
  return range_mismatch ? 2 : 0;
}

// ---------------------------------------------------------------- render
static int render(Scene* scene, const Camera& cam, uint32_t W, uint32_t H, uint32_t frames,
                  const char* path, const char* json_path) {
  DisplayFrame frame(W, H);
  uint64_t clock0 = g_fake_ms;
  auto t0 = std::chrono::steady_clock::now();
  for (uint32_t k = 0; k < frames; k++) {
    if (g_fake_clock) g_fake_ms = clock0 + 1000ull * k + 7ull * k * k;
    TraceScene(cam, scene, &frame);
  }
  auto t1 = std::chrono::steady_clock::now();
  double sec = std::chrono::duration<double>(t1 - t0).count();
  double paths = (double)W * H * frames;
  if (path && path[0]) {
    FILE* f = fopen(path, "wb");
    if (!f) return 1;
    uint32_t hdr[4] = {0x4D463246u /* "F2FM" */, W, H, frames};
    fwrite(hdr, sizeof hdr, 1, f);
    fwrite(frame.render_target_, sizeof(vector3), (size_t)W * H, f);
    fwrite(frame.count_buffer_, sizeof(uint32), (size_t)W * H, f);
    fwrite(frame.display_buffer_, 3, (size_t)W * H, f);
    fwrite(frame.normal_buffer_, sizeof(vector3), (size_t)W * H, f);
    fwrite(frame.depth_buffer_, sizeof(float32), (size_t)W * H, f);
    fwrite(frame.material_id_buffer_, sizeof(uint64), (size_t)W * H, f);
    fclose(f);
  }
  char line[512];
  snprintf(line, sizeof line,
           "{\"width\": %u, \"height\": %u, \"frames\": %u, \"threads\": %u, \"seconds\": %.6f, "
           "\"mpaths_per_s\": %.6f}",
           W, H, frames, g_threads, sec, paths / sec / 1e6);
  printf("FS2REF %s\n", line);
  if (json_path && json_path[0]) {
    FILE* f = fopen(json_path, "w");
    if (f) {
      fprintf(f, "%s\n", line);
      fclose(f);
    }
  }
  return 0;
}

#ifdef FS2_REF_CTR
// ---------------------------------------------------------------- per-path dump
// fs2_ref_ctr only: `samples` TraceScene calls of the unmodified reference with one worker
// thread, every path drawing from the device's stream of (seed, pixel, sample index)
// (ctr_rng.cpp).  File: u32 magic "F2PT", W, H, samples, then [samples][H*W] PathRecord.
static int dump_paths(Scene* scene, const Camera& cam, uint32_t W, uint32_t H, uint32_t samples,
                      uint32_t first_sample, uint32_t stride, uint64_t seed, const char* path) {
  DisplayFrame frame(W, H);
  std::vector<fs2ctr::PathRecord> out((size_t)samples * W * H);
  memset(out.data(), 0, out.size() * sizeof(fs2ctr::PathRecord));
  g_threads = 1;
  fs2ctr::g.seed = seed;
  fs2ctr::g.out = out.data();
  fs2ctr::g.active = true;
  for (uint32_t k = 0; k < samples; k++) {
    fs2ctr::g.sample = first_sample + k * stride;
    fs2ctr::g.sample_slot = k;
    TraceScene(cam, scene, &frame);
  }
  fs2ctr::g.active = false;
  FILE* f = fopen(path, "wb");
  if (!f) return 1;
  uint32_t hdr[4] = {0x54503246u /* "F2PT" */, W, H, samples};
  fwrite(hdr, sizeof hdr, 1, f);
  fwrite(out.data(), sizeof(fs2ctr::PathRecord), out.size(), f);
  fclose(f);
  printf("FS2REF paths %ux%u x %u samples written to %s; scan-order errors: %llu\n", W, H, samples,
         path, (unsigned long long)fs2ctr::g.order_errors);
  return fs2ctr::g.order_errors ? 2 : 0;
}
#endif

// ---------------------------------------------------------------- built-in API scene
// Objects and materials the .scene grammar cannot express: PlanarObject,
// DiscObject, FogMaterial, GlowMaterial, textured Metal — built through the reference's
// own C++ API (source/scene.h:94-133, source/material.h:88-328).
static void build_api_scene(Scene* scene, const char* texture_path) {
  auto sky = std::make_shared<LightMaterial>(vector3(0.35f, 0.45f, 0.6f));
  scene->SetSkyMaterial(sky);
  auto floor_mat = std::make_shared<DiffuseMaterial>(vector3(0.7f, 0.7f, 0.65f));
  if (texture_path && texture_path[0]) floor_mat->LoadDiffuseTexture(texture_path, 0.05f);
  auto fog = std::make_shared<FogMaterial>(vector3(0.8f, 0.85f, 0.9f), 0.6f);
  auto glow = std::make_shared<GlowMaterial>(vector3(0.3f, 0.2f, 0.8f), vector3(0.4f, 0.1f, 0.9f), 0.7f);
  auto metal = std::make_shared<MetalMaterial>(vector3(0.9f, 0.6f, 0.3f), 0.97f);
  auto ceramic = std::make_shared<CeramicMaterial>(vector3(0.8f, 0.3f, 0.2f), 0.4f);
  auto lamp = std::make_shared<LightMaterial>(vector3(14.0f, 13.0f, 11.0f));
  auto glass = std::make_shared<GlassMaterial>(vector3(0.95f, 0.97f, 1.0f), 1.0f, 0.05f, 0.35f);
  auto diffuse = std::make_shared<DiffuseMaterial>(vector3(0.2f, 0.6f, 0.3f));

  const char* mask_env = getenv("FS2_API_MASK");  // debugging aid: bit k keeps object k
  unsigned mask = mask_env ? (unsigned)strtoul(mask_env, nullptr, 0) : 0xFFu;
  if (mask & 1) scene->AddPlanarObject(calculate_plane(vector3(0.01f, 1.0f, 0.02f).normalize(), vector3(0, 0, 0)))
      ->SetMaterial(floor_mat);
  if (mask & 2) scene->AddDiscObject(vector3(-14.0f, 6.0f, 10.0f), vector3(0.3f, 0.2f, -1.0f).normalize(), 6.0f)
      ->SetMaterial(metal);
  if (mask & 4) {
    CuboidObject* fog_box = scene->AddCuboidObject(vector3(4.0f, 7.0f, -8.0f), 14.0f, 12.0f, 10.0f);
    fog_box->Rotate(vector3(0, 1, 0), 0.35f);
    fog_box->SetMaterial(fog);
  }
  if (mask & 8) scene->AddSphericalObject(vector3(4.0f, 6.0f, -8.0f), 3.0f)->SetMaterial(glow);
  if (mask & 16) scene->AddSphericalObject(vector3(-6.0f, 4.0f, -12.0f), 4.0f)->SetMaterial(ceramic);
  if (mask & 32) scene->AddSphericalObject(vector3(14.0f, 5.0f, 2.0f), 5.0f)->SetMaterial(glass);
  if (mask & 64) scene->AddSphericalObject(vector3(-2.0f, 32.0f, -6.0f), 5.0f)->SetMaterial(lamp);
  if (mask & 128) scene->AddQuadObject(vector3(-20.0f, 14.0f, 16.0f), vector3(12.0f, 0.0f, 3.0f),
                       vector3(0.0f, -10.0f, 0.5f))
      ->SetMaterial(diffuse);
  scene->Optimize();
}

// ---------------------------------------------------------------- main
static void usage() {
  fprintf(stderr,
          "usage: fs2_ref SCENE [--width W] [--height H] [--flatten out.fs2]\n"
          "               [--firsthit out.bin] [--render out.bin] [--frames N] [--threads T]\n"
          "               [--clock0 MS | --realclock] [--aperture A] [--focal F] [--json out.json] [--fast]\n"
          "       fs2_ref_ctr only: [--paths out.bin --samples K --first-sample F --stride T --seed S]\n");
}

int main(int argc, char** argv) {
  if (argc < 2) {
    usage();
    return 64;
  }
  std::string scene_path = argv[1];
  uint32_t W = 64, H = 64, frames = 1;
  const char *flatten_path = nullptr, *firsthit_path = nullptr, *render_path = nullptr,
             *json_path = nullptr;
  bool do_render = false, have_aperture = false, have_focal = false, fast = false;
  float aperture = 0, focal = 0;
  float fog[14];  // --fog: position, size, rotation axis + angle, colour, density
  bool have_fog = false;
  const char* paths_path = nullptr;
  uint32_t path_samples = 1, path_first = 0, path_stride = 1;
  uint64_t path_seed = 1;
  for (int i = 2; i < argc; i++) {
    std::string a = argv[i];
    auto next = [&]() -> const char* { return (i + 1 < argc) ? argv[++i] : ""; };
    if (a == "--width") W = (uint32_t)atoi(next());
    else if (a == "--height") H = (uint32_t)atoi(next());
    else if (a == "--flatten") flatten_path = next();
    else if (a == "--firsthit") firsthit_path = next();
    else if (a == "--render") { render_path = next(); do_render = true; }
    else if (a == "--time") do_render = true;
    else if (a == "--frames") frames = (uint32_t)atoi(next());
    else if (a == "--threads") g_threads = (unsigned)atoi(next());
    else if (a == "--clock0") g_fake_ms = strtoull(next(), nullptr, 10);
    else if (a == "--realclock") g_fake_clock = false;
    else if (a == "--aperture") { aperture = (float)atof(next()); have_aperture = true; }
    else if (a == "--focal") { focal = (float)atof(next()); have_focal = true; }
    else if (a == "--json") json_path = next();
    else if (a == "--fast") fast = true;  // Camera::fast_render_enabled (engine.cpp:68-70)
    else if (a == "--fog") {  // 14 comma-separated floats: a FogMaterial cuboid added through the API
      have_fog = sscanf(next(), "%f,%f,%f,%f,%f,%f,%f,%f,%f,%f,%f,%f,%f,%f", fog, fog + 1, fog + 2, fog + 3,
                        fog + 4, fog + 5, fog + 6, fog + 7, fog + 8, fog + 9, fog + 10, fog + 11, fog + 12,
                        fog + 13) == 14;
      if (!have_fog) { usage(); return 64; }
    }
    else if (a == "--paths") paths_path = next();
    else if (a == "--samples") path_samples = (uint32_t)atoi(next());
    else if (a == "--first-sample") path_first = (uint32_t)atoi(next());
    else if (a == "--stride") path_stride = (uint32_t)atoi(next());
    else if (a == "--seed") path_seed = strtoull(next(), nullptr, 10);
    else { usage(); return 64; }
  }
  if (g_threads == 0) g_threads = 1;

  // Same order as reference main.cpp:86-101: Scene ctor (sky material id),
  // LoadScene (material ids), then InitializeMaterials (the LUT continues the
  // main thread's default-seeded RNG stream).
  Scene scene;
  if (scene_path.rfind("@api", 0) == 0) {
    // "@api" or "@api:TEXTURE.bmp"
    build_api_scene(&scene, scene_path.size() > 5 ? scene_path.c_str() + 5 : "");
  } else if (!scene.LoadScene(scene_path)) {
    return 1;
  }
  if (have_fog) {  // FogMaterial is API-only (scene.cpp:285-303): same call sequence as the host mirror
    CuboidObject* box = scene.AddCuboidObject(vector3(fog[0], fog[1], fog[2]), fog[3], fog[4], fog[5]);
    if (fog[9] != 0.0f) box->Rotate(vector3(fog[6], fog[7], fog[8]), fog[9]);
    box->SetMaterial(std::make_shared<FogMaterial>(vector3(fog[10], fog[11], fog[12]), fog[13]));
    scene.Optimize();
  }
  InitializeMaterials();

  Camera cam(vector3(-5.80f, 7.05f, -47.06f), vector3(0.0f, 8.94f, 0.0f));  // main.cpp:87-88
  if (scene.GetCameraCount() > 0) cam = *scene.GetCamera(0);
  if (have_aperture) cam.aperture_size = aperture;
  if (have_focal) cam.focal_depth = focal;
  if (fast) cam.fast_render_enabled = true;

  int rc = 0;
  if (flatten_path) {
    Flat F;
    flatten_scene(&scene, &F);
    if (save_flat(F, flatten_path) != FS_OK) return 1;
    printf("FS2REF flatten: %zu nodes, %zu tri refs, %zu faces, %zu objects, %zu materials, "
           "scene_root %d -> %s\n",
           F.nodes.size(), F.tri_aux.size(), F.faces.size(), F.objects.size(),
           F.materials.size(), F.scene_root, flatten_path);
    // camera sidecar so consumers use the same camera as this run
    std::string cpath = std::string(flatten_path) + ".camera";
    FILE* f = fopen(cpath.c_str(), "wb");
    if (f) {
      fs_camera c;
      memset(&c, 0, sizeof c);
      put3(c.origin, cam.origin);
      put3(c.target, cam.target);
      c.fov_y = cam.fov_y;
      c.aperture_size = cam.aperture_size;
      c.focal_depth = cam.focal_depth;
      c.z_near = cam.z_near;
      c.z_far = cam.z_far;
      c.fast_render_enabled = cam.fast_render_enabled ? 1 : 0;
      fwrite(&c, sizeof c, 1, f);
      fclose(f);
    }
  }
  if (firsthit_path) rc |= dump_first_hit(&scene, cam, W, H, firsthit_path);
  if (do_render) rc |= render(&scene, cam, W, H, frames, render_path, json_path);
  if (paths_path) {
#ifdef FS2_REF_CTR
    rc |= dump_paths(&scene, cam, W, H, path_samples, path_first, path_stride, path_seed, paths_path);
#else
    fprintf(stderr, "--paths needs fs2_ref_ctr (the binary linked with ctr_rng.cpp)\n");
    return 64;
#endif
  }
  return rc;
}
