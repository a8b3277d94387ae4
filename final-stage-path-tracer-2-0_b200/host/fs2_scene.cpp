// fs2_scene.cpp — Scene: object factory, .scene parser, octree construction and flattening.
//   factory / Optimize   reference source/scene.cpp:142-225
//   .scene grammar       reference source/scene.cpp:251-535
// The parser is a table-driven reader written for this project; it keeps the reference's
// observable rules: '#' comment lines, a block ends at the first line containing '}',
// keys are matched per line, unknown keys are ignored, material class is chosen from
// which keys are non-zero (scene.cpp:285-303), materials must be defined before use.
#include <stdio.h>
#include <stdlib.h>

#include <fstream>
#include <thread>

#include "fs2_blob.h"
#include "fs2_impl.h"

namespace base {

using namespace fs2h;

Scene::Scene() : impl_(new SceneImpl()) {
  impl_->sky.reset(new LightMaterial(vector3(0, 0, 0)));  // scene.cpp:142-144
  memset(&impl_->desc, 0, sizeof impl_->desc);
}
Scene::~Scene() {}

template <class T>
static T* add_object(SceneImpl* s, T* obj) {
  s->tree_valid = false;
  s->flat_valid = false;
  s->objects.emplace_back(obj);
  return obj;
}

MeshObject* Scene::AddMeshObject(const std::string& filename, bool invert_normals,
                                 const vector3& translation, const vector3& scale,
                                 const vector4& rotation) {
  return add_object(impl_.get(),
                    new MeshObject(filename, invert_normals, translation, scale, rotation));
}
SphericalObject* Scene::AddSphericalObject(const vector3& origin, float32 radius) {
  return add_object(impl_.get(), new SphericalObject(origin, radius));
}
DiscObject* Scene::AddDiscObject(const vector3& origin, const vector3& normal, float32 radius) {
  return add_object(impl_.get(), new DiscObject(origin, normal, radius));
}
PlanarObject* Scene::AddPlanarObject(const plane& data) {
  return add_object(impl_.get(), new PlanarObject(data));
}
CuboidObject* Scene::AddCuboidObject(const vector3& origin, float32 w, float32 h, float32 d) {
  return add_object(impl_.get(), new CuboidObject(origin, w, h, d));
}
QuadObject* Scene::AddQuadObject(const vector3& origin, const vector3& normal, float32 w,
                                 float32 h) {
  return add_object(impl_.get(), new QuadObject(origin, normal, w, h));
}
QuadObject* Scene::AddQuadObject(const vector3& position, const vector3& u, const vector3& v) {
  return add_object(impl_.get(), new QuadObject(position, u, v));
}
void Scene::SetSkyMaterial(std::shared_ptr<LightMaterial> material) {
  impl_->sky = material;
  impl_->flat_valid = false;
}
LightMaterial* Scene::GetSkyMaterial() { return impl_->sky.get(); }
uint32 Scene::GetCameraCount() { return (uint32)impl_->cameras.size(); }
Camera* Scene::GetCamera(uint32 index) {
  return index < impl_->cameras.size() ? &impl_->cameras[index] : nullptr;
}
void Scene::Optimize() {
  // The tree is (re)built while flattening; Optimize marks it as wanted (scene.cpp:216-225).
  impl_->tree_valid = true;
  impl_->flat_valid = false;
}

// ---------------------------------------------------------------- .scene parser
namespace {

struct Block {
  // every "key value..." line of one block, in order
  std::vector<std::pair<std::string, std::string>> lines;
  // last value wins, like repeated sscanf calls
  bool floats(const char* key, int n, float* out) const {
    bool any = false;
    for (auto& kv : lines) {
      if (kv.first != key) continue;
      const char* p = kv.second.c_str();
      float tmp[4];
      int got = 0;
      for (; got < n; got++) {
        char* end;
        tmp[got] = strtof(p, &end);
        if (end == p) break;
        p = end;
      }
      // sscanf stores the fields it converted before failing
      for (int i = 0; i < got; i++) out[i] = tmp[i];
      any = any || got > 0;
    }
    return any;
  }
  bool word(const char* key, std::string* out) const {
    bool any = false;
    for (auto& kv : lines) {
      if (kv.first != key && !(std::string(key) == "texture" && kv.first == "texture_scale"))
        continue;
      // " texture %s" also matches a "texture_scale 2" line and yields "_scale"
      // (SURVEY.md App. D); keep that quirk.
      std::string v = kv.second;
      if (kv.first == "texture_scale" && std::string(key) == "texture") v = "_scale";
      size_t a = v.find_first_not_of(" \t");
      if (a == std::string::npos) continue;
      size_t b = v.find_first_of(" \t\r", a);
      *out = v.substr(a, b == std::string::npos ? std::string::npos : b - a);
      any = true;
    }
    return any;
  }
};

Block read_block(std::ifstream& in) {
  Block b;
  std::string line;
  while (std::getline(in, line)) {
    if (line.find('}') != std::string::npos) break;
    size_t a = line.find_first_not_of(" \t");
    if (a == std::string::npos) continue;
    size_t e = line.find_first_of(" \t\r", a);
    std::string key = line.substr(a, e == std::string::npos ? std::string::npos : e - a);
    std::string rest = e == std::string::npos ? "" : line.substr(e);
    b.lines.emplace_back(key, rest);
  }
  return b;
}

typedef std::map<std::string, std::shared_ptr<DiffuseMaterial>> MaterialMap;

void assign_material(Object* o, const Block& b, MaterialMap& mats) {
  std::string name;
  if (o && b.word("material", &name) && mats.count(name)) o->SetMaterial(mats[name]);
}

}  // namespace

bool Scene::LoadScene(const std::string& filename) {
  std::ifstream in(filename, std::ios::in);
  if (!in.is_open()) {
    printf("Failed to read scene file %s.\n", filename.c_str());
    return false;
  }
  MaterialMap mats;
  std::string line;
  while (std::getline(in, line)) {
    if (!line.empty() && line[0] == '#') continue;
    char name[260] = {0};
    if (sscanf(line.c_str(), " material %259s", name) == 1) {
      Block b = read_block(in);
      float color[3] = {0, 0, 0}, emission[3] = {0, 0, 0};
      float metallic = 0, roughness = 0, index = 1, tex_scale = 1.0, frost = 0.0, refl = 0.1;
      float brdf_f = 0;
      b.floats("color", 3, color);
      b.floats("emission", 3, emission);
      b.floats("metallic", 1, &metallic);
      b.floats("roughness", 1, &roughness);
      b.floats("index", 1, &index);
      b.floats("texture_scale", 1, &tex_scale);
      b.floats("brdf", 1, &brdf_f);
      b.floats("frostiness", 1, &frost);
      b.floats("reflectivity", 1, &refl);
      int brdf = (int)brdf_f;
      std::string texture;
      b.word("texture", &texture);
      vector3 c(color[0], color[1], color[2]);
      std::shared_ptr<DiffuseMaterial> m;
      if (emission[0] || emission[1] || emission[2]) {
        m = std::make_shared<LightMaterial>(vector3(emission[0], emission[1], emission[2]));
      } else if (roughness) {
        m = std::make_shared<CeramicMaterial>(c, roughness);
      } else if (metallic) {
        if (metallic == 1.0)
          m = std::make_shared<MirrorMaterial>(c);
        else
          m = std::make_shared<MetalMaterial>(c, metallic);
      } else if (brdf == 1) {
        m = std::make_shared<LiquidMaterial>(c, index, refl);
      } else if (brdf == 2) {
        m = std::make_shared<GlassMaterial>(c, index, refl, frost);
      } else {
        m = std::make_shared<DiffuseMaterial>(c);
      }
      if (!texture.empty() && texture != "None") m->LoadDiffuseTexture(texture, tex_scale);
      mats[name] = m;
    } else if (line.find("sphere") != std::string::npos) {
      Block b = read_block(in);
      float pos[3] = {0, 0, 0}, radius = 0;
      b.floats("position", 3, pos);
      b.floats("radius", 1, &radius);
      assign_material(AddSphericalObject(vector3(pos[0], pos[1], pos[2]), radius), b, mats);
    } else if (line.find("camera") != std::string::npos) {
      Block b = read_block(in);
      Camera cam;
      float pos[3] = {cam.origin.x, cam.origin.y, cam.origin.z};
      float tgt[3] = {cam.target.x, cam.target.y, cam.target.z};
      b.floats("position", 3, pos);
      b.floats("target", 3, tgt);
      b.floats("fov", 1, &cam.fov_y);
      b.floats("aperture", 1, &cam.aperture_size);
      b.floats("focal_depth", 1, &cam.focal_depth);
      cam.origin = vector3(pos[0], pos[1], pos[2]);
      cam.target = vector3(tgt[0], tgt[1], tgt[2]);
      impl_->cameras.push_back(cam);
    } else if (line.find("sky") != std::string::npos) {
      Block b = read_block(in);
      std::string mname;
      if (b.word("material", &mname) && mats.count(mname))
        // the reference casts whatever material the block names (scene.cpp:382-385); the sky is
        // only ever read through the base-class record, so a non-light material keeps its class
        SetSkyMaterial(std::static_pointer_cast<LightMaterial>(mats[mname]));
    } else if (line.find("quad") != std::string::npos) {
      Block b = read_block(in);
      float pos[3] = {0, 0, 0}, nrm[3] = {0, 0, 0}, w = 0, h = 0;
      b.floats("position", 3, pos);
      b.floats("normal", 3, nrm);
      b.floats("width", 1, &w);
      b.floats("height", 1, &h);
      assign_material(AddQuadObject(vector3(pos[0], pos[1], pos[2]),
                                    vector3(nrm[0], nrm[1], nrm[2]), w, h), b, mats);
    } else if (line.find("cuboid") != std::string::npos) {
      Block b = read_block(in);
      float pos[3] = {0, 0, 0}, w = 0, h = 0, d = 0, rot[4] = {0, 0, 0, 0};
      b.floats("position", 3, pos);
      b.floats("width", 1, &w);
      b.floats("height", 1, &h);
      b.floats("depth", 1, &d);
      b.floats("rotation", 4, rot);
      CuboidObject* o = AddCuboidObject(vector3(pos[0], pos[1], pos[2]), w, h, d);
      assign_material(o, b, mats);
      o->Rotate(vector3(rot[0], rot[1], rot[2]), rot[3]);
    } else if (line.find("mesh") != std::string::npos) {
      Block b = read_block(in);
      float tr[3] = {0, 0, 0}, sc[3] = {1, 1, 1}, rot[4] = {0, 0, 0, 0};
      std::string file;
      b.word("file", &file);
      b.floats("translation", 3, tr);
      b.floats("scale", 3, sc);
      b.floats("rotation", 4, rot);
      if (!file.empty()) {
        assign_material(AddMeshObject(file, false, vector3(tr[0], tr[1], tr[2]),
                                      vector3(sc[0], sc[1], sc[2]),
                                      vector4(rot[0], rot[1], rot[2], rot[3])), b, mats);
      }
    }
  }
  Optimize();
  printf("Scene file %s loaded successfully.\n", filename.c_str());
  return true;
}

// ---------------------------------------------------------------- flatten
static uint32_t flatten_material(FlatArrays* F, std::map<Material*, uint32_t>* seen, Material* m,
                                 std::vector<Material*>* owner) {
  if (!m) return 0xFFFFFFFFu;
  auto it = seen->find(m);
  if (it != seen->end()) return it->second;
  MaterialImpl* mi = m->impl();
  fs_material r = mi->rec;
  r.texture = -1;
  if (!mi->texture.rgb.empty() && mi->texture.width && mi->texture.height) {
    fs_texture t;
    t.width = mi->texture.width;
    t.height = mi->texture.height;
    t.offset = F->texels.size();
    F->texels.insert(F->texels.end(), mi->texture.rgb.begin(), mi->texture.rgb.end());
    r.texture = (int32_t)F->textures.size();
    F->textures.push_back(t);
  }
  uint32_t idx = (uint32_t)F->materials.size();
  F->materials.push_back(r);
  owner->push_back(m);
  (*seen)[m] = idx;
  return idx;
}

const fs_scene_desc* Scene::Flatten() {
  SceneImpl* S = impl_.get();
  if (S->flat_valid) return &S->desc;
  FlatArrays& F = S->flat;
  F = FlatArrays();
  S->flat_material_owner.clear();
  std::map<Material*, uint32_t> seen;
  unsigned threads = std::thread::hardware_concurrency();
  if (const char* e = getenv("FS2_HOST_THREADS")) threads = (unsigned)atoi(e);
  if (threads == 0) threads = 1;
  std::vector<Box> boxes;
  for (auto& up : S->objects) {
    Object* o = up.get();
    ObjectImpl* oi = o->impl();
    fs_object r = oi->rec;
    r.material = flatten_material(&F, &seen, o->GetMaterial(), &S->flat_material_owner);
    put3(r.bmin, oi->box.mn);
    put3(r.bmax, oi->box.mx);
    boxes.push_back(oi->box);
    if (r.type == FS_OBJ_MESH) {
      uint32_t face_base = (uint32_t)F.faces.size();
      r.u[1] = face_base;
      r.u[2] = (oi->normals.size() ? 1u : 0u) | (oi->texcoords.size() ? 2u : 0u);
      r.u[3] = (uint32_t)oi->faces.size();
      F.faces.resize((size_t)face_base + oi->faces.size());
      parallel_for(oi->faces.size(), threads, [&](size_t i) {
        const MeshFaceRec& f = oi->faces[i];
        fs_face_attr fa;
        memset(&fa, 0, sizeof fa);
        for (int k = 0; k < 3; k++) {
          if (oi->normals.size() && f.n[k] < oi->normals.size()) put3(fa.n[k], oi->normals[f.n[k]]);
          if (oi->texcoords.size() && f.t[k] < oi->texcoords.size() / 2) {
            fa.uv[k][0] = oi->texcoords[2 * (size_t)f.t[k]];
            fa.uv[k][1] = oi->texcoords[2 * (size_t)f.t[k] + 1];
          }
        }
        F.faces[(size_t)face_base + i] = fa;
      });
      r.u[0] = oi->mesh_loaded ? build_mesh_tree(oi->verts, oi->faces, face_base, &F, threads)
                               : 0xFFFFFFFFu;
    }
    F.objects.push_back(r);
  }
  F.sky = flatten_material(&F, &seen, S->sky.get(), &S->flat_material_owner);
  F.scene_root = S->tree_valid ? build_scene_tree(boxes, &F) : -1;
  F.lut = normal_table();

  fs_scene_desc& d = S->desc;
  memset(&d, 0, sizeof d);
  d.abi_version = FS2_ABI_VERSION;
  d.n_nodes = (uint32_t)F.nodes.size();
  d.n_scene_refs = (uint32_t)F.scene_refs.size();
  d.n_tris = (uint32_t)F.tri_aux.size();
  d.n_faces = (uint32_t)F.faces.size();
  d.n_objects = (uint32_t)F.objects.size();
  d.n_materials = (uint32_t)F.materials.size();
  d.n_textures = (uint32_t)F.textures.size();
  d.n_texels = F.texels.size();
  d.scene_root = F.scene_root;
  d.sky_material = F.sky;
  d.n_lut = (uint32_t)(F.lut.size() / 3);
  d.nodes = F.nodes.data();
  d.scene_refs = F.scene_refs.data();
  d.tri_plane = F.tri_plane.data();
  d.tri_aux = F.tri_aux.data();
  d.faces = F.faces.data();
  d.objects = F.objects.data();
  d.materials = F.materials.data();
  d.textures = F.textures.data();
  d.texels = F.texels.data();
  d.lut = F.lut.data();
  S->flat_valid = true;
  S->flat_version++;
  return &d;
}

bool Scene::SaveBlob(const std::string& path) {
  return fs_blob_save(Flatten(), path.c_str()) == FS_OK;
}

}  // namespace base
