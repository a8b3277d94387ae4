// fs2_objects.cpp — host records of cameras, materials and objects: constructors only
// (bounds, planes, tangents, transformed mesh vertices), evaluated with the reference's
// float operations so that the flattened scene equals the reference's:
//   camera     reference source/camera.cpp:6-24
//   materials  reference source/material.cpp:26-31,55-95,139-148,172,192-198,223-228,252-255,
//              284-289,303-307
//   objects    reference source/object.cpp:18-21,38-57,75-95,115-124,167-211;
//              cube: source/math/volume.cpp:234-283,434-445
//   mesh       reference source/mesh.cpp:168-271
#include <stdio.h>

#include "fs2_impl.h"

namespace fs2h {

Xorshift& main_rng() {
  static Xorshift rng;
  return rng;
}

static std::vector<float> g_normal_table;
const std::vector<float>& normal_table() { return g_normal_table; }

// reference source/math/normal.cpp:9-27
static void fill_normal_table(int count) {
  Xorshift& rng = main_rng();
  g_normal_table.resize((size_t)count * 3);
  for (int i = 0; i < count; i++) {
    Vec3 axis(10, 10, 10);
    while (length(axis) > 1.0) {
      // The reference draws inside a constructor call, vector3(r(), r(), r()); the g++ build
      // that serves as the oracle evaluates those arguments right to left: z, y, x.
      float c = rng.range(-1.0f, 1.0f);
      float b = rng.range(-1.0f, 1.0f);
      float a = rng.range(-1.0f, 1.0f);
      axis = Vec3(a, b, c);
    }
    axis = normalize(axis);
    put3(&g_normal_table[3 * (size_t)i], axis);
  }
}

// cube::compute_planes, volume.cpp:234-247
static void cube_planes(const Vec3* v, float* f24) {
  static const int idx[6][3] = {{0, 1, 2}, {6, 5, 4}, {5, 6, 2}, {7, 4, 0}, {4, 5, 1}, {6, 7, 3}};
  for (int i = 0; i < 6; i++) {
    Plane p = plane_from(v[idx[i][0]], v[idx[i][1]], v[idx[i][2]]);
    f24[4 * i + 0] = p.x;
    f24[4 * i + 1] = p.y;
    f24[4 * i + 2] = p.z;
    f24[4 * i + 3] = p.w;
  }
}

}  // namespace fs2h

namespace base {

using namespace fs2h;

// ---------------------------------------------------------------- camera
Camera::Camera()
    : origin(0, 0, -200), target(0, 0, 0), fov_y(45), aperture_size(1.5), focal_depth(80),
      z_near(1.0f), z_far(10000.0f), fast_render_enabled(false) {}
Camera::Camera(const vector3& o, const vector3& t)
    : origin(o), target(t), fov_y(45), aperture_size(1.5), focal_depth(80), z_near(1.0f),
      z_far(10000.0f), fast_render_enabled(false) {}

// ---------------------------------------------------------------- materials
Material::Material(uint32_t type) : impl_(new MaterialImpl()) {
  memset(&impl_->rec, 0, sizeof impl_->rec);
  impl_->rec.type = type;
  impl_->rec.texture = -1;
  impl_->rec.id = (uint32_t)main_rng().next();  // material.cpp:26
}
Material::~Material() {}
uint32 Material::GetID() const { return impl_->rec.id; }
const fs_material& Material::Record() const { return impl_->rec; }

DiffuseMaterial::DiffuseMaterial(uint32_t type) : Material(type) {}
DiffuseMaterial::DiffuseMaterial() : Material(FS_MAT_DIFFUSE) {}
DiffuseMaterial::DiffuseMaterial(const vector3& d) : Material(FS_MAT_DIFFUSE) {
  put3(impl_->rec.diffuse, to3(d));
}
DiffuseMaterial::DiffuseMaterial(const std::string& filename, float32 tex_scale)
    : Material(FS_MAT_DIFFUSE) {
  LoadDiffuseTexture(filename, tex_scale);
}

static bool ends_with(const std::string& s, const char* ext) {
  size_t n = strlen(ext);
  return s.size() >= n && s.compare(s.size() - n, n, ext) == 0;
}

void DiffuseMaterial::LoadDiffuseTexture(const std::string& filename, float32 tex_scale) {
  impl_->rec.tex_scale = tex_scale;
  bool ok = false;
  if (ends_with(filename, ".bmp")) {
    printf("Loading bitmap file: %s.\n", filename.c_str());
    ok = load_bmp(filename, &impl_->texture);
  } else if (ends_with(filename, ".hdr")) {
    ok = load_hdr(filename, &impl_->texture);
  } else if (ends_with(filename, ".exr")) {
    // reference source/material.cpp:70-92 (LoadEXR, then rgb of every texel in file row order)
    std::string errors;
    ok = load_exr(filename, &impl_->texture, &errors);
    if (ok)
      printf("Loaded %s with dims: <%i, %i>.\n", filename.c_str(), (int)impl_->texture.width,
             (int)impl_->texture.height);
    else
      printf("Failed to load %s. Errors: %s.\n", filename.c_str(), errors.c_str());
  }
  if (!ok) impl_->texture = Image();
}

LightMaterial::LightMaterial() : DiffuseMaterial(FS_MAT_LIGHT) {
  put3(impl_->rec.emissive, Vec3(1, 1, 1));
  put3(impl_->rec.diffuse, Vec3(1, 1, 1));
}
LightMaterial::LightMaterial(const vector3& e) : DiffuseMaterial(FS_MAT_LIGHT) {
  put3(impl_->rec.emissive, to3(e));
  put3(impl_->rec.diffuse, Vec3(1, 1, 1));
}
MetalMaterial::MetalMaterial(const vector3& d, float32 roughness) : DiffuseMaterial(FS_MAT_METAL) {
  put3(impl_->rec.diffuse, to3(d));
  impl_->rec.roughness = roughness;
}
MirrorMaterial::MirrorMaterial(const vector3& d) : DiffuseMaterial(FS_MAT_MIRROR) {
  put3(impl_->rec.diffuse, to3(d));
}
GlassMaterial::GlassMaterial(const vector3 d, float32 index, float32 reflectivity, float32 frost)
    : DiffuseMaterial(FS_MAT_GLASS) {
  put3(impl_->rec.diffuse, to3(d));
  impl_->rec.index = index;
  impl_->rec.frost = frost;
  impl_->rec.reflectivity = reflectivity;
}
LiquidMaterial::LiquidMaterial(const vector3 d, float32 index, float32 reflectivity)
    : DiffuseMaterial(FS_MAT_LIQUID) {
  put3(impl_->rec.diffuse, to3(d));
  impl_->rec.index = index;
  impl_->rec.reflectivity = reflectivity;
}
CeramicMaterial::CeramicMaterial(uint32_t type, const vector3& d, float32 shininess)
    : DiffuseMaterial(type) {
  put3(impl_->rec.diffuse, to3(d));
  impl_->rec.roughness = shininess;
}
CeramicMaterial::CeramicMaterial(const vector3& d, float32 shininess)
    : CeramicMaterial(FS_MAT_CERAMIC, d, shininess) {}
GlowMaterial::GlowMaterial(const vector3& d, const vector3& glow, float32 shininess)
    : CeramicMaterial(FS_MAT_GLOW, d, shininess) {
  put3(impl_->rec.emissive, to3(glow));
}
FogMaterial::FogMaterial(const vector3 d, float32 density) : DiffuseMaterial(FS_MAT_FOG) {
  put3(impl_->rec.diffuse, to3(d));
  impl_->rec.density = density * 1000.0f;  // material.cpp:306
}

void InitializeMaterials() { fill_normal_table(32 * 1024); }

// ---------------------------------------------------------------- objects
Object::Object() : impl_(new ObjectImpl()) { memset(&impl_->rec, 0, sizeof impl_->rec); }
Object::~Object() {}
void Object::SetMaterial(std::shared_ptr<Material> m) { material_ = m; }
void Object::GetBounds(vector3* bmin, vector3* bmax) const {
  const Box& b = impl_->box;
  if (bmin) *bmin = vector3(b.mn.x, b.mn.y, b.mn.z);
  if (bmax) *bmax = vector3(b.mx.x, b.mx.y, b.mx.z);
}

SphericalObject::SphericalObject(const vector3& origin, float32 radius) {
  fs_object& r = impl_->rec;
  r.type = FS_OBJ_SPHERE;
  Vec3 o = to3(origin);
  put3(r.f, o);
  r.f[3] = radius;
  impl_->box.add(o + Vec3(radius, radius, radius));
  impl_->box.add(o - Vec3(radius, radius, radius));
}

PlanarObject::PlanarObject(const plane& data) {
  fs_object& r = impl_->rec;
  r.type = FS_OBJ_PLANE;
  r.f[0] = data.x;
  r.f[1] = data.y;
  r.f[2] = data.z;
  r.f[3] = data.w;
  Vec3 n(data.x, data.y, data.z), up(0, 1, 0);
  float distance = -data.w;
  Vec3 p = n * distance;
  Vec3 right = cross(n, up);
  Vec3 forward = cross(n, right);
  Box& b = impl_->box;
  b.add(p + right * 1000);
  b.add(p - right * 1000);
  b.add(p + forward * 1000);
  b.add(p - forward * 1000);
  b.add(p + n * kEpsilon);
  b.add(p - n * kEpsilon);
}

DiscObject::DiscObject(const vector3& origin, const vector3& normal, float32 radius) {
  fs_object& r = impl_->rec;
  r.type = FS_OBJ_DISC;
  Vec3 o = to3(origin), n = to3(normal), up(0, 1, 0);
  Plane pl = plane_from(n, o);
  r.f[0] = pl.x;
  r.f[1] = pl.y;
  r.f[2] = pl.z;
  r.f[3] = pl.w;
  put3(r.f + 4, o);
  r.f[7] = radius;
  Vec3 right = cross(n, up);
  Vec3 forward = cross(n, right);
  Box& b = impl_->box;
  b.add(o + right * radius);
  b.add(o - right * radius);
  b.add(o + forward * radius);
  b.add(o - forward * radius);
  b.add(o + n * kEpsilon);
  b.add(o - n * kEpsilon);
}

CuboidObject::CuboidObject(const vector3& origin, float32 width, float32 height, float32 depth) {
  fs_object& r = impl_->rec;
  r.type = FS_OBJ_CUBOID;
  Vec3 o = to3(origin);
  Vec3 half((float)(width * 0.5), (float)(height * 0.5), (float)(depth * 0.5));
  Box t;
  t.add(o - half);
  t.add(o + half);
  // cube::operator=(bounds), volume.cpp:262-283
  Vec3 mn = t.mn, mx = t.mx;
  Vec3 v[8] = {mn,
               Vec3(mx.x, mn.y, mn.z),
               Vec3(mx.x, mn.y, mx.z),
               Vec3(mn.x, mn.y, mx.z),
               Vec3(mn.x, mx.y, mn.z),
               Vec3(mx.x, mx.y, mn.z),
               Vec3(mx.x, mx.y, mx.z),
               Vec3(mn.x, mx.y, mx.z)};
  for (int i = 0; i < 8; i++) impl_->cube_verts[i] = v[i];
  impl_->cube_box = t;
  impl_->box = t;
  cube_planes(impl_->cube_verts, r.f);
}

// cube::rotate, volume.cpp:434-445
void CuboidObject::Rotate(const vector3& axis, float32 angle) {
  Vec3 c = impl_->cube_box.center();
  impl_->cube_box.clear();
  for (int i = 0; i < 8; i++) {
    Vec3 t = impl_->cube_verts[i] - c;
    impl_->cube_verts[i] = rotate(t, angle, to3(axis)) + c;
    impl_->cube_box.add(impl_->cube_verts[i]);
  }
  impl_->box = impl_->cube_box;
  cube_planes(impl_->cube_verts, impl_->rec.f);
}

QuadObject::QuadObject(const vector3& origin, const vector3& normal, float32 width,
                       float32 height) {
  fs_object& r = impl_->rec;
  r.type = FS_OBJ_QUAD;
  Vec3 o = to3(origin), nrm = to3(normal), up(0, 1, 0);
  Vec3 n = normalize(nrm);
  Plane pl = plane_from(n, o);
  float half_w = width * 0.5, half_h = height * 0.5;
  Vec3 bitangent = cross(n, up);
  Vec3 tangent = cross(n, bitangent);
  r.f[0] = pl.x;
  r.f[1] = pl.y;
  r.f[2] = pl.z;
  r.f[3] = pl.w;
  put3(r.f + 4, o);
  put3(r.f + 7, tangent);
  put3(r.f + 10, bitangent);
  r.f[13] = half_w;
  r.f[14] = half_h;
  Box& b = impl_->box;
  b.add(o + bitangent * half_w);
  b.add(o - bitangent * half_w);
  b.add(o + tangent * half_h);
  b.add(o - tangent * half_h);
  b.add(o + nrm * kEpsilon);
  b.add(o - nrm * kEpsilon);
}

QuadObject::QuadObject(const vector3& position, const vector3& uu, const vector3& vv) {
  fs_object& r = impl_->rec;
  r.type = FS_OBJ_QUAD;
  Vec3 p = to3(position), u = to3(uu), v = to3(vv);
  Vec3 n = normalize(cross(u, v));
  Plane pl = plane_from(n, p);
  Vec3 o = p + u * 0.5 + v * 0.5;
  float half_w = length(u) * 0.5, half_h = length(v) * 0.5;
  Vec3 bitangent = normalize(u), tangent = normalize(v);
  r.f[0] = pl.x;
  r.f[1] = pl.y;
  r.f[2] = pl.z;
  r.f[3] = pl.w;
  put3(r.f + 4, o);
  put3(r.f + 7, tangent);
  put3(r.f + 10, bitangent);
  r.f[13] = half_w;
  r.f[14] = half_h;
  Box& b = impl_->box;
  b.add(o + n * kEpsilon);
  b.add(o - n * kEpsilon);
  b.add(o + bitangent * half_w);
  b.add(o - bitangent * half_w);
  b.add(o + tangent * half_h);
  b.add(o - tangent * half_h);
}

// reference source/mesh.cpp:168-271 (the octree itself is built when the scene flattens)
MeshObject::MeshObject(const std::string& filename, bool invert_normals,
                       const vector3& translation, const vector3& scale,
                       const vector4& rotation) {
  impl_->rec.type = FS_OBJ_MESH;
  ObjData obj;
  std::string error;
  if (!load_obj(filename, &obj, &error)) {
    printf("Error loading obj file %s: %s.\n", filename.c_str(), error.c_str());
    return;
  }
  bool transform_needed = false;
  Mat4 t = Mat4::identity(), s = Mat4::identity(), rm = Mat4::identity();
  if (translation.x || translation.y || translation.z) {
    transform_needed = true;
    t = Mat4::translation(translation.x, translation.y, translation.z);
  }
  if (scale.x || scale.y || scale.z) {
    transform_needed = true;
    s = Mat4::scale(scale.x, scale.y, scale.z);
  }
  if (rotation.x || rotation.y || rotation.z) {
    transform_needed = true;
    rm = Mat4::rotation(rotation.w, Vec3(rotation.x, rotation.y, rotation.z));
  }
  if (transform_needed) {  // sic: without any transform the reference keeps no vertices
    Mat4 m = t * rm * s;
    impl_->verts.reserve(obj.positions.size() / 3);
    for (size_t i = 0; i + 2 < obj.positions.size(); i += 3) {
      Vec3 v = m.transform_point(Vec3(obj.positions[i], obj.positions[i + 1], obj.positions[i + 2]));
      impl_->box.add(v);
      impl_->verts.push_back(v);
    }
  }
  impl_->normals.reserve(obj.normals.size() / 3);
  for (size_t i = 0; i + 2 < obj.normals.size(); i += 3) {
    Vec3 n = normalize(Vec3(obj.normals[i], obj.normals[i + 1], obj.normals[i + 2]));
    if (invert_normals) n = n * -1.0;
    impl_->normals.push_back(n);
  }
  impl_->texcoords.swap(obj.texcoords);
  // reversed winding (mesh.cpp:250-260)
  impl_->faces.reserve(obj.corners.size() / 9);
  for (size_t j = 0; j + 8 < obj.corners.size(); j += 9) {
    MeshFaceRec f;
    for (int k = 0; k < 3; k++) {
      const int* c = &obj.corners[j + 3 * (2 - k)];
      f.v[k] = (uint32_t)c[0];
      f.n[k] = (uint32_t)c[1];
      f.t[k] = (uint32_t)c[2];
    }
    impl_->faces.push_back(f);
  }
  impl_->mesh_loaded = true;
}

}  // namespace base
