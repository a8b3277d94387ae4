// Builds the "@api" golden scene (oracle/ref/harness.cpp: build_api_scene) through the host
// mirror's C++ API — the same calls, in the same order, a user of the reference would write
// against source/scene.h / source/material.h — and writes the flattened blob.
//   api_scene TEXTURE.bmp OUT.fs2
#include <stdio.h>
#include <stdlib.h>

#include "fs2_host.h"

using namespace base;

static plane make_plane(vector3 n, vector3 p) {
  return plane(n.x, n.y, n.z, -(n.x * p.x + n.y * p.y + n.z * p.z));
}

int main(int argc, char** argv) {
  if (argc < 3) return 64;
  Scene scene_storage;
  Scene* scene = &scene_storage;
  auto sky = std::make_shared<LightMaterial>(vector3(0.35f, 0.45f, 0.6f));
  scene->SetSkyMaterial(sky);
  auto floor_mat = std::make_shared<DiffuseMaterial>(vector3(0.7f, 0.7f, 0.65f));
  floor_mat->LoadDiffuseTexture(argv[1], 0.05f);
  auto fog = std::make_shared<FogMaterial>(vector3(0.8f, 0.85f, 0.9f), 0.6f);
  auto glow = std::make_shared<GlowMaterial>(vector3(0.3f, 0.2f, 0.8f), vector3(0.4f, 0.1f, 0.9f), 0.7f);
  auto metal = std::make_shared<MetalMaterial>(vector3(0.9f, 0.6f, 0.3f), 0.97f);
  auto ceramic = std::make_shared<CeramicMaterial>(vector3(0.8f, 0.3f, 0.2f), 0.4f);
  auto lamp = std::make_shared<LightMaterial>(vector3(14.0f, 13.0f, 11.0f));
  auto glass = std::make_shared<GlassMaterial>(vector3(0.95f, 0.97f, 1.0f), 1.0f, 0.05f, 0.35f);
  auto diffuse = std::make_shared<DiffuseMaterial>(vector3(0.2f, 0.6f, 0.3f));

  // normals are passed pre-normalised (the harness calls vector3::normalize, which is the
  // reference's approximate one; the values below are its results).
  if (argc < 9) return 64;
  vector3 n_floor((float)atof(argv[3]), (float)atof(argv[4]), (float)atof(argv[5]));
  vector3 n_disc((float)atof(argv[6]), (float)atof(argv[7]), (float)atof(argv[8]));
  scene->AddPlanarObject(make_plane(n_floor, vector3(0, 0, 0)))->SetMaterial(floor_mat);
  scene->AddDiscObject(vector3(-14.0f, 6.0f, 10.0f), n_disc, 6.0f)->SetMaterial(metal);
  CuboidObject* fog_box = scene->AddCuboidObject(vector3(4.0f, 7.0f, -8.0f), 14.0f, 12.0f, 10.0f);
  fog_box->Rotate(vector3(0, 1, 0), 0.35f);
  fog_box->SetMaterial(fog);
  scene->AddSphericalObject(vector3(4.0f, 6.0f, -8.0f), 3.0f)->SetMaterial(glow);
  scene->AddSphericalObject(vector3(-6.0f, 4.0f, -12.0f), 4.0f)->SetMaterial(ceramic);
  scene->AddSphericalObject(vector3(14.0f, 5.0f, 2.0f), 5.0f)->SetMaterial(glass);
  scene->AddSphericalObject(vector3(-2.0f, 32.0f, -6.0f), 5.0f)->SetMaterial(lamp);
  scene->AddQuadObject(vector3(-20.0f, 14.0f, 16.0f), vector3(12.0f, 0.0f, 3.0f),
                       vector3(0.0f, -10.0f, 0.5f))
      ->SetMaterial(diffuse);
  scene->Optimize();
  InitializeMaterials();
  return scene->SaveBlob(argv[2]) ? 0 : 1;
}
