/*
 * fs2_blob.h — on-disk form of an fs_scene_desc (".fs2" blob), header-only.
 *
 * Used to move flattened scenes between the host flattener, the CPU oracle,
 * the compiled-reference harness and the tests.  Layout: fs_blob_header, then
 * the ten arrays of fs_scene_desc in declaration order, each padded to 16
 * bytes.  Little-endian, no compression.  numpy can read it with the offsets
 * in the header (see final-stage-path-tracer-2-0_b200/__init__.py).
 */
#ifndef FS2_BLOB_H_
#define FS2_BLOB_H_

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "fs2_cabi.h"

#define FS2_BLOB_MAGIC 0x31325346u /* "FS21" */

typedef struct fs_blob_header {
  uint32_t magic;
  uint32_t abi_version;
  uint32_t n_nodes, n_scene_refs, n_tris, n_faces, n_objects, n_materials,
      n_textures, n_lut;
  uint64_t n_texels;
  int32_t scene_root;
  uint32_t sky_material;
  uint64_t offset[10]; /* byte offsets of the arrays from the file start */
  uint64_t total_bytes;
} fs_blob_header;

static inline uint64_t fs_blob_pad16(uint64_t n) { return (n + 15u) & ~(uint64_t)15u; }

static inline void fs_blob_sizes(const fs_scene_desc* s, uint64_t sz[10]) {
  sz[0] = (uint64_t)s->n_nodes * sizeof(fs_node);
  sz[1] = (uint64_t)s->n_scene_refs * sizeof(uint32_t);
  sz[2] = (uint64_t)s->n_tris * 4 * sizeof(float);
  sz[3] = (uint64_t)s->n_tris * sizeof(fs_tri_aux);
  sz[4] = (uint64_t)s->n_faces * sizeof(fs_face_attr);
  sz[5] = (uint64_t)s->n_objects * sizeof(fs_object);
  sz[6] = (uint64_t)s->n_materials * sizeof(fs_material);
  sz[7] = (uint64_t)s->n_textures * sizeof(fs_texture);
  sz[8] = (uint64_t)s->n_texels * sizeof(float);
  sz[9] = (uint64_t)s->n_lut * 3 * sizeof(float);
}

static inline int fs_blob_save(const fs_scene_desc* s, const char* path) {
  const void* ptr[10] = {s->nodes,   s->scene_refs, s->tri_plane, s->tri_aux,  s->faces,
                         s->objects, s->materials,  s->textures,  s->texels,   s->lut};
  uint64_t sz[10];
  fs_blob_header h;
  static const char zeros[16] = {0};
  FILE* f;
  int i;
  memset(&h, 0, sizeof h);
  fs_blob_sizes(s, sz);
  h.magic = FS2_BLOB_MAGIC;
  h.abi_version = FS2_ABI_VERSION;
  h.n_nodes = s->n_nodes;
  h.n_scene_refs = s->n_scene_refs;
  h.n_tris = s->n_tris;
  h.n_faces = s->n_faces;
  h.n_objects = s->n_objects;
  h.n_materials = s->n_materials;
  h.n_textures = s->n_textures;
  h.n_lut = s->n_lut;
  h.n_texels = s->n_texels;
  h.scene_root = s->scene_root;
  h.sky_material = s->sky_material;
  {
    uint64_t off = fs_blob_pad16(sizeof h);
    for (i = 0; i < 10; i++) {
      h.offset[i] = off;
      off += fs_blob_pad16(sz[i]);
    }
    h.total_bytes = off;
  }
  f = fopen(path, "wb");
  if (!f) return FS_E_IO;
  fwrite(&h, sizeof h, 1, f);
  fwrite(zeros, 1, (size_t)(fs_blob_pad16(sizeof h) - sizeof h), f);
  for (i = 0; i < 10; i++) {
    if (sz[i]) fwrite(ptr[i], 1, (size_t)sz[i], f);
    fwrite(zeros, 1, (size_t)(fs_blob_pad16(sz[i]) - sz[i]), f);
  }
  if (fclose(f) != 0) return FS_E_IO;
  return FS_OK;
}

/* Loads a blob into one malloc'ed block (*storage, free() it) and points the
 * arrays of *s into it. */
static inline int fs_blob_load(const char* path, fs_scene_desc* s, void** storage) {
  fs_blob_header h;
  FILE* f = fopen(path, "rb");
  char* mem;
  long len;
  if (!f) return FS_E_IO;
  fseek(f, 0, SEEK_END);
  len = ftell(f);
  fseek(f, 0, SEEK_SET);
  if (len < (long)sizeof h) {
    fclose(f);
    return FS_E_IO;
  }
  mem = (char*)malloc((size_t)len + 16);
  if (!mem) {
    fclose(f);
    return FS_E_NOMEM;
  }
  if (fread(mem, 1, (size_t)len, f) != (size_t)len) {
    fclose(f);
    free(mem);
    return FS_E_IO;
  }
  fclose(f);
  memcpy(&h, mem, sizeof h);
  if (h.magic != FS2_BLOB_MAGIC || h.abi_version != FS2_ABI_VERSION ||
      h.total_bytes != (uint64_t)len) {
    free(mem);
    return FS_E_INVALID;
  }
  memset(s, 0, sizeof *s);
  s->abi_version = h.abi_version;
  s->n_nodes = h.n_nodes;
  s->n_scene_refs = h.n_scene_refs;
  s->n_tris = h.n_tris;
  s->n_faces = h.n_faces;
  s->n_objects = h.n_objects;
  s->n_materials = h.n_materials;
  s->n_textures = h.n_textures;
  s->n_lut = h.n_lut;
  s->n_texels = h.n_texels;
  s->scene_root = h.scene_root;
  s->sky_material = h.sky_material;
  s->nodes = (const fs_node*)(mem + h.offset[0]);
  s->scene_refs = (const uint32_t*)(mem + h.offset[1]);
  s->tri_plane = (const float*)(mem + h.offset[2]);
  s->tri_aux = (const fs_tri_aux*)(mem + h.offset[3]);
  s->faces = (const fs_face_attr*)(mem + h.offset[4]);
  s->objects = (const fs_object*)(mem + h.offset[5]);
  s->materials = (const fs_material*)(mem + h.offset[6]);
  s->textures = (const fs_texture*)(mem + h.offset[7]);
  s->texels = (const float*)(mem + h.offset[8]);
  s->lut = (const float*)(mem + h.offset[9]);
  *storage = mem;
  return FS_OK;
}

#endif /* FS2_BLOB_H_ */
