// fs2_assets.cpp — asset readers of the host side: Wavefront .obj geometry, 24-bit .bmp
// and Radiance .hdr textures.  Written for this project (the reference uses the vendored
// tinyobjloader v1.2.2 and its own bitmap.cpp; neither is copied): the readers produce the
// arrays MeshObject's constructor consumes (reference source/mesh.cpp:168-271) and the
// float RGB texel layout of reference source/bitmap.cpp:84-102 (BGR swap, /255, rows in
// file order).
#include "fs2_assets.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>

namespace fs2h {

namespace {

bool read_file(const std::string& path, std::vector<char>* data) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  if (n < 0 || n > (1L << 34)) {  // a directory opens and reports LONG_MAX; no asset is 16 GB
    fclose(f);
    return false;
  }
  data->resize((size_t)n + 1);
  size_t got = n ? fread(data->data(), 1, (size_t)n, f) : 0;
  fclose(f);
  (*data)[got] = 0;
  data->resize(got + 1);
  return true;
}

inline const char* skip_space(const char* p) {
  while (*p == ' ' || *p == '\t') p++;
  return p;
}

inline bool parse_float(const char*& p, float* out) {
  p = skip_space(p);
  char* end;
  double v = strtod(p, &end);
  if (end == p) return false;
  *out = (float)v;
  p = end;
  return true;
}

// OBJ indices are 1-based, negative = relative to the end; missing = -1.
inline int fix_index(long idx, size_t n) {
  if (idx > 0) return (int)(idx - 1);
  if (idx == 0) return -1;
  return (int)((long)n + idx);
}

struct Corner {
  int v, t, n;
};

bool parse_corner(const char*& p, size_t nv, size_t nt, size_t nn, Corner* c) {
  p = skip_space(p);
  if (*p == 0 || *p == '\r' || *p == '\n') return false;
  c->v = c->t = c->n = -1;
  char* end;
  long a = strtol(p, &end, 10);
  if (end == p) return false;
  c->v = fix_index(a, nv);
  p = end;
  if (*p != '/') return true;
  p++;
  if (*p == '/') {  // v//vn
    p++;
    long b = strtol(p, &end, 10);
    if (end != p) c->n = fix_index(b, nn);
    p = end;
    return true;
  }
  long b = strtol(p, &end, 10);
  if (end != p) c->t = fix_index(b, nt);
  p = end;
  if (*p != '/') return true;
  p++;
  long d = strtol(p, &end, 10);
  if (end != p) c->n = fix_index(d, nn);
  p = end;
  return true;
}

}  // namespace

bool load_obj(const std::string& path, ObjData* out, std::string* error) {
  std::vector<char> data;
  if (!read_file(path, &data)) {
    if (error) *error = "cannot open " + path;
    return false;
  }
  out->positions.clear();
  out->normals.clear();
  out->texcoords.clear();
  out->corners.clear();
  std::vector<Corner> poly;
  const char* p = data.data();
  while (*p) {
    const char* line = skip_space(p);
    const char* eol = line;
    while (*eol && *eol != '\n') eol++;
    if (line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) {
      const char* q = line + 2;
      float v[3] = {0, 0, 0};
      parse_float(q, &v[0]) && parse_float(q, &v[1]) && parse_float(q, &v[2]);
      out->positions.insert(out->positions.end(), v, v + 3);
    } else if (line[0] == 'v' && line[1] == 'n' && (line[2] == ' ' || line[2] == '\t')) {
      const char* q = line + 3;
      float v[3] = {0, 0, 0};
      parse_float(q, &v[0]) && parse_float(q, &v[1]) && parse_float(q, &v[2]);
      out->normals.insert(out->normals.end(), v, v + 3);
    } else if (line[0] == 'v' && line[1] == 't' && (line[2] == ' ' || line[2] == '\t')) {
      const char* q = line + 3;
      float v[2] = {0, 0};
      parse_float(q, &v[0]) && parse_float(q, &v[1]);
      out->texcoords.insert(out->texcoords.end(), v, v + 2);
    } else if (line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) {
      const char* q = line + 2;
      poly.clear();
      Corner c;
      while (q < eol && parse_corner(q, out->positions.size() / 3, out->texcoords.size() / 2,
                                     out->normals.size() / 3, &c))
        poly.push_back(c);
      // fan triangulation: (0, k-1, k)
      for (size_t k = 2; k < poly.size(); k++) {
        const Corner tri[3] = {poly[0], poly[k - 1], poly[k]};
        for (const Corner& t : tri) {
          out->corners.push_back(t.v);
          out->corners.push_back(t.n);
          out->corners.push_back(t.t);
        }
      }
    }
    p = *eol ? eol + 1 : eol;
  }
  for (size_t i = 0; i < out->corners.size(); i += 3) {
    int v = out->corners[i];
    if (v < 0 || (size_t)v >= out->positions.size() / 3) {
      if (error) *error = path + ": face references a vertex that does not exist";
      return false;
    }
  }
  return true;
}

#pragma pack(push, 1)
struct BmpFileHeader {
  uint16_t type;
  uint32_t size;
  uint16_t r1, r2;
  uint32_t offset;
};
struct BmpInfoHeader {
  uint32_t size;
  int32_t width, height;
  uint16_t planes, bit_count;
  uint32_t compression, image_size;
  int32_t xppm, yppm;
  uint32_t used, important;
};
#pragma pack(pop)

bool load_bmp(const std::string& path, Image* out) {
  std::vector<char> data;
  if (!read_file(path, &data)) return false;
  size_t n = data.size() - 1;
  if (n < sizeof(BmpFileHeader) + sizeof(BmpInfoHeader)) return false;
  BmpInfoHeader ih;
  memcpy(&ih, data.data() + sizeof(BmpFileHeader), sizeof ih);
  if (ih.bit_count != 24 || ih.width <= 0 || ih.height <= 0) return false;
  // like the reference, pixel rows are read right after the two headers
  size_t w = (size_t)ih.width, h = (size_t)ih.height;
  size_t stride = (w * 3 + 3) & ~(size_t)3;
  const unsigned char* px = (const unsigned char*)data.data() + sizeof(BmpFileHeader) + sizeof ih;
  if (sizeof(BmpFileHeader) + sizeof ih + (h - 1) * stride + w * 3 > n) return false;
  out->width = (uint32_t)w;
  out->height = (uint32_t)h;
  out->rgb.resize(w * h * 3);
  for (size_t i = 0; i < h; i++) {
    const unsigned char* row = px + i * stride;
    for (size_t j = 0; j < w; j++) {
      out->rgb[(i * w + j) * 3 + 0] = row[j * 3 + 2] / 255.0f;
      out->rgb[(i * w + j) * 3 + 1] = row[j * 3 + 1] / 255.0f;
      out->rgb[(i * w + j) * 3 + 2] = row[j * 3 + 0] / 255.0f;
    }
  }
  return true;
}

// Radiance RGBE (.hdr): an extension over the reference (which reads .bmp and .exr only;
// BASELINE.json asks for .hdr environments).  Flat and new-style RLE scanlines.
bool load_hdr(const std::string& path, Image* out) {
  std::vector<char> data;
  if (!read_file(path, &data)) return false;
  const unsigned char* p = (const unsigned char*)data.data();
  const unsigned char* end = p + data.size() - 1;
  if (end - p < 10 || (memcmp(p, "#?RADIANCE", 10) != 0 && memcmp(p, "#?RGBE", 6) != 0)) return false;
  // header lines until the empty line, then the resolution line
  const char* q = (const char*)p;
  const char* res = strstr(q, "\n\n");
  if (!res) return false;
  res += 2;
  int w = 0, h = 0;
  if (sscanf(res, "-Y %d +X %d", &h, &w) != 2 || w <= 0 || h <= 0) return false;
  const char* nl = strchr(res, '\n');
  if (!nl) return false;
  p = (const unsigned char*)nl + 1;
  // the header is untrusted: every scanline takes at least 4 bytes of file (an RLE line starts
  // with a 4-byte marker, a flat one is 4*w), so a size the file cannot hold is refused before
  // anything is allocated
  const size_t left = (size_t)(end - p);
  if ((size_t)h > left / 4 || (size_t)w > (size_t)1 << 20 || (size_t)w * (size_t)h > (size_t)1 << 28)
    return false;
  if ((w < 8 || w >= 32768) && (size_t)w * 4 * (size_t)h > left) return false;  // flat lines only
  std::vector<unsigned char> scan((size_t)w * 4);
  out->width = (uint32_t)w;
  out->height = (uint32_t)h;
  out->rgb.assign((size_t)w * h * 3, 0.f);
  for (int y = 0; y < h; y++) {
    if (end - p >= 4 && p[0] == 2 && p[1] == 2 && ((p[2] << 8) | p[3]) == w && w >= 8 && w < 32768) {
      p += 4;
      for (int c = 0; c < 4; c++) {
        int x = 0;
        while (x < w) {
          if (p >= end) return false;
          int count = *p++;
          if (count > 128) {
            count -= 128;
            if (p >= end || x + count > w) return false;
            unsigned char v = *p++;
            while (count--) scan[(size_t)(x++) * 4 + c] = v;
          } else {
            if (end - p < count || x + count > w || count == 0) return false;
            while (count--) scan[(size_t)(x++) * 4 + c] = *p++;
          }
        }
      }
    } else {
      if (end - p < (long)w * 4) return false;
      memcpy(scan.data(), p, (size_t)w * 4);
      p += (size_t)w * 4;
    }
    for (int x = 0; x < w; x++) {
      const unsigned char* e = &scan[(size_t)x * 4];
      float f = e[3] ? ldexpf(1.0f, (int)e[3] - (128 + 8)) : 0.0f;
      float* d = &out->rgb[((size_t)y * w + x) * 3];
      d[0] = e[0] * f;
      d[1] = e[1] * f;
      d[2] = e[2] * f;
    }
  }
  return true;
}

}  // namespace fs2h
