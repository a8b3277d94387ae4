// fs2_assets.cpp — asset readers of the host side: Wavefront .obj geometry, 24-bit .bmp
// and Radiance .hdr textures.  Written for this project (the reference uses the vendored
// tinyobjloader v1.2.2 and its own bitmap.cpp; neither is copied): the readers produce the
// arrays MeshObject's constructor consumes (reference source/mesh.cpp:168-271) and the
// float RGB texel layout of reference source/bitmap.cpp:84-102 (BGR swap, /255, rows in
// file order).
#include "fs2_assets.h"

#include <fcntl.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <memory>
#include <new>
#include <string>
#include <thread>
#include <vector>

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

// What one stretch of lines contributes.  Corner indices that are positive or missing are final;
// negative (relative) ones are stored relative to this stretch's own counts and listed in
// `relative`, to be completed with the counts of everything in front of the stretch.
struct ObjChunk {
  std::vector<float> positions, normals, texcoords;
  std::vector<int> corners;           // v, n, t per corner
  std::vector<size_t> relative;       // positions in `corners` that still lack their base
};

// OBJ indices are 1-based, negative = relative to the end, missing = -1; `relative` is set when
// the result still needs the count in front of the chunk added.
inline int local_index(long idx, size_t n_local, bool* relative) {
  *relative = idx < 0;
  if (idx > 0) return (int)(idx - 1);
  if (idx == 0) return -1;
  return (int)((long)n_local + idx);
}

bool parse_corner_local(const char*& p, const ObjChunk& c, int out[3], bool rel[3]) {
  p = skip_space(p);
  if (*p == 0 || *p == '\r' || *p == '\n') return false;
  out[0] = out[1] = out[2] = -1;  // v, n, t
  rel[0] = rel[1] = rel[2] = false;
  char* end;
  long a = strtol(p, &end, 10);
  if (end == p) return false;
  out[0] = local_index(a, c.positions.size() / 3, &rel[0]);
  p = end;
  if (*p != '/') return true;
  p++;
  if (*p == '/') {  // v//vn
    p++;
    long b = strtol(p, &end, 10);
    if (end != p) out[1] = local_index(b, c.normals.size() / 3, &rel[1]);
    p = end;
    return true;
  }
  long b = strtol(p, &end, 10);
  if (end != p) out[2] = local_index(b, c.texcoords.size() / 2, &rel[2]);
  p = end;
  if (*p != '/') return true;
  p++;
  long d = strtol(p, &end, 10);
  if (end != p) out[1] = local_index(d, c.normals.size() / 3, &rel[1]);
  p = end;
  return true;
}

// Parses the lines that start in [begin, stop).  The buffer behind `stop` stays readable (it is
// the whole file, NUL-terminated), so a number parser that runs over a line end sees exactly what
// it would see in a front-to-back pass.
void parse_obj_lines(const char* begin, const char* stop, ObjChunk* c) {
  struct Poly {
    int idx[3];
    bool rel[3];
  };
  std::vector<Poly> poly;
  const char* p = begin;
  while (p < stop && *p) {
    const char* line = skip_space(p);
    const char* eol = line;
    while (*eol && *eol != '\n') eol++;
    if (line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) {
      const char* q = line + 2;
      float v[3] = {0, 0, 0};
      parse_float(q, &v[0]) && parse_float(q, &v[1]) && parse_float(q, &v[2]);
      c->positions.insert(c->positions.end(), v, v + 3);
    } else if (line[0] == 'v' && line[1] == 'n' && (line[2] == ' ' || line[2] == '\t')) {
      const char* q = line + 3;
      float v[3] = {0, 0, 0};
      parse_float(q, &v[0]) && parse_float(q, &v[1]) && parse_float(q, &v[2]);
      c->normals.insert(c->normals.end(), v, v + 3);
    } else if (line[0] == 'v' && line[1] == 't' && (line[2] == ' ' || line[2] == '\t')) {
      const char* q = line + 3;
      float v[2] = {0, 0};
      parse_float(q, &v[0]) && parse_float(q, &v[1]);
      c->texcoords.insert(c->texcoords.end(), v, v + 2);
    } else if (line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) {
      const char* q = line + 2;
      poly.clear();
      Poly k;
      while (q < eol && parse_corner_local(q, *c, k.idx, k.rel)) poly.push_back(k);
      // fan triangulation: (0, k-1, k)
      for (size_t j = 2; j < poly.size(); j++) {
        const Poly* tri[3] = {&poly[0], &poly[j - 1], &poly[j]};
        for (const Poly* t : tri)
          for (int f = 0; f < 3; f++) {
            if (t->rel[f]) c->relative.push_back(c->corners.size());
            c->corners.push_back(t->idx[f]);
          }
      }
    }
    p = *eol ? eol + 1 : eol;
  }
}

unsigned obj_threads(size_t bytes) {
  unsigned t = std::thread::hardware_concurrency();
  if (const char* e = getenv("FS2_HOST_THREADS")) t = (unsigned)atoi(e);
  if (t > 32) t = 32;
  const size_t by_size = bytes >> 16;  // at least 64 KB of text per thread
  if (t > by_size) t = (unsigned)by_size;
  return t < 1 ? 1 : t;
}

}  // namespace

// The file is cut at line ends into one stretch per host thread; every stretch is parsed on its
// own, then the arrays are joined in file order and the relative indices completed — the same
// arrays a front-to-back pass produces (tests/test_host_flatten.py compares the two).
namespace {
// The whole file plus a terminating NUL in a buffer that is not cleared first (a std::vector
// would write 160 MB of zeros for a 4 M-triangle mesh before reading over them).
std::unique_ptr<char[]> read_text(const std::string& path, size_t* size) {
  int fd = open(path.c_str(), O_RDONLY);
  if (fd < 0) return nullptr;
  struct stat st;
  if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode) || st.st_size < 0 || st.st_size > (off_t)1 << 34) {
    close(fd);  // directories, devices, absurd sizes: not an asset
    return nullptr;
  }
  const size_t n = (size_t)st.st_size;
  std::unique_ptr<char[]> buf(new (std::nothrow) char[n + 1]);
  if (!buf) {
    close(fd);
    return nullptr;
  }
  // one reader: several threads touching the fresh pages at once were measured slower (0.87 vs
  // 0.65 s for the whole load of a 163 MB file) than one front-to-back read
  size_t have = 0;
  while (have < n) {
    ssize_t r = read(fd, buf.get() + have, n - have);
    if (r <= 0) break;
    have += (size_t)r;
  }
  close(fd);
  buf[have] = 0;
  *size = have;
  return buf;
}
}  // namespace

bool load_obj(const std::string& path, ObjData* out, std::string* error) {
  size_t file_size = 0;
  std::unique_ptr<char[]> data = read_text(path, &file_size);
  if (!data) {
    if (error) *error = "cannot open " + path;
    return false;
  }
  out->positions.clear();
  out->normals.clear();
  out->texcoords.clear();
  out->corners.clear();
  const char* base = data.get();
  const size_t len = strlen(base);  // a front-to-back pass stops at the first NUL byte
  // more stretches than threads, handed out from a queue: vertex lines (three strtod calls) and
  // face lines cost differently and sit in different parts of the file
  const unsigned threads = obj_threads(len);
  const unsigned nt = threads > 1 ? threads * 8 : 1;
  std::vector<const char*> cut(nt + 1);
  cut[0] = base;
  cut[nt] = base + len;
  for (unsigned i = 1; i < nt; i++) {
    const char* q = base + len / nt * i;
    if (q < cut[i - 1]) q = cut[i - 1];
    while (*q && *q != '\n') q++;
    cut[i] = *q ? q + 1 : q;
  }
  std::vector<ObjChunk> chunks(nt);
  auto run = [&](auto&& body) {
    if (threads == 1) {
      for (unsigned i = 0; i < nt; i++) body(i);
      return;
    }
    std::atomic<unsigned> next(0);
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < threads; t++)
      pool.emplace_back([&]() {
        for (unsigned i; (i = next.fetch_add(1)) < nt;) body(i);
      });
    for (std::thread& t : pool) t.join();
  };
  run([&](unsigned i) { parse_obj_lines(cut[i], cut[i + 1], &chunks[i]); });
  // counts in front of every stretch, then the joined arrays
  std::vector<size_t> pv(nt + 1, 0), pn(nt + 1, 0), pt(nt + 1, 0), pc(nt + 1, 0);
  for (unsigned i = 0; i < nt; i++) {
    pv[i + 1] = pv[i] + chunks[i].positions.size();
    pn[i + 1] = pn[i] + chunks[i].normals.size();
    pt[i + 1] = pt[i] + chunks[i].texcoords.size();
    pc[i + 1] = pc[i] + chunks[i].corners.size();
  }
  out->positions.resize(pv[nt]);
  out->normals.resize(pn[nt]);
  out->texcoords.resize(pt[nt]);
  out->corners.resize(pc[nt]);
  auto join = [&](unsigned i) {
    ObjChunk& c = chunks[i];
    // corners hold v, n, t in turn: position % 3 says which count completes a relative index
    const long add[3] = {(long)(pv[i] / 3), (long)(pn[i] / 3), (long)(pt[i] / 2)};
    for (size_t k : c.relative) c.corners[k] = (int)((long)c.corners[k] + add[k % 3]);
    if (!c.positions.empty()) memcpy(&out->positions[pv[i]], c.positions.data(), c.positions.size() * sizeof(float));
    if (!c.normals.empty()) memcpy(&out->normals[pn[i]], c.normals.data(), c.normals.size() * sizeof(float));
    if (!c.texcoords.empty()) memcpy(&out->texcoords[pt[i]], c.texcoords.data(), c.texcoords.size() * sizeof(float));
    if (!c.corners.empty()) memcpy(&out->corners[pc[i]], c.corners.data(), c.corners.size() * sizeof(int));
    std::vector<float>().swap(c.positions);
    std::vector<int>().swap(c.corners);
  };
  run(join);
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
