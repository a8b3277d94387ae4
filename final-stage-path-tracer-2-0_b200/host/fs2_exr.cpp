// fs2_exr.cpp — OpenEXR texture reader of the host side.
//
// The reference reads .exr textures through tinyexr's LoadEXR (call site: reference
// source/material.cpp:70-92: RGBA float out, rgb copied texel by texel in file row order).
// tinyexr is a third-party header that is NOT in /root/reference (un-vendored), so this is a
// restatement of the published OpenEXR 2 file layout, written for this project:
//   magic 20000630, version 2 (tiled / deep / multi-part files are rejected),
//   attribute list (channels, compression, dataWindow, lineOrder are read; the rest skipped),
//   chunk offset table, scan-line chunks { int32 y, int32 size, data }.
// Supported: pixel types HALF and FLOAT, compression NONE, RLE, ZIPS, ZIP (zlib + byte-delta
// predictor + even/odd de-interleave), either line order, any dataWindow origin.  Channel
// selection follows LoadEXR: R, G, B by name; a single-channel file becomes grey.
// Row 0 of the result is the top scan line (dataWindow.min.y), as tinyexr returns it.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <zlib.h>

#include <string>
#include <vector>

#include "fs2_assets.h"

namespace fs2h {

namespace {

struct Reader {
  const unsigned char* p;
  const unsigned char* end;
  bool ok = true;
  bool need(size_t n) {
    if ((size_t)(end - p) < n) ok = false;
    return ok;
  }
  uint32_t u32() {
    if (!need(4)) return 0;
    uint32_t v = p[0] | p[1] << 8 | p[2] << 16 | (uint32_t)p[3] << 24;
    p += 4;
    return v;
  }
  uint64_t u64() {
    uint64_t lo = u32(), hi = u32();
    return lo | hi << 32;
  }
  uint8_t u8() {
    if (!need(1)) return 0;
    return *p++;
  }
  std::string str() {  // zero-terminated
    std::string s;
    while (need(1) && *p) s.push_back((char)*p++);
    if (ok) p++;
    return s;
  }
};

float half_to_float(uint16_t h) {
  uint32_t sign = (uint32_t)(h >> 15) << 31, exp = (h >> 10) & 31u, man = h & 1023u, w;
  if (exp == 0) {
    if (man == 0) {
      w = sign;
    } else {  // subnormal half: normalise
      int e = -1;
      do {
        man <<= 1;
        e++;
      } while (!(man & 1024u));
      w = sign | (uint32_t)(127 - 15 - e) << 23 | (man & 1023u) << 13;
    }
  } else if (exp == 31) {
    w = sign | 0x7F800000u | man << 13;
  } else {
    w = sign | (exp + 127 - 15) << 23 | man << 13;
  }
  float f;
  memcpy(&f, &w, 4);
  return f;
}

struct Channel {
  std::string name;
  int type;       // 1 half, 2 float
  size_t offset;  // byte offset of the channel's run inside one scan line
};

// OpenEXR's post-processing of RLE / ZIP data: running byte sum, then even/odd interleave.
void unpredict(std::vector<unsigned char>& t, std::vector<unsigned char>& out) {
  for (size_t i = 1; i < t.size(); i++) t[i] = (unsigned char)(t[i - 1] + t[i] - 128);
  const size_t n = t.size(), half = (n + 1) / 2;
  out.resize(n);
  for (size_t i = 0; i < half; i++) {
    out[2 * i] = t[i];
    if (2 * i + 1 < n) out[2 * i + 1] = t[half + i];
  }
}

bool rle_decode(const unsigned char* s, size_t n, std::vector<unsigned char>& out, size_t want) {
  out.clear();
  const unsigned char* e = s + n;
  while (s < e) {
    int count = (signed char)*s++;
    if (count < 0) {
      size_t k = (size_t)(-count);
      if ((size_t)(e - s) < k) return false;
      out.insert(out.end(), s, s + k);
      s += k;
    } else {
      if (s >= e) return false;
      out.insert(out.end(), (size_t)count + 1, *s++);
    }
    if (out.size() > want) return false;
  }
  return out.size() == want;
}

}  // namespace

bool load_exr(const std::string& path, Image* out, std::string* error) {
  auto fail = [&](const char* why) {
    if (error) *error = why;
    return false;
  };
  std::vector<unsigned char> data;
  {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return fail("cannot open file");
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (n < 0 || n > (1L << 34)) {  // a directory opens and reports LONG_MAX
      fclose(f);
      return fail("not a regular file");
    }
    data.resize((size_t)n);
    size_t got = data.empty() ? 0 : fread(data.data(), 1, data.size(), f);
    fclose(f);
    data.resize(got);
  }
  Reader r{data.data(), data.data() + data.size()};
  if (r.u32() != 20000630u) return fail("not an OpenEXR file");
  uint32_t version = r.u32();
  if ((version & 0xFFu) != 2u) return fail("unsupported OpenEXR version");
  if (version & (0x200u | 0x800u | 0x1000u)) return fail("tiled, deep and multi-part files are not supported");
  std::vector<Channel> channels;
  int compression = -1, line_order = 0;
  int32_t win[4] = {0, 0, -1, -1};
  bool have_window = false;
  for (;;) {  // attribute list, closed by an empty name
    std::string name = r.str();
    if (!r.ok) return fail("truncated header");
    if (name.empty()) break;
    std::string type = r.str();
    uint32_t size = r.u32();
    if (!r.need(size)) return fail("truncated header");
    Reader a{r.p, r.p + size};
    r.p += size;
    if (name == "channels") {
      for (;;) {
        std::string cn = a.str();
        if (!a.ok) return fail("bad channel list");
        if (cn.empty()) break;
        Channel c;
        c.name = cn;
        c.type = (int)a.u32();
        a.u8();  // pLinear
        a.u8(), a.u8(), a.u8();
        uint32_t xs = a.u32(), ys = a.u32();
        if (!a.ok) return fail("bad channel list");
        if (xs != 1 || ys != 1) return fail("sub-sampled channels are not supported");
        if (c.type != 1 && c.type != 2) return fail("only HALF and FLOAT channels are supported");
        c.offset = 0;
        channels.push_back(c);
      }
    } else if (name == "compression") {
      compression = a.u8();
    } else if (name == "dataWindow") {
      for (int i = 0; i < 4; i++) win[i] = (int32_t)a.u32();
      have_window = a.ok;
    } else if (name == "lineOrder") {
      line_order = a.u8();
    }
  }
  if (channels.empty() || compression < 0 || !have_window) return fail("missing required attribute");
  if (compression > 3) return fail("only NONE, RLE, ZIPS and ZIP compression are supported");
  if (line_order > 1) return fail("unsupported line order");
  const long w = (long)win[2] - win[0] + 1, h = (long)win[3] - win[1] + 1;
  if (w <= 0 || h <= 0 || w > 65536 || h > 65536 || w * h > (1L << 28)) return fail("bad data window");
  size_t line_bytes = 0;
  for (Channel& c : channels) {
    c.offset = line_bytes;
    line_bytes += (size_t)w * (c.type == 1 ? 2 : 4);
  }
  // LoadEXR's channel choice
  int pick[3] = {-1, -1, -1};
  for (size_t i = 0; i < channels.size(); i++) {
    if (channels[i].name == "R") pick[0] = (int)i;
    if (channels[i].name == "G") pick[1] = (int)i;
    if (channels[i].name == "B") pick[2] = (int)i;
  }
  if (pick[0] < 0 || pick[1] < 0 || pick[2] < 0) {
    if (channels.size() != 1) return fail("R, G or B channel not found");
    pick[0] = pick[1] = pick[2] = 0;
  }
  const int lines_per_block = compression == 3 ? 16 : 1;
  const size_t blocks = ((size_t)h + lines_per_block - 1) / lines_per_block;
  std::vector<uint64_t> offsets(blocks);
  for (uint64_t& o : offsets) o = r.u64();
  if (!r.ok) return fail("truncated offset table");
  // every scan line needs at least its chunk header in the file: refuse windows the file cannot hold
  if (blocks * 8 > data.size()) return fail("data window larger than the file");
  out->width = (uint32_t)w;
  out->height = (uint32_t)h;
  out->rgb.assign((size_t)w * h * 3, 0.f);
  std::vector<unsigned char> tmp, raw;
  for (size_t b = 0; b < blocks; b++) {
    if (offsets[b] > data.size()) return fail("bad chunk offset");
    Reader c{data.data() + offsets[b], data.data() + data.size()};
    int32_t y0 = (int32_t)c.u32();
    uint32_t size = c.u32();
    if (!c.ok || !c.need(size)) return fail("truncated chunk");
    long row0 = (long)y0 - win[1];
    if (row0 < 0 || row0 >= h) return fail("chunk outside the data window");
    long rows = h - row0 < lines_per_block ? h - row0 : lines_per_block;
    const size_t want = line_bytes * (size_t)rows;
    const unsigned char* px = c.p;
    if (compression != 0 && size < want) {  // a chunk that did not shrink is stored raw
      if (compression == 1) {
        if (!rle_decode(c.p, size, tmp, want)) return fail("bad RLE data");
      } else {
        tmp.resize(want);
        uLongf got = (uLongf)want;
        if (uncompress(tmp.data(), &got, c.p, size) != Z_OK || got != want) return fail("bad ZIP data");
      }
      unpredict(tmp, raw);
      px = raw.data();
    } else if (size != want) {
      return fail("bad chunk size");
    }
    for (long k = 0; k < rows; k++) {
      const unsigned char* line = px + (size_t)k * line_bytes;
      float* dst = &out->rgb[(size_t)(row0 + k) * w * 3];
      for (int ch = 0; ch < 3; ch++) {
        const Channel& cc = channels[pick[ch]];
        const unsigned char* s = line + cc.offset;
        for (long x = 0; x < w; x++) {
          float v;
          if (cc.type == 1) {
            v = half_to_float((uint16_t)(s[2 * x] | s[2 * x + 1] << 8));
          } else {
            memcpy(&v, s + 4 * x, 4);
          }
          dst[3 * x + ch] = v;
        }
      }
    }
  }
  return true;
}

// Writer for the headless front end: FLOAT B, G, R channels, ZIP
// compression (16-line blocks; a block that does not shrink is stored raw), increasing line
// order.  rgb: 3 floats per pixel; `bottom_up` says row 0 of rgb is the bottom scan line (the
// film's order) and is flipped so that the file's first line is the top of the picture.
bool save_exr(const std::string& path, uint32_t w, uint32_t h, const float* rgb, bool bottom_up) {
  if (!w || !h || !rgb) return false;
  std::vector<unsigned char> head;
  auto put32 = [&](std::vector<unsigned char>& v, uint32_t x) {
    for (int i = 0; i < 4; i++) v.push_back((unsigned char)(x >> (8 * i)));
  };
  auto puts = [&](const char* t) {
    while (*t) head.push_back((unsigned char)*t++);
    head.push_back(0);
  };
  auto attr = [&](const char* name, const char* type, const std::vector<unsigned char>& payload) {
    puts(name);
    puts(type);
    put32(head, (uint32_t)payload.size());
    head.insert(head.end(), payload.begin(), payload.end());
  };
  put32(head, 20000630u);
  put32(head, 2u);
  {
    std::vector<unsigned char> ch;
    for (const char* n : {"B", "G", "R"}) {
      ch.push_back((unsigned char)n[0]);
      ch.push_back(0);
      put32(ch, 2u);  // FLOAT
      put32(ch, 0u);  // pLinear + reserved
      put32(ch, 1u);
      put32(ch, 1u);
    }
    ch.push_back(0);
    attr("channels", "chlist", ch);
  }
  attr("compression", "compression", {3});
  std::vector<unsigned char> box;
  put32(box, 0u), put32(box, 0u), put32(box, w - 1), put32(box, h - 1);
  attr("dataWindow", "box2i", box);
  attr("displayWindow", "box2i", box);
  attr("lineOrder", "lineOrder", {0});
  const float one = 1.0f, zero = 0.0f;
  std::vector<unsigned char> f1((const unsigned char*)&one, (const unsigned char*)&one + 4);
  std::vector<unsigned char> f0((const unsigned char*)&zero, (const unsigned char*)&zero + 4);
  attr("pixelAspectRatio", "float", f1);
  std::vector<unsigned char> v2(f0);
  v2.insert(v2.end(), f0.begin(), f0.end());
  attr("screenWindowCenter", "v2f", v2);
  attr("screenWindowWidth", "float", f1);
  head.push_back(0);
  const uint32_t blocks = (h + 15) / 16;
  const size_t line_bytes = (size_t)w * 12;
  std::vector<std::vector<unsigned char>> chunks(blocks);
  std::vector<unsigned char> raw, split, packed;
  for (uint32_t b = 0; b < blocks; b++) {
    const uint32_t r0 = b * 16, rows = h - r0 < 16 ? h - r0 : 16;
    raw.resize(line_bytes * rows);
    for (uint32_t k = 0; k < rows; k++) {
      const uint32_t src_row = bottom_up ? h - 1 - (r0 + k) : r0 + k;
      const float* src = rgb + (size_t)src_row * w * 3;
      float* line = (float*)(raw.data() + (size_t)k * line_bytes);
      for (uint32_t x = 0; x < w; x++) {  // channel runs in alphabetical order: B, G, R
        line[x] = src[3 * x + 2];
        line[w + x] = src[3 * x + 1];
        line[2 * (size_t)w + x] = src[3 * x];
      }
    }
    // even/odd split, byte delta, deflate
    const size_t n = raw.size(), half = (n + 1) / 2;
    split.resize(n);
    for (size_t i = 0; i < n; i++) split[(i & 1) ? half + i / 2 : i / 2] = raw[i];
    for (size_t i = n - 1; i > 0; i--) split[i] = (unsigned char)(split[i] - split[i - 1] + 128);
    uLongf cap = compressBound((uLong)n);
    packed.resize(cap);
    const bool shrunk = compress2(packed.data(), &cap, split.data(), (uLong)n, 6) == Z_OK && cap < n;
    const std::vector<unsigned char>& body = shrunk ? packed : raw;
    const size_t body_n = shrunk ? (size_t)cap : n;
    put32(chunks[b], r0);
    put32(chunks[b], (uint32_t)body_n);
    chunks[b].insert(chunks[b].end(), body.begin(), body.begin() + body_n);
  }
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) return false;
  bool ok = fwrite(head.data(), 1, head.size(), f) == head.size();
  uint64_t pos = head.size() + 8ull * blocks;
  for (uint32_t b = 0; b < blocks && ok; b++) {
    ok = fwrite(&pos, 8, 1, f) == 1;  // little-endian host
    pos += chunks[b].size();
  }
  for (uint32_t b = 0; b < blocks && ok; b++) ok = fwrite(chunks[b].data(), 1, chunks[b].size(), f) == chunks[b].size();
  return fclose(f) == 0 && ok;
}

}  // namespace fs2h
