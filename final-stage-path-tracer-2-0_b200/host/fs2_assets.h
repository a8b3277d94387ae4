// fs2_assets.h — .obj / .bmp / .hdr readers (fs2_assets.cpp) and the .exr reader (fs2_exr.cpp).
#ifndef FS2_ASSETS_H_
#define FS2_ASSETS_H_

#include <stdint.h>

#include <string>
#include <vector>

namespace fs2h {

struct ObjData {
  std::vector<float> positions;  // 3 per vertex
  std::vector<float> normals;    // 3 per normal
  std::vector<float> texcoords;  // 2 per texcoord
  std::vector<int> corners;      // (vertex, normal, texcoord) per triangle corner; -1 = absent
};

struct Image {
  uint32_t width = 0, height = 0;
  std::vector<float> rgb;  // 3 floats per texel, rows in file order
};

bool load_obj(const std::string& path, ObjData* out, std::string* error);
bool load_bmp(const std::string& path, Image* out);
bool load_hdr(const std::string& path, Image* out);
// OpenEXR scan-line files (fs2_exr.cpp): HALF / FLOAT, NONE / RLE / ZIPS / ZIP; rows top to bottom.
bool load_exr(const std::string& path, Image* out, std::string* error);
// FLOAT B/G/R, ZIP-compressed scan-line file; bottom_up: row 0 of rgb is the bottom scan line.
bool save_exr(const std::string& path, uint32_t w, uint32_t h, const float* rgb, bool bottom_up);

}  // namespace fs2h
#endif
