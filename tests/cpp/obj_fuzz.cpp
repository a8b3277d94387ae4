// obj_fuzz.cpp — loads every file named on the command line with the .obj reader (host/fs2_assets.cpp)
// and prints a one-line summary; built with ASan + UBSan by tests/test_obj_reader.py.
#include <stdio.h>
#include <string>
#include "fs2_assets.h"
int main(int argc, char** argv) {
  size_t ok = 0, bad = 0, tris = 0;
  for (int i = 1; i < argc; i++) {
    fs2h::ObjData d;
    std::string e;
    if (fs2h::load_obj(argv[i], &d, &e)) {
      ok++;
      tris += d.corners.size() / 9;
      // touch everything the mesh constructor would index
      for (size_t k = 0; k + 2 < d.corners.size(); k += 3) {
        int v = d.corners[k];
        if (v < 0 || (size_t)v >= d.positions.size() / 3) return 3;  // load_obj promised this
      }
    } else {
      bad++;
    }
  }
  printf("loaded %zu, refused %zu, %zu triangles\n", ok, bad, tris);
  return 0;
}
