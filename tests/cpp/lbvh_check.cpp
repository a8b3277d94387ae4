// lbvh_check.cpp — CPU check of the leaf-BVH builder and of the walk's exactness argument:
// for random segments and every mesh-tree leaf they can touch, the 4-wide BVH walk must
// return the same (parameter bits, original record index, u bits, v bits) as the linear leaf
// loop of reference source/mesh.cpp:89-110.  The walk below restates csrc/fs2_lbvh.cuh for
// the host; the triangle test restates mesh_leaf() of csrc/fs2_traverse.cuh.
//
// usage: lbvh_check scene.fs2 n_rays seed   (prints "ok <leaf queries> <mismatches>")
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "fs2_blob.h"
#include "fs2_lbvh_build.h"

using namespace fs2;

static bool near_zero(float a) { return a > -1.0e-5f && a < 1.0e-5f; }

struct Best {
  float t, u, v;
  uint32_t idx;  // original packed index
  bool found;
};

static bool tri_inside(const fs_tri_aux& a, const float p[3], float& u, float& v) {
  float vp[3] = {p[0] - a.p1[0], p[1] - a.p1[1], p[2] - a.p1[2]};
  uint32_t drop = a.face >> 30;
  float vpa = (drop == 0u) ? vp[1] : vp[0];
  float vpb = (drop == 2u) ? vp[1] : vp[2];
  float dot02 = a.v0a * vpa + a.v0b * vpb;
  float dot12 = a.v1a * vpa + a.v1b * vpb;
  float uu = (a.dot11 * dot02 - a.dot01 * dot12) * a.inv_denom;
  float vv = (a.dot00 * dot12 - a.dot01 * dot02) * a.inv_denom;
  if (uu >= 0.0f && vv >= 0.0f && uu + vv <= 1.0f) {
    u = uu;
    v = vv;
    return true;
  }
  return false;
}

static bool plane_t(const float* pl, const float s[3], const float d[3], float& t) {
  float ts = pl[0] * d[0] + pl[1] * d[1] + pl[2] * d[2];
  if (near_zero(ts)) return false;
  float ns = -(pl[0] * s[0] + pl[1] * s[1] + pl[2] * s[2] + pl[3]);
  t = ns / ts;
  return t >= 0.0f && t <= 1.0f;
}

// the reference's sequential loop on the original arrays
static Best linear(const fs_scene_desc& sc, uint32_t first, uint32_t count, const float s[3],
                   const float d[3], float limit) {
  Best b{limit, 0, 0, 0, false};
  for (uint32_t i = 0; i < count; i++) {
    uint32_t k = first + i;
    float t;
    if (!plane_t(sc.tri_plane + 4 * (size_t)k, s, d, t)) continue;
    if (!(t < b.t)) continue;
    float p[3] = {s[0] + d[0] * t, s[1] + d[1] * t, s[2] + d[2] * t}, u, v;
    if (tri_inside(sc.tri_aux[k], p, u, v)) b = Best{t, u, v, k, true};
  }
  return b;
}

static void test_tris(const LbvhHost& lb, uint32_t k0, uint32_t n, const float s[3],
                      const float d[3], Best& b, uint64_t& n_tests) {
  n_tests += n;
  for (uint32_t j = 0; j < n; j++) {
    uint32_t k = k0 + j;
    float t;
    if (!plane_t(&lb.plane[4 * (size_t)k], s, d, t)) continue;
    uint32_t ord = lb.order[k];
    if (!(t < b.t)) {
      if (!(t == b.t) || !b.found) continue;
      if (!(ord < b.idx)) continue;
    }
    float p[3] = {s[0] + d[0] * t, s[1] + d[1] * t, s[2] + d[2] * t}, u, v;
    if (tri_inside(lb.aux[k], p, u, v)) b = Best{t, u, v, ord, true};
  }
}

static float inv(float d) {
  const float tiny = 1.0e-30f;
  if (fabsf(d) < tiny) d = (d < 0.0f) ? -tiny : tiny;
  return 1.0f / d;
}

static float child(const LbNode& n, int c, const float s[3], const float id[3], float lim) {
  float ax = (n.lox[c] - s[0]) * id[0], bx = (n.hix[c] - s[0]) * id[0];
  float ay = (n.loy[c] - s[1]) * id[1], by = (n.hiy[c] - s[1]) * id[1];
  float az = (n.loz[c] - s[2]) * id[2], bz = (n.hiz[c] - s[2]) * id[2];
  float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
  float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
  float lo = tn * 0.99998f;
  return (lo <= tf * 1.00002f && lo <= lim) ? lo : 3.0f;
}

static Best walk(const LbvhHost& lb, uint32_t node, uint32_t first, uint32_t count,
                 const float s[3], const float d[3], float limit, uint64_t& n_nodes,
                 uint64_t& n_tests, int& max_sp) {
  Best b{limit, 0, 0, 0, false};
  uint32_t root = lb.leaf[node].root;
  if (root == 0xFFFFFFFFu) {
    test_tris(lb, first, count, s, d, b, n_tests);
    return b;
  }
  float id[3] = {inv(d[0]), inv(d[1]), inv(d[2])};
  struct E {
    uint32_t n;
    float t;
  } stack[64];
  int sp = 0;
  uint32_t cur = root;
  for (;;) {
    const LbNode& n = lb.nodes[cur];
    n_nodes++;
    float lim = fminf(b.t, 1.0f);
    float t[4];
    uint32_t c[4];
    for (int k = 0; k < 4; k++) {
      t[k] = child(n, k, s, id, lim);
      c[k] = n.child + k;
    }
    bool descended = false;
    if (n.kind) {
      for (uint32_t k = 0; k < 4; k++)
        if (t[k] < 3.0f && t[k] <= fminf(b.t, 1.0f)) {
          uint32_t rem = n.count - 4 * k;
          test_tris(lb, n.child + 4 * k, rem < 4 ? rem : 4, s, d, b, n_tests);
        }
    } else {
      auto cswap = [&](int a, int bb) {
        if (t[bb] < t[a]) {
          std::swap(t[a], t[bb]);
          std::swap(c[a], c[bb]);
        }
      };
      cswap(0, 1);
      cswap(2, 3);
      cswap(0, 2);
      cswap(1, 3);
      cswap(1, 2);
      for (int k = 3; k >= 1; k--)
        if (t[k] < 3.0f) stack[sp++] = E{c[k], t[k]};
      if (sp > max_sp) max_sp = sp;
      if (t[0] < 3.0f) {
        cur = c[0];
        descended = true;
      }
    }
    if (descended) continue;
    bool more = false;
    while (sp > 0) {
      E e = stack[--sp];
      if (e.t <= fminf(b.t, 1.0f)) {
        cur = e.n;
        more = true;
        break;
      }
    }
    if (!more) break;
  }
  return b;
}

static uint64_t rng_state = 88172645463325252ull;
static double rnd() {
  rng_state ^= rng_state << 13;
  rng_state ^= rng_state >> 7;
  rng_state ^= rng_state << 17;
  return (double)(rng_state >> 11) / 9007199254740992.0;
}

static bool seg_hits_box(const fs_node& n, const float s[3], const float d[3]) {
  float tn = 0.f, tf = 1.f;
  for (int k = 0; k < 3; k++) {
    float id = inv(d[k]);
    float a = (n.bmin[k] - s[k]) * id, b = (n.bmax[k] - s[k]) * id;
    tn = fmaxf(tn, fminf(a, b));
    tf = fminf(tf, fmaxf(a, b));
  }
  return tn <= tf * 1.001f + 1e-6f;
}

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  fs_scene_desc sc;
  void* storage = nullptr;
  if (fs_blob_load(argv[1], &sc, &storage) != FS_OK) {
    fprintf(stderr, "cannot load %s\n", argv[1]);
    return 2;
  }
  int n_rays = atoi(argv[2]);
  rng_state ^= (uint64_t)atoll(argv[3]) * 0x9E3779B97F4A7C15ull;
  LbvhHost lb;
  build_lbvh(&sc, &lb, 9, 4);
  // permutation sanity: order[] is a permutation within each leaf and records moved with it
  std::vector<uint32_t> leaves;
  float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
  for (uint32_t i = 0; i < sc.n_nodes; i++) {
    const fs_node& n = sc.nodes[i];
    if (!lb.mesh_leaf[i] || n.count == FS_NODE_INTERNAL || n.count == 0) continue;
    if ((uint64_t)n.first + n.count > sc.n_tris) continue;
    leaves.push_back(i);
    std::vector<uint8_t> seen(n.count, 0);
    for (uint32_t j = 0; j < n.count; j++) {
      uint32_t o = lb.order[n.first + j];
      if (o < n.first || o >= n.first + n.count || seen[o - n.first]) {
        printf("FAIL permutation leaf %u\n", i);
        return 1;
      }
      seen[o - n.first] = 1;
      if (memcmp(&lb.aux[n.first + j], &sc.tri_aux[o], sizeof(fs_tri_aux)) ||
          memcmp(&lb.plane[4 * (size_t)(n.first + j)], sc.tri_plane + 4 * (size_t)o, 16)) {
        printf("FAIL record move leaf %u\n", i);
        return 1;
      }
    }
    for (int k = 0; k < 3; k++) {
      lo[k] = fminf(lo[k], n.bmin[k]);
      hi[k] = fmaxf(hi[k], n.bmax[k]);
    }
  }
  if (leaves.empty()) {
    printf("ok 0 0 (no mesh leaves)\n");
    return 0;
  }
  // segments that start beyond origin_limit never reach the walk (the kernels scan such leaves
  // linearly, fs2_phase.cuh PH_F_FAR); origins up to the limit must be exact
  const float origin_limit = lb.origin_limit();
  uint64_t queries = 0, mismatches = 0, hits = 0, n_nodes = 0, n_tests = 0, n_linear = 0, n_far = 0;
  int max_sp = 0;
  for (int r = 0; r < n_rays; r++) {
    float s[3], e[3], d[3];
    int mode = r % 4;
    for (int k = 0; k < 3; k++) {
      float ext = hi[k] - lo[k];
      // origins inside and well outside the meshes; targets inside the mesh bounds
      s[k] = lo[k] + ext * (float)(mode == 0 ? rnd() : (rnd() * 5.0 - 2.0));
      e[k] = lo[k] + ext * (float)rnd();
    }
    if (r % 7 == 6) {
      // origin on the surface of the cube |s|_inf = origin_limit: the worst rounding of the hit
      // point the box pads have to cover
      int ax = (int)(rnd() * 3) % 3;
      for (int k = 0; k < 3; k++) s[k] = origin_limit * (float)(rnd() * 2.0 - 1.0);
      s[ax] = (rnd() < 0.5 ? -origin_limit : origin_limit);
    }
    if (fmaxf(fmaxf(fabsf(s[0]), fabsf(s[1])), fabsf(s[2])) > origin_limit) {
      n_far++;
      continue;
    }
    if (mode == 3) e[r % 3] = s[r % 3];  // axis-parallel component
    if (r % 5 == 4) {
      // aim exactly at a vertex or an edge midpoint of a random record: grazing hits, ties
      // between the triangles that share it
      uint32_t k = (uint32_t)(rnd() * sc.n_tris) % sc.n_tris;
      float v[4][3];
      lbvh_detail::record_vertices(sc.tri_plane + 4 * (size_t)k, sc.tri_aux[k], v);
      int a = (int)(rnd() * 3) % 3, b = (int)(rnd() * 3) % 3;
      for (int q = 0; q < 3; q++) e[q] = (a == b) ? v[a][q] : 0.5f * (v[a][q] + v[b][q]);
    }
    float len = 0.f;
    for (int k = 0; k < 3; k++) len += (e[k] - s[k]) * (e[k] - s[k]);
    len = sqrtf(len);
    if (len == 0.f) continue;
    float scale = (mode == 1 ? 1.0f : 10000.0f / len);  // engine rays are ~z_far long
    for (int k = 0; k < 3; k++) d[k] = (e[k] - s[k]) * scale;
    for (uint32_t li : leaves) {
      const fs_node& n = sc.nodes[li];
      if (!seg_hits_box(n, s, d)) continue;
      float limit = (rnd() < 0.3) ? (float)(rnd() * 1.2 * len / 10000.0f * 2.0) : 2.0f;
      Best a = linear(sc, n.first, n.count, s, d, limit);
      Best b = walk(lb, li, n.first, n.count, s, d, limit, n_nodes, n_tests, max_sp);
      n_linear += n.count;
      queries++;
      hits += a.found;
      bool same = a.found == b.found;
      if (same && a.found)
        same = memcmp(&a.t, &b.t, 4) == 0 && a.idx == b.idx && memcmp(&a.u, &b.u, 4) == 0 &&
               memcmp(&a.v, &b.v, 4) == 0;
      if (!same) {
        if (mismatches < 10)
          printf("MISMATCH leaf %u: linear found=%d t=%.9g idx=%u | walk found=%d t=%.9g idx=%u\n", li,
                 a.found, a.t, a.idx, b.found, b.t, b.idx);
        mismatches++;
      }
    }
  }
  printf("%s %llu queries, %llu hits, %llu mismatches; per query: %.1f linear tests vs %.1f nodes + %.1f tests; "
         "max stack %d; origin limit %g (%llu segments beyond it skipped)\n",
         mismatches ? "FAIL" : "ok", (unsigned long long)queries, (unsigned long long)hits,
         (unsigned long long)mismatches, (double)n_linear / queries, (double)n_nodes / queries,
         (double)n_tests / queries, max_sp, origin_limit, (unsigned long long)n_far);
  free(storage);
  return mismatches ? 1 : 0;
}
