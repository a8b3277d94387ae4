// TEST INFRASTRUCTURE — known-answer probe of the UNMODIFIED reference's math (linked from the
// objects oracle/ref/Makefile compiles where the sources lie).  Prints the 32-bit words of
// SURVEY.md Appendix C's probes in the layout of csrc/fs2_selftest.cuh (FS2_SELFTEST_WORDS),
// one hex word per line; tests/golden/make_kat.py stores them as tests/golden/kat_words.txt.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "math/intersect.h"
#include "math/random.h"
#include "frame.h"

using namespace base;

static uint32_t W[48];
static void put(int i, float f) { memcpy(&W[i], &f, 4); }
static void put3(int i, const vector3& v) {
  put(i, v[0]);
  put(i + 1, v[1]);
  put(i + 2, v[2]);
}

int main() {
  put(0, inv_sqrt(2.0f));
  put(1, base::sqrtf(2.0f));
  put(2, vector3(3, 4, 0).length());
  const vector3 eye(-5.8f, 7.05f, -47.06f);
  {
    ray r(eye, vector3(120.5f, -33.25f, 9000.f));
    collision c;
    plane pl = calculate_plane(vector3(0, 1, 0), vector3(0, 0, 0));
    bool hit = ray_intersect_plane(pl, r, &c);
    put(3, hit ? c.param : -1.f);
    put3(4, c.point);
    vector3 p0(-10, 0, 0), p1(30, -2, 10), p2(5, 40, 20);
    collision t;
    vector2 bary;
    plane tp = calculate_plane(p0, p1, p2);
    hit = ray_intersect_triangle(p0, p1, p2, tp, r, &t, &bary);
    put(7, hit ? t.param : -1.f);
    put(8, bary[0]);
    put(9, bary[1]);
    put3(10, vector3(tp[0], tp[1], tp[2]));
  }
  {
    vector3 stop = eye + (vector3(4.5f, 6.f, 1.f) - eye) * 200.f;
    ray r(eye, stop);
    collision c;
    bool hit = ray_intersect_sphere(vector3(4, 5, 0), 5.f, r, &c);
    put(13, hit ? c.param : -1.f);
    put3(14, c.point);
    put3(17, stop);
    bounds b;
    b.set_min(vector3(-1, 0, -1));
    b.set_max(vector3(9, 10, 9));
    collision bc;
    hit = ray_intersect_bounds(b, r, &bc);
    put(20, hit ? bc.param : -1.f);
    put3(21, bc.point);
  }
  {
    vector3 up(0, 1, 0);
    vector3 v = vector3(0.3f, -0.8f, 0.52f).normalize();
    put3(24, v);
    put3(27, v.reflect(up));
    put3(30, v.refract(up, 0.75f));
    put3(33, v.rotate(0.7f, up));
    vector2 uv = sphere_map_texcoords(v);
    put(36, uv[0]);
    put(37, uv[1]);
    put3(38, vector3(0.9f, -0.1f, 0.42f).normalize().refract(up, 1.5f));
  }
  set_seed(521288629ull);
  for (int i = 0; i < 3; i++) {
    uint64_t x = random_integer();
    W[41 + 2 * i] = (uint32_t)x;
    W[42 + 2 * i] = (uint32_t)(x >> 32);
  }
  {
    DisplayFrame f(1, 1);
    f.WritePixel(vector3(0.2f, 0.5f, 1.7f), 0, 0);
    f.WritePixel(vector3(0.4f, 0.1f, 0.0f), 0, 0);
    const uint8* d = f.GetDisplayBuffer();
    W[47] = d[0] | d[1] << 8 | d[2] << 16;
  }
  for (int i = 0; i < 48; i++) printf("%08x\n", W[i]);
  return 0;
}
