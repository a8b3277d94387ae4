// ctr_rng.cpp — feeds the CUDA path's per-path random streams to the UNMODIFIED reference.
//
// TEST INFRASTRUCTURE (oracle side).  Linked, ahead of the reference's own objects, into
// oracle/_ref/fs2_ref_ctr only (never into fs2_ref, the timing / golden-film binary).  It
// replaces five symbols of the reference at link time (-Wl,--allow-multiple-definition, first
// definition wins); no reference source is edited or copied:
//
//   base::set_seed, random_float, random_integer_range, random_float_range
//       (reference source/math/random.cpp:8-47).  While a path dump is active they draw from
//       the device generator (csrc/fs2_shade.cuh: xorshift64* started from a splitmix64 hash of
//       (seed, pixel, sample); float = top 24 bits / 2^24; range = (top 32 bits * span) >> 32).
//       Outside a dump they forward to the reference's own base::random_integer() with the
//       reference's expressions, so scene loading, material ids and the normal LUT are what
//       fs2_ref produces (tests/golden/make_golden.py checks the flattened blobs are identical).
//   base::DisplayFrame::WritePixel(const TraceResult&, x, y)   (source/frame.cpp:78-85)
//       records the path's TraceResult instead of averaging it, then starts the stream of the
//       next pixel in the reference's scan order.
//
// Everything between those hooks — ray generation with jitter and thin lens, TraceStep,
// Scene::Trace, the octrees, every intersector and material — is the reference's own code,
// driven through its own TraceScene (engine.cpp:166-296) with one worker thread.
//
// One draw cannot be reproduced by a forward-evaluating path tracer in the reference's order:
// FogMaterial::Sample at depth 0 draws *after* the whole sub-path returned (material.cpp:334).
// Here it comes, like every other draw, from the path's stream at that point; the CPU
// restatement reproduces that order with FS2O_FOG_SHARED_STREAM (pinned against this binary),
// and the device's dedicated fog stream is compared against the restatement's other mode.
#include "ctr_rng.h"

#include "frame.h"
#include "math/random.h"

namespace fs2ctr {
State g;

static inline uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

void begin_pixel(uint32_t pixel) {
  g.pixel = pixel;
  uint64_t s = splitmix64(splitmix64(g.seed ^ ((uint64_t)g.sample << 32 | pixel)) + 0x632BE59BD9B4E019ull);
  g.stream = s ? s : 521288629ull;
}

static inline uint64_t next() {
  uint64_t x = g.stream;
  x ^= x << 12;
  x ^= x >> 25;
  x ^= x << 27;
  g.stream = x;
  return x * 0x2545F4914F6CDD1Dull;
}
}  // namespace fs2ctr

namespace base {

extern thread_local uint64 g_random_seed;  // reference source/math/random.cpp:5

void set_seed(uint64 seed) {
  if (fs2ctr::g.active) {
    fs2ctr::begin_pixel(0);  // TraceThreadFunction, once per frame (engine.cpp:179); one band
    return;
  }
  g_random_seed = seed;
}

float32 random_float() {
  if (fs2ctr::g.active) return (float32)(fs2ctr::next() >> 40) * (1.0f / 16777216.0f);
  return ((float32)(random_integer() % BASE_MAX_INT32)) / BASE_MAX_INT32;
}

int64 random_integer_range(int32 imin, int32 imax) {
  if (imax < imin) return 0;
  if (imin == imax) return imin;
  uint32 span = (imax - imin) + 1;
  if (fs2ctr::g.active) return (int32)imin + (int64)(((fs2ctr::next() >> 32) * (uint64)span) >> 32);
  return ((int32)imin + (random_integer() % span));
}

float32 random_float_range(float32 imin, float32 imax) {
  if (imax < imin) return 0.0f;
  return imin + random_float() * (imax - imin);
}

void DisplayFrame::WritePixel(const TraceResult& result, uint32 x, uint32 y) {
  fs2ctr::State& g = fs2ctr::g;
  if (g.active && g.out) {
    fs2ctr::PathRecord& r = g.out[(size_t)g.sample_slot * width_ * height_ + (size_t)y * width_ + x];
    r.rgb[0] = result.color.x;
    r.rgb[1] = result.color.y;
    r.rgb[2] = result.color.z;
    r.normal[0] = result.normal.x;
    r.normal[1] = result.normal.y;
    r.normal[2] = result.normal.z;
    r.depth = result.depth;
    r.ray_count = (uint32_t)result.ray_count;
    r.mat_id = result.material_id;
    if (y * width_ + x != g.pixel) g.order_errors++;
  }
  fs2ctr::begin_pixel(y * width_ + x + 1);
}

}  // namespace base
