// ctr_rng.h — shared state between oracle/ref/harness.cpp (--paths) and ctr_rng.cpp.
// TEST INFRASTRUCTURE (oracle side), see ctr_rng.cpp.
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace fs2ctr {

struct PathRecord {  // 48 bytes, one per (sample, pixel)
  float rgb[3];
  float normal[3];
  float depth;
  uint32_t ray_count;
  uint64_t mat_id;
  uint64_t pad;
};

struct State {
  bool active = false;
  uint64_t seed = 0;
  uint32_t sample = 0;       // sample index of the frame being traced
  uint32_t sample_slot = 0;  // which [w*h] slab of `out` that frame fills
  uint32_t pixel = 0;        // pixel whose stream is current
  uint64_t stream = 0;
  PathRecord* out = nullptr;
  uint64_t order_errors = 0;  // WritePixel calls that did not follow the scan order
};

extern State g;
void begin_pixel(uint32_t pixel);

}  // namespace fs2ctr
