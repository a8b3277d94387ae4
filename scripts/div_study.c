/* Study for DESIGN.md §9 item 2: can the octree slab / split-plane divisions a / b (IEEE div.rn,
 * ~10 SASS instructions each) be replaced by the 3-instruction sequence
 *     q0 = a * r;  e = fma(-b, q0, a);  q = fma(e, r, q0)        with r = RN(1 / b) per ray
 * without changing a single result bit?  Counts mismatches against a / b over random operands in
 * the magnitudes the traversal sees (|b| in 2^-20 .. 2^14, |a| in 2^-24 .. 2^14) and over
 * divisors with special significands.   gcc -O2 -ffp-contract=off -fopenmp div_study.c -lm */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

static inline uint64_t next(uint64_t* s) {
  uint64_t x = *s;
  x ^= x << 12;
  x ^= x >> 25;
  x ^= x << 27;
  *s = x;
  return x * 0x2545F4914F6CDD1Dull;
}
static inline float mk(uint32_t sign, int exp, uint32_t mant) {
  uint32_t w = sign << 31 | (uint32_t)(exp + 127) << 23 | (mant & 0x7FFFFF);
  float f;
  memcpy(&f, &w, 4);
  return f;
}

int main(void) {
  const long per_thread = 400000000L;
  long bad = 0, total = 0, bad_ones = 0;
#pragma omp parallel for reduction(+ : bad, total, bad_ones)
  for (int t = 0; t < 16; t++) {
    uint64_t s = 0x9E3779B97F4A7C15ull * (uint64_t)(t + 1);
    for (long i = 0; i < per_thread; i++) {
      uint64_t x = next(&s), y = next(&s);
      uint32_t mb = (uint32_t)x & 0x7FFFFF;
      if ((i & 15) == 0) mb = 0x7FFFFF - ((uint32_t)(x >> 40) & 7);  /* significands near all-ones */
      float b = mk((uint32_t)(x >> 23) & 1, (int)((x >> 24) % 35) - 20, mb);
      float a = mk((uint32_t)(y >> 23) & 1, (int)((y >> 24) % 39) - 24, (uint32_t)y);
      float r = 1.0f / b;
      float q0 = a * r;
      float e = fmaf(-b, q0, a);
      float q = fmaf(e, r, q0);
      float ref = a / b;
      total++;
      if (memcmp(&q, &ref, 4)) {
        bad++;
        if (mb >= 0x7FFFF8) bad_ones++;
      }
    }
  }
  printf("pairs %ld, mismatches %ld (%.3g), of which divisor significand near all-ones: %ld\n", total, bad,
         (double)bad / (double)total, bad_ones);
  return 0;
}
