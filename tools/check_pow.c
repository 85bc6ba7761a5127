// check_pow.c — pow_core.h (the device's table-driven powf) compiled for the host and compared with libquadmath's powq.
//   gcc -O2 -ffp-contract=off -I master_b200/csrc tools/check_pow.c -o /tmp/check_pow -lquadmath -lm && /tmp/check_pow [n]
// Prints how many of n random (x, y) — x in (0, 1] weighted towards 1 like a Phong cosine, y a Phong exponent or 1 / (exponent + 1) —
// round to a float other than the correctly rounded one, for the tables and for float(exp2(y * log2(x))) in host double arithmetic.
#include <math.h>
#include <quadmath.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pow_tables.inc"
static const uint64_t kLog[128][2] = {B2PT_POW_LOG_TABLE};
static const uint64_t kExp[64] = {B2PT_POW_EXP_TABLE};
static const uint64_t kLc[6] = {B2PT_POW_LOG_COEFFS};
static const uint64_t kEc[5] = {B2PT_POW_EXP_COEFFS};
#define B2PT_POW_QUAL static inline
#define B2PT_POW_LOG(i, w) kLog[i][w]
#define B2PT_POW_EXP(j) kExp[j]
#define B2PT_POW_LC(k) kLc[k]
#define B2PT_POW_EC(k) kEc[k]
static inline double b2pt_pow_u2d(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }
static inline uint64_t b2pt_pow_d2u(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
static inline uint32_t b2pt_pow_f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float b2pt_pow_u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline double b2pt_pow_fma(double a, double b, double c) { return fma(a, b, c); }
#include "pow_core.h"

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t next64(void) { uint64_t z = (rng_state += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
static double u01(void) { return (double)(next64() >> 11) * (1.0 / 9007199254740992.0); }

// --eval in out: `in` holds n (x, y) float pairs; `out` gets n (table result, correctly rounded result) float pairs — for tests/test_pow_gpu.py,
// which sets the DEVICE's evaluation of the same header beside them
static int eval_file(const char* in, const char* out) {
  FILE* f = fopen(in, "rb");
  if (!f) return 2;
  fseek(f, 0, SEEK_END);
  const long n = ftell(f) / 8;
  fseek(f, 0, SEEK_SET);
  float* xy = (float*)malloc((size_t)n * 8);
  float* r = (float*)malloc((size_t)n * 8);
  if (fread(xy, 8, (size_t)n, f) != (size_t)n) return 2;
  fclose(f);
  for (long i = 0; i < n; ++i) {
    float tab = 0.0f;
    if (!b2pt_pow_core(xy[2 * i], xy[2 * i + 1], &tab)) tab = __builtin_nanf("");
    float cr = (float)powq((__float128)xy[2 * i], (__float128)xy[2 * i + 1]);
    if (fabsf(cr) < 1.17549435e-38f) cr = 0.0f;   // the device flushes subnormal floats
    if (fabsf(tab) < 1.17549435e-38f) tab = 0.0f;
    r[2 * i] = tab; r[2 * i + 1] = cr;
  }
  f = fopen(out, "wb");
  if (!f) return 2;
  fwrite(r, 8, (size_t)n, f);
  fclose(f);
  return 0;
}

int main(int argc, char** argv) {
  if (argc == 4 && strcmp(argv[1], "--eval") == 0) return eval_file(argv[2], argv[3]);
  long n = argc > 1 ? atol(argv[1]) : 20000000;
  long bad_tab = 0, bad_lib = 0, differ = 0, outside = 0;
  double worst = 0.0;
  for (long it = 0; it < n; ++it) {
    float x, y;
    switch (it & 3) {
      case 0: x = (float)u01(); y = (float)(1.0 + 999.0 * u01()); break;               // any cosine, Phong exponents
      case 1: x = (float)(1.0 - 0.05 * u01() * u01()); y = (float)(1.0 + 999.0 * u01()); break;  // inside a narrow lobe
      case 2: x = (float)u01(); y = 1.0f / ((float)(1.0 + 999.0 * u01()) + 1.0f); break; // sample_phong
      default: x = b2pt_pow_u2f((uint32_t)(next64() >> 33) % 0x7f000000u + 0x00800000u); y = (float)(40.0 * (u01() - 0.5)); break;  // any normal x
    }
    if (x == 0.0f) continue;
    float tab;
    if (!b2pt_pow_core(x, y, &tab)) { ++outside; continue; }
    const float lib = (float)exp2((double)y * log2((double)x));
    const __float128 exact = powq((__float128)x, (__float128)y);
    float cr = (float)exact;  // one rounding from 113 bits
    if (fabsf(cr) < 1.17549435e-38f) { cr = 0.0f; if (fabsf(tab) < 1.17549435e-38f) tab = 0.0f; }  // the device flushes subnormal floats
    const float libf = fabsf(lib) < 1.17549435e-38f ? 0.0f : lib;
    if (tab != cr) ++bad_tab;
    if (libf != cr) ++bad_lib;
    if (tab != libf) ++differ;
  }
  printf("n %ld  outside the fast domain %ld  tables != correctly rounded: %ld (%.3g)  exp2(y*log2 x) != correctly rounded: %ld (%.3g)  tables != exp2(y*log2 x): %ld\n",
         n, outside, bad_tab, (double)bad_tab / n, bad_lib, (double)bad_lib / n, differ);
  (void)worst;
  return bad_tab > bad_lib + n / 100000 ? 1 : 0;
}
