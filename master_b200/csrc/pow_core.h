// pow_core.h — x^y for positive normal float x and finite float y, correctly rounded to float in all but a vanishing fraction of
// cases: 2^(y log2 x) in double from two small tables (tools/make_pow_tables.py -> pow_tables.inc).
//
// Why: the shading stages call powf for every Phong lobe they sample or query (PhongBSDF::_query, BSDF.cpp:349-391;
// sample_phong, Sample.inl:137-149).  The reference gets glibc's float powf, which is within 0.52 ulp; to land on the same float
// the device evaluates in double and rounds once (common.cuh).  Through the CUDA math library that is exp2(y * log2(x)) = 192
// instructions, a quarter of them moves of 64-bit polynomial constants: 18 % of the logic stage's and 31 % of the connect stage's
// warp instructions (profiles/r02_shading_lines_v5.txt).  The same value from tables takes about 50.
//
//   log2 x:  x = 2^k z with z in [0.699, 1.398) (bit pattern rebased by 0x3f330000, so x near 1 has k = 0 and no cancellation);
//            the top 7 bits below pick an interval with centre c:  log2 z = log2 c + log2(1 + r),  r = fma(z, 1/c, -1), |r| < 2^-8 (exact to 2^-61),
//            log2(1 + r) = r (L1 + r (L2 + ... + L6 r^5)),  truncation < 2^-58.
//   2^E:     E = y log2 x = n + j/64 + t, |t| <= 2^-7:  2^E = 2^n 2^(j/64) (1 + t (E1 + ... + E5 t^4)),  truncation < 2^-54.
// Absolute error of E: 2^-53 (|k + log2 c| + |E|), i.e. < 2^-46 for the exponents shading has (|E| < 64) — the relative error of the
// result before the one rounding to float; exp2(y * log2(x)) through the math library is no better (its log2 is rounded to double too).
// tools/check_pow.c compares this function, compiled for the host, with powq (libquadmath) on 10^8 arguments.
//
// The including side defines: B2PT_POW_QUAL (function qualifiers), B2PT_POW_LOG(i, which) / B2PT_POW_EXP(j) / B2PT_POW_LC(k) /
// B2PT_POW_EC(k) (table words as uint64), b2pt_pow_u2d / b2pt_pow_d2u / b2pt_pow_f2u / b2pt_pow_u2f (bit casts), b2pt_pow_fma.
#pragma once
#include <stdint.h>

// returns 0 when (x, y) is outside the fast domain (x not a positive normal float, y not finite) — the caller falls back
B2PT_POW_QUAL int b2pt_pow_core(float x, float y, float* out) {
  const uint32_t ix = b2pt_pow_f2u(x), iy = b2pt_pow_f2u(y);
  if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) return 0;
  if ((iy & 0x7fffffffu) >= 0x7f800000u) return 0;
  // ---- log2 x
  const uint32_t tmp = ix - 0x3f330000u;
  const uint32_t i = (tmp >> 16) & 127u;
  const uint32_t top = tmp & 0xff800000u;
  const int32_t k = (int32_t)top >> 23;
  const double z = (double)b2pt_pow_u2f(ix - top);
  const double invc = b2pt_pow_u2d(B2PT_POW_LOG(i, 0)), logc = b2pt_pow_u2d(B2PT_POW_LOG(i, 1));
  const double r = b2pt_pow_fma(z, invc, -1.0);
  double p = b2pt_pow_u2d(B2PT_POW_LC(5));
  p = b2pt_pow_fma(p, r, b2pt_pow_u2d(B2PT_POW_LC(4)));
  p = b2pt_pow_fma(p, r, b2pt_pow_u2d(B2PT_POW_LC(3)));
  p = b2pt_pow_fma(p, r, b2pt_pow_u2d(B2PT_POW_LC(2)));
  p = b2pt_pow_fma(p, r, b2pt_pow_u2d(B2PT_POW_LC(1)));
  p = b2pt_pow_fma(p, r, b2pt_pow_u2d(B2PT_POW_LC(0)));
  const double l = b2pt_pow_fma(p, r, logc + (double)k);
  // ---- 2^(y l)
  const double e = (double)y * l;
  if (!(e > -200.0)) { *out = 0.0f; return 1; }            // far below the smallest float (and e = -inf never gets here: y, l finite)
  if (!(e < 200.0)) { *out = b2pt_pow_u2f(0x7f800000u); return 1; }
  const double shift = 105553116266496.0;                  // 1.5 * 2^52 / 64: adding it rounds e to a multiple of 1/64
  double kd = e + shift;
  const uint64_t ki = b2pt_pow_d2u(kd);
  kd -= shift;
  const double t = e - kd;
  const uint64_t sbits = B2PT_POW_EXP(ki & 63u) + ((ki >> 6) << 52);  // the low bits of ki are round(64 e) in two's complement
  const double s = b2pt_pow_u2d(sbits);
  double q = b2pt_pow_u2d(B2PT_POW_EC(4));
  q = b2pt_pow_fma(q, t, b2pt_pow_u2d(B2PT_POW_EC(3)));
  q = b2pt_pow_fma(q, t, b2pt_pow_u2d(B2PT_POW_EC(2)));
  q = b2pt_pow_fma(q, t, b2pt_pow_u2d(B2PT_POW_EC(1)));
  q = b2pt_pow_fma(q, t, b2pt_pow_u2d(B2PT_POW_EC(0)));
  *out = (float)b2pt_pow_fma(s * t, q, s);
  return 1;
}
