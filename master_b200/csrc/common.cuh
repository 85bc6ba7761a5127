// common.cuh — small float3/mat3 algebra, error handling and launch helpers shared by the kernels.
//
// The library is compiled with --fmad=false -ftz=true: the reference runs on x86-64 SSE without
// FMA contraction and with FTZ/DAZ on (main.cpp:70-71), and the per-path 1e-4 parity level of
// north_star is easiest to hold when a*b+c rounds twice here as it does there.  Where fusing is
// wanted for speed and does not touch parity (BVH slab tests) the code calls __fmaf_rn itself.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/b2pt.h"

namespace b2pt {

// ---------------------------------------------------------------- errors
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define B2_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t e__ = (call);                                                \
    if (e__ != cudaSuccess) return b2pt::cuda_fail(e__, #call, __FILE__, __LINE__); \
  } while (0)

#define B2_CHECK(cond, code, msg)                  \
  do {                                             \
    if (!(cond)) { b2pt::set_error(msg); return (code); } \
  } while (0)

constexpr int kNumSMs = 148;  // B200; grids are sized in multiples of this

// ---------------------------------------------------------------- float3 algebra (glm semantics)
struct mat3 { float3 c0, c1, c2; };  // column-major like glm::mat3

#define B2_HD __host__ __device__ __forceinline__

B2_HD float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
B2_HD float3 f3(float s) { return make_float3(s, s, s); }
B2_HD float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
B2_HD float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
B2_HD float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
B2_HD float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
B2_HD float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
B2_HD float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
B2_HD float3 operator/(float3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
B2_HD float3 operator/(float3 a, float3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
B2_HD float3& operator+=(float3& a, float3 b) { a = a + b; return a; }
B2_HD float3& operator*=(float3& a, float s) { a = a * s; return a; }
B2_HD float3& operator/=(float3& a, float s) { a = a / s; return a; }
// glm::dot for vec3 (compute_dot<vec3>): tmp = a*b; tmp.x + tmp.y + tmp.z
B2_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
B2_HD float3 cross(float3 a, float3 b) {
  return f3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
B2_HD float l1norm(float3 a) { return fabsf(a.x) + fabsf(a.y) + fabsf(a.z); }
B2_HD float length2(float3 a) { return dot(a, a); }
B2_HD float distance2(float3 a, float3 b) { return length2(b - a); }

// glm::inversesqrt is 1/sqrt(x); glm::normalize is v * inversesqrt(dot(v, v))
__device__ __forceinline__ float3 normalize(float3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
__device__ __forceinline__ float length(float3 a) { return sqrtf(dot(a, a)); }

// glm mat3 * vec3 :  m[0]*v.x + m[1]*v.y + m[2]*v.z  (component-wise, left to right)
B2_HD float3 mul(const mat3& m, float3 v) {
  return f3(m.c0.x * v.x + m.c1.x * v.y + m.c2.x * v.z,
            m.c0.y * v.x + m.c1.y * v.y + m.c2.y * v.z,
            m.c0.z * v.x + m.c1.z * v.y + m.c2.z * v.z);
}
// glm vec3 * mat3 :  (dot(v, m[0]), dot(v, m[1]), dot(v, m[2]))
B2_HD float3 mul(float3 v, const mat3& m) { return f3(dot(v, m.c0), dot(v, m.c1), dot(v, m.c2)); }

B2_HD float sign_glm(float x) { return float((0.0f < x) - (x < 0.0f)); }  // glm::sign
B2_HD float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

constexpr float kPi = 3.14159265358979323846264338327950288f;
constexpr float kOneOverPi = 0.318309886183790671537767526745028724f;
constexpr float kHalfPi = 1.57079632679489661923132169163975144f;
constexpr float kTwoPi = 6.28318530717958647692528676655900576f;
constexpr float kFltEpsilon = 1.192092896e-07f;
constexpr float kFltMin = 1.175494351e-38f;

// ---------------------------------------------------------------- libm parity
// The reference calls glibc's float sin / cos / asin / atan2 / pow, which return the correctly rounded result in all
// but a vanishing fraction of cases; CUDA's float versions are 1-2 ulp functions.  A last-bit difference in a sampled
// direction flips discrete decisions further down a path (which triangle, which side of a roulette threshold), and in
// VCM / UPG a flipped photon path changes every eye path that merges with it.  Evaluating in double and rounding once
// reproduces the correctly rounded float result (double error ~1e-16 relative, far below half a float ulp except on
// exact ties) and removes most of those flips.  Shading is a small share of the step; the FP64 pipe is otherwise idle.
#ifndef B2PT_MATH_INLINE
#define B2PT_MATH_INLINE static __noinline__  // one copy per kernel: the shading stages stall on instruction fetch (40 % of warp-state samples), code size is what they pay for
#endif
__device__ __forceinline__ float cr_sinf(float x) { return float(sin(double(x))); }
__device__ __forceinline__ float cr_cosf(float x) { return float(cos(double(x))); }
__device__ __forceinline__ float cr_asinf(float x) { return float(asin(double(x))); }
__device__ __forceinline__ float cr_atan2f(float y, float x) { return float(atan2(double(y), double(x))); }
// pow for the arguments shading has (0 < x, finite): exp2(y log2 x) in double is ~2^-42 relative for the exponents used here — still
// far inside half a float ulp — at a third of the instructions of double pow (which carries an extended-precision logarithm)
__device__ B2PT_MATH_INLINE float cr_powf_lib(float x, float y) {
  if (x > 0.0f && x < INFINITY) return float(exp2(double(y) * log2(double(x))));
  return float(pow(double(x), double(y)));
}
// The same value from two small tables in a quarter of the instructions (pow_core.h; tables: tools/make_pow_tables.py).
// -DB2PT_POW_LIB=1 keeps the math-library form for A/B runs.
#ifndef B2PT_POW_LIB
#define B2PT_POW_LIB 0
#endif
#include "pow_tables.inc"
static __device__ const uint64_t kPowLogTable[128][2] = {B2PT_POW_LOG_TABLE};
static __device__ const uint64_t kPowExpTable[64] = {B2PT_POW_EXP_TABLE};
static __constant__ uint64_t kPowLogCoeffs[6] = {B2PT_POW_LOG_COEFFS};
static __constant__ uint64_t kPowExpCoeffs[5] = {B2PT_POW_EXP_COEFFS};
#define B2PT_POW_QUAL __device__ __forceinline__
#define B2PT_POW_LOG(i, w) b2pt_pow_log_word(i, w)
#define B2PT_POW_EXP(j) __ldg(&kPowExpTable[j])
#define B2PT_POW_LC(k) kPowLogCoeffs[k]
#define B2PT_POW_EC(k) kPowExpCoeffs[k]
__device__ __forceinline__ uint64_t b2pt_pow_log_word(uint32_t i, int w) {  // both words of an entry come with one 128-bit load
  const ulonglong2 e = __ldg(reinterpret_cast<const ulonglong2*>(&kPowLogTable[i][0]));
  return w ? e.y : e.x;
}
__device__ __forceinline__ double b2pt_pow_u2d(uint64_t u) { return __longlong_as_double((long long)u); }
__device__ __forceinline__ uint64_t b2pt_pow_d2u(double d) { return (uint64_t)__double_as_longlong(d); }
__device__ __forceinline__ uint32_t b2pt_pow_f2u(float f) { return __float_as_uint(f); }
__device__ __forceinline__ float b2pt_pow_u2f(uint32_t u) { return __uint_as_float(u); }
__device__ __forceinline__ double b2pt_pow_fma(double a, double b, double c) { return fma(a, b, c); }
#include "pow_core.h"
__device__ B2PT_MATH_INLINE float cr_powf_slow(float x, float y) {
#if !B2PT_POW_LIB
  float r;
  if (b2pt_pow_core(x, y, &r)) return r;
#endif
  return cr_powf_lib(x, y);
}
// The Phong lobe clamps its cosine to [0, 1] and lands ON the ends for every direction outside the lobe's hemisphere (about half of
// all BSDF queries of a connection): pow(+0, y > 0) = +0 and pow(1, y) = 1 exactly (C11 F.10.4.4), without the call — the general
// double pow those arguments used to fall into was 6 % of the connect stage's warp instructions at 9 lanes.
__device__ __forceinline__ float cr_powf(float x, float y) {
  if (x == 1.0f || y == 0.0f) return 1.0f;  // pow(1, y) = pow(x, +-0) = 1 for every x and y, NaN included: a Phong exponent of 0 is a legal material
  if (x == 0.0f && y > 0.0f && __float_as_uint(x) == 0u) return 0.0f;
  return cr_powf_slow(x, y);
}
__device__ __forceinline__ float cr_cubef(float x) { const double d = double(x); return float(d * d * d); }  // powf(x, 3.0f)
// sin and cos of one angle: one range reduction
__device__ B2PT_MATH_INLINE void cr_sincosf(float x, float& s, float& c) { double ds, dc; sincos(double(x), &ds, &dc); s = float(ds); c = float(dc); }
__device__ __forceinline__ float3 cr_polar(float r, float y, float phi) { float s, c; cr_sincosf(phi, s, c); return f3(r * c, y, r * s); }

// ---------------------------------------------------------------- warp-aggregated queue append
// One atomic per warp: ballot the lanes that append, lane 0 of the group reserves, everyone
// computes its slot from the prefix popcount.  Returns the slot (valid only where `want`).
__device__ __forceinline__ unsigned warp_append(unsigned* counter, bool want) {
  unsigned active = __activemask();
  unsigned mask = __ballot_sync(active, want);
  if (!want) return 0xFFFFFFFFu;
  unsigned lane = threadIdx.x & 31u;
  int leader = __ffs(mask) - 1;
  unsigned base = 0;
  if (lane == unsigned(leader)) base = atomicAdd(counter, unsigned(__popc(mask)));
  base = __shfl_sync(mask, base, leader);
  return base + __popc(mask & ((1u << lane) - 1u));
}

}  // namespace b2pt
