// rng.cuh — per-(pixel, sample) counter-based random streams.
//
// Replaces haste::random_generator_t (Sample.hpp:9-31, Sample.inl:249-278): the reference
// drives one std::mt19937 sequentially over all pixels (Technique.cpp:333-351), which cannot be
// split over thousands of paths in flight.  Here every path owns (key, counter); draw i of a
// path is a pure function of (key, i), so stages may run in any order and on any GPU and still
// consume the numbers in the reference's draw order.  The stream definition is shared with the
// oracle (oracle/rng_shim.h drives the unmodified reference with it; DESIGN.md "RNG").
//
// Real-valued draws reproduce libstdc++ on a 32-bit URBG:
//   uniform_real_distribution<float>  -> min(float(u) * 2^-32, nextafter(1, 0))      (1 draw)
//   piecewise_constant_distribution   -> one double from two draws (light pick)      (2 draws)
#pragma once
#include "common.cuh"

namespace b2pt {

B2_HD uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

enum { kDomainPixel = 0, kDomainPass = 1, kDomainPhoton = 2 };

B2_HD uint32_t stream_key(uint32_t seed, uint32_t domain, uint32_t index, uint32_t sample) {
  uint32_t k = mix32(seed + 0x9E3779B9U * (domain + 1U));
  k = mix32(k ^ index);
  k = mix32(k + 0x85EBCA6BU * sample);
  return k;
}

struct Rng {
  uint32_t key, ctr;
  B2_HD uint32_t next_u32() {
    uint32_t r = mix32(key ^ mix32(ctr * 0x9E3779B9U + 0x7F4A7C15U));
    ++ctr;
    return r;
  }
  // std::generate_canonical<float, 24> over a 32-bit engine, then the libstdc++ 1.0 fix-up
  __device__ __forceinline__ float sample() {
    float f = __uint2float_rn(next_u32()) * 2.3283064365386962890625e-10f;  // 2^-32, exact scale
    return f >= 1.0f ? 0.99999994f : f;
  }
  // std::generate_canonical<double, 53>: two engine calls, sum = u0 + u1 * 2^32, / 2^64
  __device__ __forceinline__ double sample_double() {
    double u0 = double(next_u32());
    double u1 = double(next_u32());
    double r = (u0 + u1 * 4294967296.0) * 5.42101086242752217003726400434970855712890625e-20;  // 2^-64
    return r >= 1.0 ? 0.99999999999999988897769753748 : r;
  }
};

}  // namespace b2pt
