// rng_shim.h — TEST INFRASTRUCTURE (oracle/_ref build only; never part of the product).
//
// Force-included (-include) in front of every reference translation unit when
// oracle/build_ref.sh compiles the reference's own sources.  It swaps the engine
// inside haste::random_generator_t (Sample.hpp:25 `std::mt19937 engine;`) for a
// counter-based 32-bit generator, so that every (pixel, sample) can be given its
// own stream: the reference has no per-pixel streams (Technique.cpp:333-351 drives
// one mt19937 across all pixels), and north_star's parity levels need them.
//
// The engine's contract towards libstdc++ is unchanged: a 32-bit URBG with
// min() = 0, max() = 2^32-1.  std::uniform_real_distribution<float> then yields
// min(float(u) * 2^-32, 0x1.fffffep-1f) and std::piecewise_constant_distribution
// draws one double from two engine calls — both restated in master_b200/csrc/rng.cuh and
// pinned by tests/test_oracle_pinning.py (numpy restatement vs `ref_harness pinning`).
//
// Stream definition (shared by the three implementations; see DESIGN.md "RNG"):
//   mix32(x):  x ^= x>>16; x *= 0x7feb352d; x ^= x>>15; x *= 0x846ca68b; x ^= x>>16
//   key(seed, domain, index, sample) =
//       mix32( mix32( mix32(seed + 0x9E3779B9*(domain+1)) ^ index ) + 0x85EBCA6B*sample )
//   draw(key, ctr) = mix32( key ^ mix32(ctr*0x9E3779B9 + 0x7F4A7C15) ),  ctr = 0, 1, 2, ...
//
// Trial mode (trial_draws = k > 0), used only by `ref_harness density`: BSDF::gathering_density
// (BSDF.cpp:42-66,264-289,393-434) is a rejection loop that makes exactly k draws per trial
// (DiffuseBSDF 2, PhongBSDF 3) from ONE generator.  The GPU runs the trials of one loop on
// many lanes at once, so it keys trial t (1, 2, ...) by its number:
//   trial_key(key, t) = mix32( key ^ mix32(t*0x9E3779B9 + 0x85EBCA6B) ),  draw j of trial t = draw(trial_key, j).
// In trial mode the engine hands the reference's sequential loop exactly those numbers
// (draw number c belongs to trial c / k + 1, position c % k), so the trial count N of a loop
// can be compared with the device's for equality.
#pragma once
#include <cstddef>
#include <cstdint>
#include <random>

namespace std {

struct b2pt_counter_engine {
  typedef uint32_t result_type;
  uint32_t key = 0;
  uint32_t ctr = 0;
  uint32_t trial_draws = 0;  // 0 = one sequential stream (everything except `ref_harness density`)

  static constexpr result_type min() { return 0u; }
  static constexpr result_type max() { return 0xFFFFFFFFu; }

  static inline uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
  }
  static inline uint32_t stream_key(uint32_t seed, uint32_t domain, uint32_t index, uint32_t sample) {
    uint32_t k = mix32(seed + 0x9E3779B9U * (domain + 1U));
    k = mix32(k ^ index);
    k = mix32(k + 0x85EBCA6BU * sample);
    return k;
  }

  b2pt_counter_engine() {}
  explicit b2pt_counter_engine(size_t s) { seed(s); }
  // the harness passes a ready-made stream key; low 32 bits are the key
  void seed(size_t s) { key = uint32_t(s); ctr = 0; }
  result_type operator()() {
    uint32_t k = key, c = ctr;
    if (trial_draws) {
      k = mix32(key ^ mix32((ctr / trial_draws + 1U) * 0x9E3779B9U + 0x85EBCA6BU));
      c = ctr % trial_draws;
    }
    uint32_t r = mix32(k ^ mix32(c * 0x9E3779B9U + 0x7F4A7C15U));
    ++ctr;
    return r;
  }
};

}  // namespace std

// every later spelling of std::mt19937 in the reference now names the engine above
#define mt19937 b2pt_counter_engine
