// grid.cuh — the photon index: a sorted cell grid in HBM with a hash table of cell ranges.
//
// Replaces v3::HashGrid3D<T> (HashGrid3D.hpp:168-325).  The reference sorts the points by (z, y, x) cell with
// std::sort on one thread, merges x-neighbour cells into ranges and keeps them in a ska::flat_hash_map keyed by the
// cell's vec3; rQuery probes the 3 x 3 (y, z) rows around the query and scans each range with a strict
// distance2 < radius^2 test (:181-204).  Here:
//   * cell = floor(p * (1 / radius)) exactly as HashGrid3D.hpp:253, packed 3 x 21 bits into one 64-bit key;
//   * the points are radix-sorted by key on the device (cub), so a cell is one contiguous run of float4
//     (x, y, z, bits(index)) — a scan of a cell is a coalesced stream of 128-bit loads;
//   * cell -> [begin, end) lives in an open-addressing table (linear probing, <= 50 % load) built with one
//     64-bit CAS per non-empty cell;
//   * a query visits the 27 cells around it (the reference's 9 x-merged ranges cover the same 27 cells) and
//     reports exactly the points with distance2 < radius^2 among them, which is the reference's result set.
// Keys wrap modulo 2^21 cells per axis; a wrapped collision can only add candidates, never results, because every
// candidate is distance-tested.
#pragma once
#include "common.cuh"

namespace b2pt {

struct DeviceGrid {
  uint32_t num_points = 0;
  uint32_t table_mask = 0;             // table size - 1 (power of two)
  float radius = 0.f, radius_inv = 0.f;
  const float4* points = nullptr;      // sorted by cell key: xyz, bits(caller's index)
  const unsigned long long* table_keys = nullptr;  // kEmptyCell where unused
  const uint2* table_ranges = nullptr; // begin, end into `points`
};

constexpr unsigned long long kEmptyCell = ~0ull;

__host__ __device__ __forceinline__ unsigned long long cell_key(int x, int y, int z) {
  return (unsigned long long)(uint32_t(x) & 0x1FFFFFu) | ((unsigned long long)(uint32_t(y) & 0x1FFFFFu) << 21) |
         ((unsigned long long)(uint32_t(z) & 0x1FFFFFu) << 42);
}

__host__ __device__ __forceinline__ uint32_t cell_hash(unsigned long long k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
  return uint32_t(k);
}

// floor(p * radius_inv) per component (HashGrid3D.hpp:188,253), as integers
__device__ __forceinline__ int3 grid_cell(const DeviceGrid& g, float3 p) {
  return make_int3(int(floorf(p.x * g.radius_inv)), int(floorf(p.y * g.radius_inv)), int(floorf(p.z * g.radius_inv)));
}

__device__ __forceinline__ uint2 grid_lookup(const DeviceGrid& g, unsigned long long key) {
  uint32_t slot = cell_hash(key) & g.table_mask;
  for (;;) {
    const unsigned long long k = __ldg(&g.table_keys[slot]);
    if (k == key) return __ldg(&g.table_ranges[slot]);
    if (k == kEmptyCell) return make_uint2(0u, 0u);
    slot = (slot + 1u) & g.table_mask;
  }
}

// rQuery (HashGrid3D.hpp:181-204): f(index, position) for every point with distance2(query, point) < radius_sq
// in the 27 cells around the query.  Row order z, y like the reference; inside a row x - 1, x, x + 1.
template <class F>
__device__ __forceinline__ void grid_query(const DeviceGrid& g, float3 q, float radius_sq, F&& f) {
  if (g.num_points == 0) return;
  const int3 c = grid_cell(g, q);
  for (int dz = -1; dz <= 1; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        const uint2 r = grid_lookup(g, cell_key(c.x + dx, c.y + dy, c.z + dz));
        for (uint32_t i = r.x; i < r.y; ++i) {
          const float4 p = __ldg(&g.points[i]);
          if (distance2(q, f3(p.x, p.y, p.z)) < radius_sq) f(__float_as_uint(p.w), f3(p.x, p.y, p.z));
        }
      }
}

// host-side owner; build is asynchronous on `stream`, buffers grow on demand and are reused between builds
struct GridStorage {
  DeviceGrid dev;
  size_t capacity = 0, table_capacity = 0, temp_bytes = 0;
  unsigned long long *keys_in = nullptr, *keys_out = nullptr, *table_keys = nullptr;
  float4 *pts_out = nullptr;
  uint32_t *idx_in = nullptr, *idx_out = nullptr;
  uint2* table_ranges = nullptr;
  void* temp = nullptr;
  void release();
};

// points_in: n float4 (xyz, bits(index)) in DEVICE memory; entries whose w is kInvalidPoint are left out of the grid
constexpr uint32_t kInvalidPoint = 0xFFFFFFFFu;
int grid_build_device(GridStorage& st, const float4* points_in, uint32_t n, float radius, cudaStream_t stream);

}  // namespace b2pt
