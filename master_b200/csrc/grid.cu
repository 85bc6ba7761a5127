// grid.cu — build of the sorted cell grid (grid.cuh) and the b2pt_grid_* entry points
// (v3::HashGrid3D<T>::build / rQuery, HashGrid3D.hpp:181-325).
#include <cub/cub.cuh>

#include "grid.cuh"

namespace b2pt {

void GridStorage::release() {
  cudaFree(keys_in); cudaFree(keys_out); cudaFree(table_keys); cudaFree(pts_out); cudaFree(idx_in); cudaFree(idx_out);
  cudaFree(table_ranges); cudaFree(temp);
  *this = GridStorage();
}

namespace {

__global__ void grid_keys_kernel(uint32_t n, const float4* __restrict__ pts, float radius_inv, unsigned long long* keys, uint32_t* idx) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = pts[i];
  idx[i] = i;
  if (__float_as_uint(p.w) == kInvalidPoint) { keys[i] = kEmptyCell; return; }
  // floor(position * radius_inv) (HashGrid3D.hpp:253)
  keys[i] = cell_key(int(floorf(p.x * radius_inv)), int(floorf(p.y * radius_inv)), int(floorf(p.z * radius_inv)));
}

__global__ void grid_gather_kernel(uint32_t n, const float4* __restrict__ pts, const uint32_t* __restrict__ order, float4* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = pts[order[i]];
}

// one 64-bit CAS per non-empty cell: the thread at the first point of a run claims a slot and writes `begin`
__global__ void grid_insert_kernel(uint32_t n, const unsigned long long* __restrict__ keys, unsigned long long* table_keys,
                                   uint2* table_ranges, uint32_t mask) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long k = keys[i];
  if (k == kEmptyCell || (i > 0 && keys[i - 1] == k)) return;
  uint32_t end = i + 1;  // runs are short (a cell holds tens of points): walk to the end of the run
  while (end < n && keys[end] == k) ++end;
  uint32_t slot = cell_hash(k) & mask;
  for (;;) {
    const unsigned long long prev = atomicCAS(&table_keys[slot], kEmptyCell, k);
    if (prev == kEmptyCell) { table_ranges[slot] = make_uint2(i, end); return; }
    slot = (slot + 1u) & mask;
  }
}

template <class T>
int grow(T** p, size_t count) {
  if (*p) cudaFree(*p);
  *p = nullptr;
  B2_CUDA(cudaMalloc((void**)p, count * sizeof(T)));
  return B2PT_OK;
}

}  // namespace

int grid_build_device(GridStorage& st, const float4* points_in, uint32_t n, float radius, cudaStream_t stream) {
  B2_CHECK(radius > 0.0f, B2PT_ERR_INVALID, "grid: radius must be positive");
  st.dev.radius = radius;
  st.dev.radius_inv = 1.0f / radius;  // HashGrid3D.hpp:247
  st.dev.num_points = n;
  if (n == 0) { st.dev.points = nullptr; st.dev.table_mask = 0; return B2PT_OK; }
  if (n > st.capacity) {
    const size_t cap = size_t(n) + n / 4 + 1024;
    int rc;
    if ((rc = grow(&st.keys_in, cap))) return rc;
    if ((rc = grow(&st.keys_out, cap))) return rc;
    if ((rc = grow(&st.idx_in, cap))) return rc;
    if ((rc = grow(&st.idx_out, cap))) return rc;
    if ((rc = grow(&st.pts_out, cap))) return rc;
    st.capacity = cap;
    size_t bytes = 0;
    B2_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, bytes, st.keys_in, st.keys_out, st.idx_in, st.idx_out, int(cap)));
    if (bytes > st.temp_bytes) {
      if (st.temp) cudaFree(st.temp);
      st.temp = nullptr;
      B2_CUDA(cudaMalloc(&st.temp, bytes));
      st.temp_bytes = bytes;
    }
    size_t table = 1024;
    while (table < 2 * cap) table <<= 1;
    if ((rc = grow(&st.table_keys, table))) return rc;
    if ((rc = grow(&st.table_ranges, table))) return rc;
    st.table_capacity = table;
  }
  // table sized for this build: a power of two >= 2 n, within the allocation
  size_t table = 1024;
  while (table < 2 * size_t(n)) table <<= 1;
  const uint32_t mask = uint32_t(table - 1);
  const int B = 256;
  const uint32_t G = (n + B - 1) / B;
  grid_keys_kernel<<<G, B, 0, stream>>>(n, points_in, st.dev.radius_inv, st.keys_in, st.idx_in);
  size_t bytes = st.temp_bytes;
  B2_CUDA(cub::DeviceRadixSort::SortPairs(st.temp, bytes, st.keys_in, st.keys_out, st.idx_in, st.idx_out, int(n), 0, 64, stream));
  grid_gather_kernel<<<G, B, 0, stream>>>(n, points_in, st.idx_out, st.pts_out);
  B2_CUDA(cudaMemsetAsync(st.table_keys, 0xFF, table * sizeof(unsigned long long), stream));
  grid_insert_kernel<<<G, B, 0, stream>>>(n, st.keys_out, st.table_keys, st.table_ranges, mask);
  B2_CUDA(cudaGetLastError());
  st.dev.points = st.pts_out;
  st.dev.table_keys = st.table_keys;
  st.dev.table_ranges = st.table_ranges;
  st.dev.table_mask = mask;
  return B2PT_OK;
}

namespace {

__global__ void grid_query_kernel(DeviceGrid g, const float* __restrict__ queries, uint32_t nq, uint32_t max_per_query, uint32_t* counts,
                                  uint32_t* indices) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  const float3 q = f3(queries[i * 3 + 0], queries[i * 3 + 1], queries[i * 3 + 2]);
  uint32_t found = 0;
  grid_query(g, q, g.radius * g.radius, [&](uint32_t index, float3) {
    if (found < max_per_query) indices[size_t(i) * max_per_query + found] = index;
    ++found;
  });
  counts[i] = found;
}

__global__ void grid_pack_kernel(uint32_t n, const float* __restrict__ xyz, float4* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = make_float4(xyz[i * 3 + 0], xyz[i * 3 + 1], xyz[i * 3 + 2], __uint_as_float(i));
}

}  // namespace

}  // namespace b2pt

struct b2pt_grid_s {
  b2pt::GridStorage storage;
  float build_ms = 0.f;  // device time of the build (keys, radix sort, gather, table inserts), points already in HBM
};

extern "C" {

int b2pt_grid_build(const float* points, size_t n, float radius, b2pt_grid_t* out) {
  using namespace b2pt;
  B2_CHECK(out && (n == 0 || points), B2PT_ERR_INVALID, "b2pt_grid_build: null argument");
  B2_CHECK(n < 0xFFFFFFFFull, B2PT_ERR_CAPACITY, "b2pt_grid_build: too many points");
  B2_CHECK(radius > 0.0f, B2PT_ERR_INVALID, "b2pt_grid_build: radius must be positive");
  int devices = 0;
  B2_CHECK(cudaGetDeviceCount(&devices) == cudaSuccess && devices > 0, B2PT_ERR_NO_DEVICE,
           "no CUDA device: the B200 path has no CPU fallback");
  b2pt_grid_s* g = new b2pt_grid_s();
  int rc = B2PT_OK;
  if (n > 0) {
    float* d_xyz = nullptr; float4* d_pts = nullptr;
    cudaError_t e = cudaMalloc(&d_xyz, n * 12);
    if (e == cudaSuccess) e = cudaMalloc(&d_pts, n * 16);
    if (e == cudaSuccess) e = cudaMemcpy(d_xyz, points, n * 12, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
      grid_pack_kernel<<<unsigned((n + 255) / 256), 256>>>(uint32_t(n), d_xyz, d_pts);
      cudaEvent_t e0 = nullptr, e1 = nullptr;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0, nullptr);
      rc = grid_build_device(g->storage, d_pts, uint32_t(n), radius, nullptr);
      cudaEventRecord(e1, nullptr);
      if (rc == B2PT_OK) e = cudaDeviceSynchronize();
      if (rc == B2PT_OK && e == cudaSuccess) cudaEventElapsedTime(&g->build_ms, e0, e1);
      cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
    cudaFree(d_xyz); cudaFree(d_pts);
    if (e != cudaSuccess) rc = cuda_fail(e, "b2pt_grid_build", __FILE__, __LINE__);
  } else {
    rc = grid_build_device(g->storage, nullptr, 0, radius, nullptr);
  }
  if (rc != B2PT_OK) { g->storage.release(); delete g; return rc; }
  *out = g;
  return B2PT_OK;
}

void b2pt_grid_destroy(b2pt_grid_t grid) {
  if (!grid) return;
  grid->storage.release();
  delete grid;
}

int b2pt_grid_info(b2pt_grid_t grid, uint64_t* num_points, float* radius, float* build_ms) {
  B2_CHECK(grid, B2PT_ERR_INVALID, "b2pt_grid_info: null grid");
  if (num_points) *num_points = grid->storage.dev.num_points;
  if (radius) *radius = grid->storage.dev.radius;
  if (build_ms) *build_ms = grid->build_ms;
  return B2PT_OK;
}

int b2pt_grid_query_device(b2pt_grid_t grid, const float* queries_dev, size_t nq, uint32_t max_per_query, uint32_t* counts_dev,
                           uint32_t* indices_dev, void* stream) {
  using namespace b2pt;
  B2_CHECK(grid && (nq == 0 || (queries_dev && counts_dev)) && (max_per_query == 0 || indices_dev), B2PT_ERR_INVALID, "b2pt_grid_query_device: null argument");
  if (nq == 0) return B2PT_OK;
  B2_CHECK(nq < 0xFFFFFFFFull, B2PT_ERR_CAPACITY, "b2pt_grid_query_device: too many queries");
  grid_query_kernel<<<unsigned((nq + 127) / 128), 128, 0, (cudaStream_t)stream>>>(grid->storage.dev, queries_dev, uint32_t(nq), max_per_query, counts_dev, indices_dev);
  B2_CUDA(cudaGetLastError());
  return B2PT_OK;
}

int b2pt_grid_query(b2pt_grid_t grid, const float* queries, size_t nq, uint32_t max_per_query, uint32_t* counts, uint32_t* indices) {
  using namespace b2pt;
  B2_CHECK(grid && counts && (nq == 0 || queries) && (max_per_query == 0 || indices), B2PT_ERR_INVALID, "b2pt_grid_query: null argument");
  if (nq == 0) return B2PT_OK;
  B2_CHECK(nq < 0xFFFFFFFFull, B2PT_ERR_CAPACITY, "b2pt_grid_query: too many queries");
  float* d_q = nullptr; uint32_t *d_counts = nullptr, *d_idx = nullptr;
  const size_t slots = nq * size_t(max_per_query ? max_per_query : 1);
  cudaError_t e = cudaMalloc(&d_q, nq * 12);
  if (e == cudaSuccess) e = cudaMalloc(&d_counts, nq * 4);
  if (e == cudaSuccess) e = cudaMalloc(&d_idx, slots * 4);
  if (e == cudaSuccess) e = cudaMemcpy(d_q, queries, nq * 12, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    grid_query_kernel<<<unsigned((nq + 127) / 128), 128>>>(grid->storage.dev, d_q, uint32_t(nq), max_per_query, d_counts, d_idx);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(counts, d_counts, nq * 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && max_per_query) e = cudaMemcpy(indices, d_idx, slots * 4, cudaMemcpyDeviceToHost);
  cudaFree(d_q); cudaFree(d_counts); cudaFree(d_idx);
  if (e != cudaSuccess) return cuda_fail(e, "b2pt_grid_query", __FILE__, __LINE__);
  return B2PT_OK;
}

}  // extern "C"
