// film_errors_static.cu — b2pt_film_errors_static: the error record of a pass (b2pt_film_errors, render.cu) for callers that
// may not allocate: CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// b2pt_film_errors takes its three partial sums from the driver's stream-ordered pool (cudaMallocAsync) and reads them back to the
// host.  This variant allocates nothing and never touches the host: the per-tile partial sums live in ONE file-scope __device__
// array (below), the last CTA to finish adds them in a fixed order and writes (rms, abs) to a DEVICE buffer of the caller.
// The array is part of the module image: every GPU that loads libb2pt.so holds its 98 560 bytes until the process exits, every call on
// a device shares it, and it keeps the last call's contents — so two calls that overlap on one GPU (different streams, no event between
// them) race, and the arrival counter, the only word that must start at zero, is cleared on the caller's stream before every launch.
#include <mutex>

#include "common.cuh"

namespace b2pt {
namespace {

constexpr uint32_t kTilePixels = 4096;                                   // pixels one CTA sums: 256 threads x 16
constexpr uint32_t kMaxTiles = B2PT_FILM_ERRORS_STATIC_MAX_PIXELS / kTilePixels;  // 4096
constexpr uint32_t kCounterAt = kMaxTiles * 3;                           // doubles; the arrival counter sits in its own 256-byte line
constexpr uint32_t kScratchDoubles = kCounterAt + 32;                    // 12 320 doubles = 98 560 bytes

// [tile][sum of dot(d, d), sum of d.x + d.y + d.z, pixels], then the arrival counter.  In HOST code this name is a host-side
// stand-in, not a device address: the kernels name it directly, the host only uses its address (cudaGetSymbolAddress below) to clear
// the counter.
static __device__ __align__(256) double g_film_errors_scratch[kScratchDoubles];

// The per-pixel terms and the final division are b2pt_film_errors's (render.cu: film_errors_kernel, ImageView.cpp:60-117).
__global__ void __launch_bounds__(256) film_errors_static_kernel(const double* __restrict__ accum, const float* __restrict__ ref, uint32_t width,
                                                                 uint32_t height, int x0, int y0, int x1, int y1, int windowed,
                                                                 float* __restrict__ out) {
  const uint32_t w = uint32_t(x1 - x0), n = w * uint32_t(y1 - y0);
  double sq = 0.0, ab = 0.0, cnt = 0.0;
  const uint32_t tile_end = min(n, (blockIdx.x + 1u) * kTilePixels);
  for (uint32_t i = blockIdx.x * kTilePixels + threadIdx.x; i < tile_end; i += blockDim.x) {
    const size_t p = size_t(y0 + int(i / w)) * width + size_t(x0 + int(i % w));
    const double2 rg = *reinterpret_cast<const double2*>(accum + p * 4), bw = *reinterpret_cast<const double2*>(accum + p * 4 + 2);
    const float dx = fabsf(float(rg.x / bw.y) - ref[p * 3 + 0]);
    const float dy = fabsf(float(rg.y / bw.y) - ref[p * 3 + 1]);
    const float dz = fabsf(float(bw.x / bw.y) - ref[p * 3 + 2]);
    ab += double(__fadd_rn(__fadd_rn(dx, dy), dz));
    sq += double(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    cnt += 1.0;
  }
  __shared__ double part[3][8];
  __shared__ bool last;
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  for (int o = 16; o > 0; o >>= 1) {
    sq += __shfl_down_sync(0xFFFFFFFFu, sq, o); ab += __shfl_down_sync(0xFFFFFFFFu, ab, o); cnt += __shfl_down_sync(0xFFFFFFFFu, cnt, o);
  }
  if (lane == 0) { part[0][warp] = sq; part[1][warp] = ab; part[2][warp] = cnt; }
  __syncthreads();
  unsigned* counter = reinterpret_cast<unsigned*>(g_film_errors_scratch + kCounterAt);
  if (threadIdx.x == 0) {
    for (int k = 0; k < 3; ++k) {
      double t = 0.0;
      for (int j = 0; j < 8; ++j) t += part[k][j];
      g_film_errors_scratch[blockIdx.x * 3u + k] = t;
    }
    __threadfence();
    last = atomicAdd(counter, 1u) == gridDim.x - 1u;
  }
  __syncthreads();
  if (!last) return;
  // ---- the last CTA to arrive: every tile's partial sums, in tile order per thread, then the same tree as above
  __threadfence();
  sq = ab = cnt = 0.0;
  for (uint32_t t = threadIdx.x; t < gridDim.x; t += blockDim.x) {
    sq += __ldcg(g_film_errors_scratch + t * 3u + 0); ab += __ldcg(g_film_errors_scratch + t * 3u + 1); cnt += __ldcg(g_film_errors_scratch + t * 3u + 2);
  }
  for (int o = 16; o > 0; o >>= 1) {
    sq += __shfl_down_sync(0xFFFFFFFFu, sq, o); ab += __shfl_down_sync(0xFFFFFFFFu, ab, o); cnt += __shfl_down_sync(0xFFFFFFFFu, cnt, o);
  }
  __syncthreads();
  if (lane == 0) { part[0][warp] = sq; part[1][warp] = ab; part[2][warp] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s[3] = {0.0, 0.0, 0.0};
    for (int k = 0; k < 3; ++k) for (int j = 0; j < 8; ++j) s[k] += part[k][j];
    // whole image: num = width * height * 3 (ImageView.cpp:81-84); window: num = pixels (ImageView.cpp:109-116); an empty window is 0 / 0
    const float num = windowed ? float(s[2]) : float(size_t(width) * height * 3);
    out[0] = sqrtf(float(s[0]) / num);
    out[1] = float(s[1]) / num;
    *counter = 0u;  // leave the counter as the next call expects it (the host clears it as well)
  }
}

// device address of the scratch array per device ordinal, resolved once with that device current
std::mutex g_scratch_mutex;
double* g_scratch_ptr[64] = {};

}  // namespace
}  // namespace b2pt

extern "C" int b2pt_film_errors_static(const double* accum_dev, const float* reference_dev, uint32_t width, uint32_t height,
                                       const int32_t* window, float* out_dev, void* stream_) {
  using namespace b2pt;
  B2_CHECK(accum_dev && reference_dev && out_dev, B2PT_ERR_INVALID, "b2pt_film_errors_static: null argument");
  B2_CHECK(width > 0 && height > 0, B2PT_ERR_INVALID, "b2pt_film_errors_static: empty image");
  int x0 = 0, y0 = 0, x1 = int(width), y1 = int(height);
  if (window) {  // ImageView.cpp:103-104
    x0 = max(0, window[0] - window[2] + 1); x1 = min(window[0] + window[2], int(width));
    y0 = max(0, window[1] - window[2] + 1); y1 = min(window[1] + window[2], int(height));
  }
  const bool empty = !(x1 > x0 && y1 > y0);
  const size_t n = empty ? 0 : size_t(x1 - x0) * size_t(y1 - y0);
  if (n > size_t(B2PT_FILM_ERRORS_STATIC_MAX_PIXELS)) {
    char msg[200];
    snprintf(msg, sizeof(msg), "b2pt_film_errors_static: %zu pixels to sum, the static scratch array holds partial sums for at most %u "
             "(use b2pt_film_errors, or a window)", n, unsigned(B2PT_FILM_ERRORS_STATIC_MAX_PIXELS));
    B2_CHECK(false, B2PT_ERR_CAPACITY, msg);
  }
  int dev_count = 0;
  if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0) { cudaGetLastError(); B2_CHECK(false, B2PT_ERR_NO_DEVICE, "b2pt_film_errors_static: no CUDA device"); }
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  B2_CHECK(dev >= 0 && dev < 64, B2PT_ERR_INVALID, "b2pt_film_errors_static: device ordinal out of range");
  double* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[dev]) {
      double* p = nullptr;
      B2_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_film_errors_scratch));  // this device's copy: `dev` is current
      g_scratch_ptr[dev] = p;
    }
    scratch = g_scratch_ptr[dev];
  }
  cudaStream_t stream = (cudaStream_t)stream_;
  if (empty) { x0 = y0 = 0; x1 = y1 = 0; }
  const unsigned grid = unsigned(n ? (n + kTilePixels - 1) / kTilePixels : 1);
  B2_CUDA(cudaMemsetAsync(scratch + kCounterAt, 0, 256, stream));  // the arrival counter must start at zero whatever the last call left
  film_errors_static_kernel<<<grid, 256, 0, stream>>>(accum_dev, reference_dev, width, height, x0, y0, x1, y1, window ? 1 : 0, out_dev);
  B2_CUDA(cudaGetLastError());
  return B2PT_OK;
}
