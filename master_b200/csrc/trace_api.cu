// trace_api.cu — b2pt_trace / b2pt_trace_device: batches of explicit rays.
//
// The ray-level boundary: what Scene::intersect / intersectMesh / occluded hand to Embree
// (an RTCRay with org, dir, tnear, tfar and a mask, Scene.cpp:162-171,186-196,209-219) and what
// they read back (tfar, u, v, geomID, primID).  Used by the ray-level parity tests and the
// Mrays/s micro-benchmark; the renderers call trace_ray() from their own stages.
#include <stdlib.h>

#include "trace.cuh"

namespace b2pt {

// explicit rays in, RTCRay-style hits out
template <bool kAnyHit, bool kMeshOnly>
struct ApiIo {
  const DeviceScene* sc;
  const float4* rays;
  B2ptHit* hits;
  __device__ __forceinline__ uint32_t load(uint32_t i, float3& O, float3& D, float& tnear, float& tfar, uint32_t& mask) const {
    const float4 a = __ldg(&rays[size_t(i) * 2 + 0]);  // org.xyz, tnear
    const float4 b = __ldg(&rays[size_t(i) * 2 + 1]);  // dir.xyz, tfar
    O = f3(a.x, a.y, a.z); D = f3(b.x, b.y, b.z); tnear = a.w; tfar = b.w;
    mask = kMeshOnly ? kMaskMesh : kMaskAll;
    return i;
  }
  __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
    B2ptHit o;
    o.t = h.t; o.u = h.u; o.v = h.v;
    if (h.tri == kInvalid) { o.geom = kInvalid; o.prim = kInvalid; }
    else if (kAnyHit) { o.geom = 0u; o.prim = kInvalid; }  // rtcOccluded reports geomID = 0 (Scene.cpp:179)
    else { o.geom = tri_geom(*sc, h.tri); o.prim = tri_prim(*sc, h.tri); }
    hits[i] = o;
  }
};

// persistent warps, dynamic ray fetch, warp-synchronous rounds (trace.cuh: trace_stream8); B2PT_TRACE_STREAM=1
template <bool kAnyHit, bool kMeshOnly>
__global__ void __launch_bounds__(256) trace_stream8_kernel(DeviceScene sc, const float4* __restrict__ rays, uint32_t n,
                                                            B2ptHit* __restrict__ hits, unsigned* ticket) {
  ApiIo<kAnyHit, kMeshOnly> io{&sc, rays, hits};
  trace_stream8<kAnyHit>(sc, n, ticket, io);
}

// persistent warps, dynamic ray fetch, free-running lanes with postponed triangle groups (trace.cuh: trace_dyn8); B2PT_TRACE_STREAM=2
template <bool kAnyHit, bool kMeshOnly>
__global__ void __launch_bounds__(256, 4) trace_dyn8_kernel(DeviceScene sc, const float4* __restrict__ rays, uint32_t n,
                                                            B2ptHit* __restrict__ hits, unsigned* ticket) {
  ApiIo<kAnyHit, kMeshOnly> io{&sc, rays, hits};
  trace_dyn8<kAnyHit>(sc, n, ticket, io);
}

// persistent warps, warp-synchronous rounds, shared triangle queue (trace.cuh: trace_queue8); B2PT_TRACE_STREAM=3
template <bool kAnyHit, bool kMeshOnly>
__global__ void __launch_bounds__(256, 4) trace_queue8_kernel(DeviceScene sc, const float4* __restrict__ rays, uint32_t n,
                                                              B2ptHit* __restrict__ hits, unsigned* ticket) {
  ApiIo<kAnyHit, kMeshOnly> io{&sc, rays, hits};
  trace_queue8<kAnyHit>(sc, n, ticket, io);
}

// one ray per thread, run to completion (trace.cuh: trace_ray8): small batches
template <bool kAnyHit, bool kMeshOnly>
__global__ void __launch_bounds__(256, 5) trace_rays_kernel(DeviceScene sc, const float4* __restrict__ rays, uint32_t n,
                                                            B2ptHit* __restrict__ hits) {
  ApiIo<kAnyHit, kMeshOnly> io{&sc, rays, hits};
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float3 O, D; float tnear, tfar; uint32_t mask;
    io.load(i, O, D, tnear, tfar, mask);
    io.store(i, trace_any_bvh<kAnyHit>(sc, O, D, tnear, tfar, mask));
  }
}

int launch_trace(b2pt_scene_s* scene, const B2ptRay* rays, size_t n, unsigned mode, B2ptHit* hits, cudaStream_t stream) {
  const DeviceScene& sc = scene->dev;
  if (n == 0) return B2PT_OK;
  B2_CHECK(n < (size_t(1) << 31), B2PT_ERR_CAPACITY, "b2pt_trace: more than 2^31 rays in one call");
  B2_CHECK(mode <= B2PT_TRACE_OCCLUDED_MESH, B2PT_ERR_INVALID, "b2pt_trace: unknown mode");
  const int block = 256;
  scene_l2_window(scene, stream);
  const float4* r = reinterpret_cast<const float4*>(rays);
  // The default is trace_queue8 for batches that fill the persistent grid a few times over, one ray per thread below that;
  // B2PT_TRACE_STREAM = 0 | 1 | 2 | 3 forces one form (A/B: tools/trace_bench.py)
  const char* env = getenv("B2PT_TRACE_STREAM");
  const bool queue_ok = sc.nodes8 && sc.num_tris < (1u << 27);
  char form[2] = {env ? env[0] : (queue_ok && n >= (size_t(1) << 18) ? '3' : '0'), 0};
  if (form[0] == '3' && !queue_ok) form[0] = '2';
  const char* stream_env = form;
  if (!(stream_env[0] == '1' || stream_env[0] == '2' || stream_env[0] == '3') || !sc.nodes8) {
    const int grid = grid_for_trace(n, block);
    switch (mode) {
      case B2PT_TRACE_CLOSEST_ALL: trace_rays_kernel<false, false><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits); break;
      case B2PT_TRACE_CLOSEST_MESH: trace_rays_kernel<false, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits); break;
      default: trace_rays_kernel<true, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits); break;
    }
  } else if (stream_env[0] == '3') {
    unsigned* ticket = scene->dev.error_flags + 1;
    B2_CUDA(cudaMemsetAsync(ticket, 0, 4, stream));
    const int grid = persistent_grid(n, block);
    switch (mode) {
      case B2PT_TRACE_CLOSEST_ALL: trace_queue8_kernel<false, false><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
      case B2PT_TRACE_CLOSEST_MESH: trace_queue8_kernel<false, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
      default: trace_queue8_kernel<true, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
    }
  } else if (stream_env[0] == '2' || stream_env[0] == '3') {
    unsigned* ticket = scene->dev.error_flags + 1;
    B2_CUDA(cudaMemsetAsync(ticket, 0, 4, stream));
    const int grid = persistent_grid(n, block);
    switch (mode) {
      case B2PT_TRACE_CLOSEST_ALL: trace_dyn8_kernel<false, false><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
      case B2PT_TRACE_CLOSEST_MESH: trace_dyn8_kernel<false, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
      default: trace_dyn8_kernel<true, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
    }
  } else {
    unsigned* ticket = scene->dev.error_flags + 1;  // the scene's ticket counter, zeroed per launch
    B2_CUDA(cudaMemsetAsync(ticket, 0, 4, stream));
    const int grid = persistent_grid(n, block);
    switch (mode) {
      case B2PT_TRACE_CLOSEST_ALL: trace_stream8_kernel<false, false><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
      case B2PT_TRACE_CLOSEST_MESH: trace_stream8_kernel<false, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
      default: trace_stream8_kernel<true, true><<<grid, block, 0, stream>>>(sc, r, uint32_t(n), hits, ticket); break;
    }
  }
  B2_CUDA(cudaGetLastError());
  return B2PT_OK;
}

}  // namespace b2pt

extern "C" {

int b2pt_trace_device(b2pt_scene_t scene, const B2ptRay* rays_dev, size_t n, unsigned mode, B2ptHit* hits_dev, void* stream) {
  B2_CHECK(scene && (n == 0 || (rays_dev && hits_dev)), B2PT_ERR_INVALID, "b2pt_trace_device: null argument");
  return b2pt::launch_trace(scene, rays_dev, n, mode, hits_dev, (cudaStream_t)stream);
}

int b2pt_trace(b2pt_scene_t scene, const B2ptRay* rays, size_t n, unsigned mode, B2ptHit* hits) {
  B2_CHECK(scene && (n == 0 || (rays && hits)), B2PT_ERR_INVALID, "b2pt_trace: null argument");
  if (n == 0) return B2PT_OK;
  B2ptRay* d_rays = nullptr; B2ptHit* d_hits = nullptr;
  B2_CUDA(cudaMalloc(&d_rays, n * sizeof(B2ptRay)));
  cudaError_t e = cudaMalloc(&d_hits, n * sizeof(B2ptHit));
  if (e != cudaSuccess) { cudaFree(d_rays); return b2pt::cuda_fail(e, "cudaMalloc", __FILE__, __LINE__); }
  int rc = B2PT_OK;
  e = cudaMemcpy(d_rays, rays, n * sizeof(B2ptRay), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    rc = b2pt::launch_trace(scene, d_rays, n, mode, d_hits, 0);
    if (rc == B2PT_OK) e = cudaMemcpy(hits, d_hits, n * sizeof(B2ptHit), cudaMemcpyDeviceToHost);
  }
  cudaFree(d_rays); cudaFree(d_hits);
  if (e != cudaSuccess) return b2pt::cuda_fail(e, "b2pt_trace copy", __FILE__, __LINE__);
  return rc;
}

}  // extern "C"
