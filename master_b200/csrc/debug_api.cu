// debug_api.cu — parity probes (include/b2pt_debug.h): single device functions of the hot path exposed so that tests/ can set
// them beside the reference's own functions on the same inputs.  Not part of the drop-in boundary; nothing here is on the
// render path.  b2pt_debug_density lives in vcm.cu next to the density stages it shares its loop with.
#include <vector>

#include "../../include/b2pt_debug.h"
#include "render_common.cuh"

namespace b2pt {
namespace {

struct DevBuf {  // RAII device buffer for the probes (a few KB .. MB, host-synchronous calls)
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16); }
  template <class T> T* as() const { return (T*)p; }
};

__global__ void debug_primary_kernel(B2ptRenderParams P, const uint32_t* pixels, const uint32_t* samples, uint32_t n, float* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  RenderConfig cfg{};
  cfg.params = P;
  cfg.width_f = float(P.width);
  cfg.res_y_inv = 1.0f / float(P.height);
  Rng rng; rng.key = stream_key(P.seed, kDomainPixel, pixels[i], samples[i]); rng.ctr = 0;
  const float3 d = shoot_primary(cfg, rng, pixels[i]);
  out[i * 3 + 0] = d.x; out[i * 3 + 1] = d.y; out[i * 3 + 2] = d.z;
}

__global__ void debug_camera_kernel(B2ptRenderParams P, const float* pos, uint32_t n, float* dirs, float* pix) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float width = float(P.width), height = float(P.height), res_y_inv = 1.0f / float(P.height);
  const float3 d = camera_ray_direction(P.camera, pos[i * 2], pos[i * 2 + 1], width, res_y_inv);
  dirs[i * 3 + 0] = d.x; dirs[i * 3 + 1] = d.y; dirs[i * 3 + 2] = d.z;
  mat3 w2v;
  const float* m = P.camera.world_to_view;
  w2v.c0 = f3(m[0], m[1], m[2]); w2v.c1 = f3(m[3], m[4], m[5]); w2v.c2 = f3(m[6], m[7], m[8]);
  const float2 pp = pixel_position(mul(w2v, d), width, height, res_y_inv, P.camera.focal_length_y);  // Technique::_accumulate, Technique.cpp:283-287
  pix[i * 2] = pp.x; pix[i * 2 + 1] = pp.y;
}

__global__ void debug_light_kernel(DeviceScene sc, const uint32_t* keys, uint32_t n, float* pos, uint32_t* mat, uint32_t* draws) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Rng rng; rng.key = keys[i]; rng.ctr = 0;
  const LightSample ls = sample_light(sc, rng);
  pos[i * 3 + 0] = ls.surface.pos.x; pos[i * 3 + 1] = ls.surface.pos.y; pos[i * 3 + 2] = ls.surface.pos.z;
  mat[i] = ls.surface.material_id;
  draws[i] = rng.ctr;
}

__global__ void debug_draws_kernel(uint32_t key, uint32_t n, uint32_t* raw, float* floats, double* doubles) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Rng a; a.key = key; a.ctr = i;
  raw[i] = a.next_u32();
  Rng b; b.key = key; b.ctr = i;
  floats[i] = b.sample();
  Rng c; c.key = key; c.ctr = 2u * i;
  doubles[i] = c.sample_double();
}

// ---- tree statistics: the boxes a traversal actually tests are the QUANTISED child boxes, so the cost is taken from them
__device__ __forceinline__ void node8_child_box(const float4* nd, int slot, float3& lo, float3& hi) {
  const float4 n0 = nd[0], n2 = nd[2], n3 = nd[3], n4 = nd[4];
  const uint32_t hdr = __float_as_uint(n0.w);
  const float sx = __uint_as_float((hdr & 0xFFu) << 23), sy = __uint_as_float(((hdr >> 8) & 0xFFu) << 23), sz = __uint_as_float(((hdr >> 16) & 0xFFu) << 23);
  auto byte_of = [&](float a, float b) { const uint32_t w = __float_as_uint(slot < 4 ? a : b); return float((w >> (8 * (slot & 3))) & 0xFFu); };
  lo = f3(n0.x + byte_of(n2.x, n2.y) * sx, n0.y + byte_of(n2.z, n2.w) * sy, n0.z + byte_of(n3.x, n3.y) * sz);
  hi = f3(n0.x + byte_of(n3.z, n3.w) * sx, n0.y + byte_of(n4.x, n4.y) * sy, n0.z + byte_of(n4.z, n4.w) * sz);
}
__device__ __forceinline__ double box_area(float3 lo, float3 hi) {
  const double dx = double(hi.x) - lo.x, dy = double(hi.y) - lo.y, dz = double(hi.z) - lo.z;
  return dx < 0 || dy < 0 || dz < 0 ? 0.0 : 2.0 * (dx * dy + dy * dz + dz * dx);
}

// sums[0] = sum over inner CHILD boxes of area, [1] = inner nodes, [2] = leaf children, [3] = triangle references,
// [4] = sum over leaf children of area * triangles, [5] = area of the root's box (union of its children)
__global__ void tree_stats_kernel(DeviceScene sc, double* sums) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= sc.num_nodes) return;
  const float4* nd = sc.nodes8 + size_t(i) * kNode8Float4s;
  const uint32_t imask = __float_as_uint(nd[0].w) >> 24;
  const uint32_t meta[2] = {__float_as_uint(nd[1].z), __float_as_uint(nd[1].w)};
  double inner_area = 0.0, leaf_area = 0.0, leaves = 0.0, tris = 0.0;
  float3 rlo = f3(INFINITY), rhi = f3(-INFINITY);
  for (int s = 0; s < 8; ++s) {
    const uint32_t m = (meta[s >> 2] >> (8 * (s & 3))) & 0xFFu;
    if (m == 0u) continue;  // empty slot
    float3 lo, hi;
    node8_child_box(nd, s, lo, hi);
    const double a = box_area(lo, hi);
    if (i == 0) { rlo = f3(fminf(rlo.x, lo.x), fminf(rlo.y, lo.y), fminf(rlo.z, lo.z)); rhi = f3(fmaxf(rhi.x, hi.x), fmaxf(rhi.y, hi.y), fmaxf(rhi.z, hi.z)); }
    if (imask & (1u << s)) { inner_area += a; }
    else { const double t = double(__popc(m >> 5)); leaf_area += a * t; leaves += 1.0; tris += t; }
  }
  atomicAdd(&sums[0], inner_area); atomicAdd(&sums[1], 1.0); atomicAdd(&sums[2], leaves); atomicAdd(&sums[3], tris); atomicAdd(&sums[4], leaf_area);
  if (i == 0) sums[5] = box_area(rlo, rhi);
}

template <bool kAnyHit>
__global__ void __launch_bounds__(256) trace_counts_kernel(DeviceScene sc, const float4* __restrict__ rays, uint32_t n, uint32_t mask, uint2* counts) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 a = rays[size_t(i) * 2], b = rays[size_t(i) * 2 + 1];
    uint2 c = make_uint2(0u, 0u);
    trace_ray8<kAnyHit, true>(sc, f3(a.x, a.y, a.z), f3(b.x, b.y, b.z), a.w, b.w, mask, &c);
    counts[i] = c;
  }
}

__global__ void debug_pow_kernel(const float* x, const float* y, uint32_t n, float* table, float* lib) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  table[i] = cr_powf(x[i], y[i]);
  lib[i] = cr_powf_lib(x[i], y[i]);
}

int need_device(const char* who) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); set_error(std::string(who) + ": no CUDA device (there is no CPU fallback)"); return B2PT_ERR_NO_DEVICE; }
  return B2PT_OK;
}

}  // namespace
}  // namespace b2pt

extern "C" {

int b2pt_debug_pow(const float* x, const float* y, size_t n, float* table, float* lib) {
  using namespace b2pt;
  B2_CHECK(x && y && table && lib, B2PT_ERR_INVALID, "b2pt_debug_pow: null argument");
  if (int rc = need_device("b2pt_debug_pow")) return rc;
  if (n == 0) return B2PT_OK;
  DevBuf dx, dy, dt, dl;
  B2_CUDA(dx.alloc(n * 4)); B2_CUDA(dy.alloc(n * 4)); B2_CUDA(dt.alloc(n * 4)); B2_CUDA(dl.alloc(n * 4));
  B2_CUDA(cudaMemcpy(dx.p, x, n * 4, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(dy.p, y, n * 4, cudaMemcpyHostToDevice));
  debug_pow_kernel<<<unsigned((n + 255) / 256), 256>>>(dx.as<float>(), dy.as<float>(), uint32_t(n), dt.as<float>(), dl.as<float>());
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaMemcpy(table, dt.p, n * 4, cudaMemcpyDeviceToHost));
  B2_CUDA(cudaMemcpy(lib, dl.p, n * 4, cudaMemcpyDeviceToHost));
  return B2PT_OK;
}

int b2pt_debug_primary(const B2ptRenderParams* params, const uint32_t* pixels, const uint32_t* samples, size_t n, float* directions) {
  using namespace b2pt;
  B2_CHECK(params && (n == 0 || (pixels && samples && directions)), B2PT_ERR_INVALID, "b2pt_debug_primary: null argument");
  if (int rc = need_device("b2pt_debug_primary")) return rc;
  if (n == 0) return B2PT_OK;
  DevBuf px, sm, out;
  B2_CUDA(px.alloc(n * 4)); B2_CUDA(sm.alloc(n * 4)); B2_CUDA(out.alloc(n * 12));
  B2_CUDA(cudaMemcpy(px.p, pixels, n * 4, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(sm.p, samples, n * 4, cudaMemcpyHostToDevice));
  debug_primary_kernel<<<unsigned((n + 127) / 128), 128>>>(*params, px.as<uint32_t>(), sm.as<uint32_t>(), uint32_t(n), out.as<float>());
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaMemcpy(directions, out.p, n * 12, cudaMemcpyDeviceToHost));
  return B2PT_OK;
}

int b2pt_debug_camera(const B2ptRenderParams* params, const float* positions, size_t n, float* directions, float* pixels) {
  using namespace b2pt;
  B2_CHECK(params && (n == 0 || (positions && directions && pixels)), B2PT_ERR_INVALID, "b2pt_debug_camera: null argument");
  if (int rc = need_device("b2pt_debug_camera")) return rc;
  if (n == 0) return B2PT_OK;
  DevBuf pos, dir, pix;
  B2_CUDA(pos.alloc(n * 8)); B2_CUDA(dir.alloc(n * 12)); B2_CUDA(pix.alloc(n * 8));
  B2_CUDA(cudaMemcpy(pos.p, positions, n * 8, cudaMemcpyHostToDevice));
  debug_camera_kernel<<<unsigned((n + 127) / 128), 128>>>(*params, pos.as<float>(), uint32_t(n), dir.as<float>(), pix.as<float>());
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaMemcpy(directions, dir.p, n * 12, cudaMemcpyDeviceToHost));
  B2_CUDA(cudaMemcpy(pixels, pix.p, n * 8, cudaMemcpyDeviceToHost));
  return B2PT_OK;
}

int b2pt_debug_light_samples(b2pt_scene_t scene, const uint32_t* keys, size_t n, float* positions, uint32_t* material_ids, uint32_t* draws) {
  using namespace b2pt;
  B2_CHECK(scene && (n == 0 || (keys && positions && material_ids && draws)), B2PT_ERR_INVALID, "b2pt_debug_light_samples: null argument");
  B2_CHECK(scene->dev.num_lights > 0, B2PT_ERR_INVALID, "b2pt_debug_light_samples: the scene has no lights (AreaLights.cpp:217)");
  if (n == 0) return B2PT_OK;
  DevBuf k, pos, mat, dr;
  B2_CUDA(k.alloc(n * 4)); B2_CUDA(pos.alloc(n * 12)); B2_CUDA(mat.alloc(n * 4)); B2_CUDA(dr.alloc(n * 4));
  B2_CUDA(cudaMemcpy(k.p, keys, n * 4, cudaMemcpyHostToDevice));
  debug_light_kernel<<<unsigned((n + 127) / 128), 128>>>(scene->dev, k.as<uint32_t>(), uint32_t(n), pos.as<float>(), mat.as<uint32_t>(), dr.as<uint32_t>());
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaMemcpy(positions, pos.p, n * 12, cudaMemcpyDeviceToHost));
  B2_CUDA(cudaMemcpy(material_ids, mat.p, n * 4, cudaMemcpyDeviceToHost));
  B2_CUDA(cudaMemcpy(draws, dr.p, n * 4, cudaMemcpyDeviceToHost));
  return B2PT_OK;
}

int b2pt_debug_draws(uint32_t key, size_t n, uint32_t* raw, float* floats, double* doubles) {
  using namespace b2pt;
  B2_CHECK(n == 0 || (raw && floats && doubles), B2PT_ERR_INVALID, "b2pt_debug_draws: null argument");
  if (int rc = need_device("b2pt_debug_draws")) return rc;
  if (n == 0) return B2PT_OK;
  DevBuf r, f, d;
  B2_CUDA(r.alloc(n * 4)); B2_CUDA(f.alloc(n * 4)); B2_CUDA(d.alloc(n * 8));
  debug_draws_kernel<<<unsigned((n + 127) / 128), 128>>>(key, uint32_t(n), r.as<uint32_t>(), f.as<float>(), d.as<double>());
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaMemcpy(raw, r.p, n * 4, cudaMemcpyDeviceToHost));
  B2_CUDA(cudaMemcpy(floats, f.p, n * 4, cudaMemcpyDeviceToHost));
  B2_CUDA(cudaMemcpy(doubles, d.p, n * 8, cudaMemcpyDeviceToHost));
  return B2PT_OK;
}

int b2pt_debug_tree_stats(b2pt_scene_t scene, double* stats6) {
  using namespace b2pt;
  B2_CHECK(scene && stats6, B2PT_ERR_INVALID, "b2pt_debug_tree_stats: null argument");
  B2_CHECK(scene->dev.nodes8, B2PT_ERR_INVALID, "b2pt_debug_tree_stats: the scene has no 8-wide tree");
  DevBuf s;
  B2_CUDA(s.alloc(6 * 8));
  B2_CUDA(cudaMemset(s.p, 0, 6 * 8));
  tree_stats_kernel<<<(scene->dev.num_nodes + 127) / 128, 128>>>(scene->dev, s.as<double>());
  B2_CUDA(cudaGetLastError());
  double h[6];
  B2_CUDA(cudaMemcpy(h, s.p, sizeof h, cudaMemcpyDeviceToHost));
  // SAH cost with unit costs: every inner node visited tests its 8 child boxes once (cost 1 per node visit), every leaf child
  // entered tests its triangles (cost 1 each); probabilities = child box area / root area
  const double root = h[5] > 0 ? h[5] : 1.0;
  stats6[0] = (root + h[0] + h[4]) / root;  // expected node visits (root + inner children) + expected triangle tests
  stats6[1] = h[1]; stats6[2] = h[2]; stats6[3] = h[3];
  stats6[4] = h[4] / root;            // expected triangle tests of a random long ray
  stats6[5] = (root + h[0]) / root;   // expected node visits of a random long ray
  return B2PT_OK;
}

int b2pt_debug_l2_window(b2pt_scene_t scene, double* out6) {
  using namespace b2pt;
  B2_CHECK(scene && out6, B2PT_ERR_INVALID, "b2pt_debug_l2_window: null argument");
  int dev = 0; cudaDeviceProp prop{};
  B2_CUDA(cudaGetDevice(&dev));
  B2_CUDA(cudaGetDeviceProperties(&prop, dev));
  out6[0] = double(scene->bvh_block_bytes); out6[1] = double(scene->l2_window_bytes); out6[2] = scene->l2_hit_ratio;
  out6[3] = double(prop.persistingL2CacheMaxSize); out6[4] = double(prop.accessPolicyMaxWindowSize); out6[5] = double(prop.l2CacheSize);
  return B2PT_OK;
}

int b2pt_debug_trace_counts(b2pt_scene_t scene, const B2ptRay* rays, size_t n, unsigned mode, uint32_t* counts) {
  using namespace b2pt;
  B2_CHECK(scene && (n == 0 || (rays && counts)), B2PT_ERR_INVALID, "b2pt_debug_trace_counts: null argument");
  B2_CHECK(mode <= B2PT_TRACE_OCCLUDED_MESH, B2PT_ERR_INVALID, "b2pt_debug_trace_counts: unknown mode");
  B2_CHECK(scene->dev.nodes8, B2PT_ERR_INVALID, "b2pt_debug_trace_counts: the scene has no 8-wide tree");
  if (n == 0) return B2PT_OK;
  DevBuf r, c;
  B2_CUDA(r.alloc(n * sizeof(B2ptRay))); B2_CUDA(c.alloc(n * 8));
  B2_CUDA(cudaMemcpy(r.p, rays, n * sizeof(B2ptRay), cudaMemcpyHostToDevice));
  const int grid = grid_for_trace(n, 256);
  const uint32_t mask = mode == B2PT_TRACE_CLOSEST_ALL ? kMaskAll : kMaskMesh;
  if (mode == B2PT_TRACE_OCCLUDED_MESH) trace_counts_kernel<true><<<grid, 256>>>(scene->dev, r.as<float4>(), uint32_t(n), mask, c.as<uint2>());
  else trace_counts_kernel<false><<<grid, 256>>>(scene->dev, r.as<float4>(), uint32_t(n), mask, c.as<uint2>());
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaMemcpy(counts, c.p, n * 8, cudaMemcpyDeviceToHost));
  return B2PT_OK;
}

}  // extern "C"
