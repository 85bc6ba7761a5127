// render_common.cuh — pieces shared by the wavefront integrators (pt.cu, bpt.cu, vcm.cu).
#pragma once
#include <stdlib.h>
#include <vector>

#include "shading.cuh"
#include "trace.cuh"

namespace b2pt {

// Device-side counters of one render call.
struct RenderCounters {
  unsigned long long next_work;       // next (sample, pixel) index to hand out
  unsigned long long basic_rays;      // closest-hit rays traced (Scene::numNormalRays)
  unsigned long long shadow_rays;     // occlusion rays traced (Scene::numShadowRays)
  unsigned long long numeric_errors;  // non-finite path results dropped (Technique.cpp:224-230)
  unsigned long long capacity_events;
  unsigned int live_paths;            // paths in flight
  unsigned int queue_count[2][4];     // [iteration parity][queue id]
  unsigned int shadow_ticket[2];      // ray ticket of the persistent any-hit stage, [iteration parity]
};

// Everything a stage kernel needs to know about the render call. Passed by value.
struct RenderConfig {
  DeviceScene scene;
  B2ptRenderParams params;
  Surface camera;           // Technique::_camera_surface
  float width_f, height_f, res_y_inv, focal_factor_y;
  unsigned long long total_work;  // width * height * num_samples
  uint32_t pixels;          // width * height
  uint32_t pool_size;
  double* accum;            // dvec4 per pixel
  // Per-pass films (the host pipeline of render.cu renders ahead of the caller): pass p of the call (0-based, in the call's own
  // numbering) commits into accum + p * accum_pass_stride doubles; 0 = every pass into the one film.  A lane that renders every
  // lanes-th pass sees its local pass l as call pass pass_index0 + l * pass_index_step.
  size_t accum_pass_stride;
  uint32_t pass_index0, pass_index_step;
  float* path_radiance;     // nullable
  RenderCounters* counters;
};

// Technique::_commit_images (Technique.cpp:194-244) for one pixel-sample: view[p] += (rgb, 1) unless non-finite.
// The reference adds one dvec3 per pixel per pass into a dvec4 accumulator; here samples of the same
// pixel from different passes may be in flight together, so the add is a double-precision atomic
// (red.global.add.f64): order-independent to ~1e-16 relative.
__device__ __forceinline__ void commit_sample(const RenderConfig& cfg, unsigned long long work, float3 radiance) {
  const uint32_t pixel = uint32_t(work % cfg.pixels);
  if (cfg.path_radiance) {
    float* o = cfg.path_radiance + work * 3ull;
    o[0] = radiance.x; o[1] = radiance.y; o[2] = radiance.z;
  }
  double* a = cfg.accum + size_t(pixel) * 4 + size_t(cfg.pass_index0 + uint32_t(work / cfg.pixels) * cfg.pass_index_step) * cfg.accum_pass_stride;
  if (isfinite(l1norm(radiance))) {
    atomicAdd(a + 0, double(radiance.x));
    atomicAdd(a + 1, double(radiance.y));
    atomicAdd(a + 2, double(radiance.z));
    atomicAdd(a + 3, 1.0);
  } else {
    atomicAdd(&cfg.counters->numeric_errors, 1ull);
  }
}

// Technique::_for_each_ray's `shoot` (Technique.cpp:321-331): two jitter draws, then ray_direction.
// `vec2(x + sample(), y + sample())` is evaluated right to left by the reference build, so the first
// draw jitters y (pinned by tests/test_oracle_pinning.py).
__device__ __forceinline__ float3 shoot_primary(const RenderConfig& cfg, Rng& rng, uint32_t pixel) {
  const uint32_t px = pixel % cfg.params.width, py = pixel / cfg.params.width;
  const float jy = rng.sample();
  const float jx = rng.sample();
  return camera_ray_direction(cfg.params.camera, float(px) + jx, float(py) + jy, cfg.width_f, cfg.res_y_inv);
}

// ---------------------------------------------------------------- connections to directional (sun) lights
// BPTBase::_connect_directional (BPT.cpp:245-269) / UPGBase::_connect_directional (UPG.cpp:480-525): instead of an occlusion
// test towards a sampled point, a CLOSEST-hit ray is shot from the eye vertex against the light's direction; the light is
// seen iff the first thing hit is the light's own quad (material ids match), and the MIS weight depends on where it was hit.
// The logic stage queues what it knows (4 float4 per entry); this stage traces and finishes the arithmetic.
//   e0 = ray origin (already offset), bits(target path slot)     e1 = eye position, |dot(light normal, eye normal)|
//   e2 = unweighted contribution, KC                             e3 = bits(light id), KD, 0, 0
//   Cp = KC * beta(cos / d^2),  Dp = KD * beta(cos / d^2),  weightInv = (Cp + 1) * vc + beta_n * Dp * vm
// (BPT: vc = 1, vm = 0.)
struct DirectionalParams { float beta, vc, vm, beta_n; };

static __global__ void __launch_bounds__(256) directional_kernel(DeviceScene sc, const float4* __restrict__ queue, const unsigned* __restrict__ count,
                                                          uint32_t capacity, DirectionalParams P, float4* rad, uint32_t rad_stride, unsigned long long* basic_rays) {
  const uint32_t n = min(*count, capacity);
  if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(basic_rays, (unsigned long long)n);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 e0 = queue[size_t(i) * 4 + 0], e1 = queue[size_t(i) * 4 + 1], e2 = queue[size_t(i) * 4 + 2], e3 = queue[size_t(i) * 4 + 3];
    const B2ptLight& l = sc.lights[__float_as_uint(e3.x)];
    const float3 dir = -f3(l.tangent[3], l.tangent[4], l.tangent[5]);
    const float3 org = f3(e0.x, e0.y, e0.z);
    const Hit h = trace_ray<false, false>(sc, org, dir, 0.0f, INFINITY);  // Scene::intersect
    if (h.tri == kInvalid) continue;
    const uint32_t material = __float_as_uint(__ldg(&sc.tri_frames[size_t(h.tri) * kFrameFloat4s + 6]).w);
    if (material != l.material_id) continue;
    const float3 isect = org + dir * h.t;
    const float coeff = beta_fn(e1.w / distance2(isect, f3(e1.x, e1.y, e1.z)), P.beta);
    const float Cp = e2.w * coeff, Dp = e3.y * coeff;
    const float weight_inv = (Cp + 1.0f) * P.vc + P.beta_n * Dp * P.vm;
    const float3 result = f3(e2.x, e2.y, e2.z);
    if (l1norm(result) < kFltEpsilon) continue;
    const float3 c = result / weight_inv;
    float* r = &rad[size_t(__float_as_uint(e0.w)) * rad_stride].x;  // rad_stride: float4s between the radiance words of two paths
    atomicAdd(r + 0, c.x); atomicAdd(r + 1, c.y); atomicAdd(r + 2, c.z);
  }
}

// Per-stage device time (B2ptStats::trace_ms / shadow_ms / shade_ms): a CUDA-event pair around every
// launch of a stage, on the stream the kernels are launched on.  Counts launches always; records
// events only when B2PT_PROFILE_STAGES=1, because the pairs serialise nothing but do cost host time.
enum { kStageTrace = 0, kStageShadow = 1, kStageShade = 2, kNumStages = 3 };
struct StageTimer {
  bool enabled = false;
  cudaStream_t stream = nullptr;
  float ms[kNumStages] = {0.f, 0.f, 0.f};
  uint64_t launches[kNumStages] = {0, 0, 0};
  struct Pending { cudaEvent_t a, b; int stage; };
  std::vector<Pending> pending;
  std::vector<cudaEvent_t> spare;
  int open_stage = -1;
  cudaEvent_t open_a = nullptr;

  explicit StageTimer(cudaStream_t s) : stream(s) {
    const char* e = getenv("B2PT_PROFILE_STAGES");
    enabled = e && e[0] == '1';
  }
  ~StageTimer() {
    for (auto& p : pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto e : spare) cudaEventDestroy(e);
  }
  cudaEvent_t take() {
    if (!spare.empty()) { cudaEvent_t e = spare.back(); spare.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
  }
  void begin(int stage) {
    ++launches[stage];
    if (!enabled) return;
    open_stage = stage; open_a = take();
    cudaEventRecord(open_a, stream);
  }
  void end() {
    if (!enabled) return;
    cudaEvent_t b = take();
    cudaEventRecord(b, stream);
    pending.push_back(Pending{open_a, b, open_stage});
  }
  // call after the stream has been synchronised
  void resolve() {
    for (auto& p : pending) {
      float t = 0.f;
      if (cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) ms[p.stage] += t;
      spare.push_back(p.a); spare.push_back(p.b);
    }
    pending.clear();
  }
  uint64_t total_launches() const { return launches[0] + launches[1] + launches[2]; }
  void fill(B2ptStats* st) const {
    st->trace_ms = ms[kStageTrace]; st->shadow_ms = ms[kStageShadow]; st->shade_ms = ms[kStageShade];
    st->trace_launches = launches[kStageTrace]; st->shadow_launches = launches[kStageShadow];
    st->kernel_launches = total_launches();
  }
};

// Stored light-subpath vertices provisioned per path in the vertex pools of BPT and VCM / UPG.  The reference keeps up to 1024 per
// path on the stack (BPT.hpp:29, UPG.hpp:54); a pool in HBM is sized from what paths actually store: a subpath survives each bounce
// with probability `roulette`, so a closed scene stores 1 / (1 - roulette) vertices per path on average (10 at the default 0.9), and
// the sum over >= 10^5 paths concentrates tightly around the mean.  1.2 x the mean, at least 12; B2PT_LV_PER_PATH overrides.
// Exhausting the pool is an error (B2PT_ERR_CAPACITY), never a silently truncated subpath.
inline uint32_t lv_per_path(float roulette) {
  if (const char* e = getenv("B2PT_LV_PER_PATH")) { const int v = atoi(e); if (v >= 1 && v <= 1024) return uint32_t(v); }
  const double mean = roulette < 1.0f ? 1.0 / (1.0 - double(roulette)) : 1024.0;
  const double want = 1.2 * mean + 0.999;
  return want < 12.0 ? 12u : want > 1024.0 ? 1024u : uint32_t(want);
}
inline int capacity_error(unsigned long long events, const char* who) {
  char buf[320];
  snprintf(buf, sizeof buf, "%s: %llu light vertices / connections did not fit the device pools (sized for 1.2 / (1 - roulette) stored vertices "
           "per path); the film of this call is incomplete — set B2PT_LV_PER_PATH higher", who, events);
  set_error(buf);
  return B2PT_ERR_CAPACITY;
}

inline int grid_for(size_t n, int block) {
  size_t want = (n + block - 1) / block;
  size_t cap = size_t(kNumSMs) * 8;
  if (want < 1) want = 1;
  return int(want < cap ? want : cap);
}

}  // namespace b2pt
