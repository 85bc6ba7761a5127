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
  double* a = cfg.accum + size_t(pixel) * 4;
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

inline int grid_for(size_t n, int block) {
  size_t want = (n + block - 1) / block;
  size_t cap = size_t(kNumSMs) * 8;
  if (want < 1) want = 1;
  return int(want < cap ? want : cap);
}

}  // namespace b2pt
