// render.cu — b2pt_render / b2pt_render_host: the Technique::render boundary (Technique.cpp:15-77).
//
// Builds the render_context_t equivalent (Technique.cpp:37-45), dispatches to the wavefront
// integrator the Options select (make_technique.cpp:112-170) and fills the statistics_t fields a
// device technique can know (Technique.cpp:52-76).
#include <vector>

#include "render_common.cuh"

namespace b2pt {

int render_pt(b2pt_scene_s* scene, const RenderConfig& cfg, cudaStream_t stream, B2ptStats* stats);
int render_bpt(b2pt_scene_s* scene, const RenderConfig& cfg, cudaStream_t stream, B2ptStats* stats);
int render_vcm(b2pt_scene_s* scene, const RenderConfig& cfg, cudaStream_t stream, B2ptStats* stats);

// rms_abs_errors / rms_abs_errors_windowed (ImageView.cpp:60-117): per-pixel terms as the reference computes them, summed in double.
// sums[0] += dot(d, d), sums[1] += d.x + d.y + d.z, sums[2] += 1 per pixel of the window
__global__ void __launch_bounds__(256) film_errors_kernel(const double* __restrict__ accum, const float* __restrict__ ref, uint32_t width,
                                                          int x0, int y0, int x1, int y1, double* sums) {
  const uint32_t w = uint32_t(x1 - x0), n = w * uint32_t(y1 - y0);
  double sq = 0.0, ab = 0.0, cnt = 0.0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const size_t p = size_t(y0 + int(i / w)) * width + size_t(x0 + int(i % w));
    const double2 rg = *reinterpret_cast<const double2*>(accum + p * 4), bw = *reinterpret_cast<const double2*>(accum + p * 4 + 2);
    const float dx = fabsf(float(rg.x / bw.y) - ref[p * 3 + 0]);
    const float dy = fabsf(float(rg.y / bw.y) - ref[p * 3 + 1]);
    const float dz = fabsf(float(bw.x / bw.y) - ref[p * 3 + 2]);
    ab += double(__fadd_rn(__fadd_rn(dx, dy), dz));
    sq += double(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));  // glm::dot: tmp = a * b; tmp.x + tmp.y + tmp.z
    cnt += 1.0;
  }
  for (int o = 16; o > 0; o >>= 1) {
    sq += __shfl_down_sync(0xFFFFFFFFu, sq, o); ab += __shfl_down_sync(0xFFFFFFFFu, ab, o); cnt += __shfl_down_sync(0xFFFFFFFFu, cnt, o);
  }
  __shared__ double part[3][8];
  const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  if (lane == 0) { part[0][warp] = sq; part[1][warp] = ab; part[2][warp] = cnt; }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += part[threadIdx.x][k];
    atomicAdd(sums + threadIdx.x, t);
  }
}

static int make_config(b2pt_scene_s* scene, const B2ptRenderParams* p, double* accum, float* path_radiance, RenderConfig* cfg) {
  B2_CHECK(p->width > 0 && p->height > 0, B2PT_ERR_INVALID, "b2pt_render: empty window");
  B2_CHECK(p->technique <= B2PT_TECHNIQUE_UPG, B2PT_ERR_INVALID, "b2pt_render: unknown technique");
  B2_CHECK(p->roulette > 0.0f && p->roulette <= 1.0f, B2PT_ERR_INVALID, "b2pt_render: roulette must be in (0, 1]");
  // AreaLights::_sampleLight: runtime_assert(num_lights() != 0) (AreaLights.cpp:217)
  B2_CHECK(scene->dev.num_lights > 0, B2PT_ERR_INVALID, "b2pt_render: the scene has no lights (AreaLights.cpp:217)");
  cfg->scene = scene->dev;
  cfg->params = *p;
  cfg->camera = Surface();
  // host copy of camera_surface(): same arithmetic as the device function (negations only)
  const float* v = p->camera.view_to_world;
  cfg->camera.pos = f3(p->camera.position[0], p->camera.position[1], p->camera.position[2]);
  cfg->camera.tan.c0 = f3(v[3], v[4], v[5]);
  cfg->camera.tan.c1 = f3(-v[6], -v[7], -v[8]);
  cfg->camera.tan.c2 = f3(v[0], v[1], v[2]);
  cfg->camera.gnormal = f3(-v[6], -v[7], -v[8]);
  cfg->camera.material_id = 0u << 2 | B2PT_ENTITY_CAMERA;
  cfg->width_f = float(p->width);
  cfg->height_f = float(p->height);
  cfg->res_y_inv = 1.0f / float(p->height);
  cfg->focal_factor_y = p->camera.focal_length_y * p->camera.focal_length_y * 0.25f;
  cfg->pixels = p->width * p->height;
  cfg->total_work = (unsigned long long)cfg->pixels * p->num_samples;
  cfg->pool_size = 0;
  cfg->accum = accum;
  cfg->accum_pass_stride = 0; cfg->pass_index0 = 0; cfg->pass_index_step = 1;
  cfg->path_radiance = path_radiance;
  cfg->counters = nullptr;
  return B2PT_OK;
}

}  // namespace b2pt

namespace b2pt {

// b2pt_render with the film layout made explicit: pass_stride = 0 adds every pass into the one film at accum_dev (b2pt_render);
// pass_stride = width * height * 4 gives every pass of the call its own film (the host pipeline renders ahead of its caller).
int render_passes(b2pt_scene_s* scene, const B2ptRenderParams* params, double* accum_dev, size_t pass_stride, float* path_radiance_dev,
                  cudaStream_t stream, B2ptStats* stats) {
  RenderConfig cfg;
  if (int rc = make_config(scene, params, accum_dev, path_radiance_dev, &cfg)) return rc;
  cfg.accum_pass_stride = pass_stride;
  B2ptStats local{};
  if (params->num_samples == 0) { if (stats) *stats = local; return B2PT_OK; }
  scene_l2_window(scene, stream);  // BVH block -> persisting L2 for the kernels of this stream (scene.cu)
  // a stack overflow of an earlier call must not fail this one: the flag is per call
  B2_CUDA(cudaMemsetAsync(scene->dev.error_flags, 0, 4, stream));
  if (!scene->ev[0]) { B2_CUDA(cudaEventCreate(&scene->ev[0])); }
  if (!scene->ev[1]) { B2_CUDA(cudaEventCreate(&scene->ev[1])); }
  cudaEvent_t e0 = scene->ev[0], e1 = scene->ev[1];  // kept with the scene: one pair per handle, not per call
  B2_CUDA(cudaEventRecord(e0, stream));
  int rc;
  switch (params->technique) {
    case B2PT_TECHNIQUE_PT: rc = render_pt(scene, cfg, stream, &local); break;
    case B2PT_TECHNIQUE_BPT: rc = render_bpt(scene, cfg, stream, &local); break;
    default: rc = render_vcm(scene, cfg, stream, &local); break;
  }
  if (rc == B2PT_OK) {
    cudaError_t e = cudaEventRecord(e1, stream);
    if (e == cudaSuccess) e = cudaEventSynchronize(e1);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&local.device_ms, e0, e1);
    if (e != cudaSuccess) rc = cuda_fail(e, "b2pt_render timing", __FILE__, __LINE__);
  }
  if (rc != B2PT_OK) return rc;
  uint32_t flags = 0;
  B2_CUDA(cudaMemcpy(&flags, scene->dev.error_flags, 4, cudaMemcpyDeviceToHost));
  B2_CHECK((flags & 1u) == 0, B2PT_ERR_CAPACITY, "BVH traversal stack overflow");
  local.num_samples = params->num_samples;
  local.num_paths = cfg.total_work;
  if (stats) *stats = local;
  return B2PT_OK;
}

}  // namespace b2pt

extern "C" {

int b2pt_render(b2pt_scene_t scene, const B2ptRenderParams* params, double* accum_dev, float* path_radiance_dev, void* stream_,
                B2ptStats* stats) {
  using namespace b2pt;
  B2_CHECK(scene && params && accum_dev, B2PT_ERR_INVALID, "b2pt_render: null argument");
  // one render at a time per scene handle: passes the host pipeline is still rendering ahead for b2pt_render_host finish first
  if (scene->pipeline_quiesce) scene->pipeline_quiesce(scene);
  return render_passes(scene, params, accum_dev, 0, path_radiance_dev, (cudaStream_t)stream_, stats);
}

int b2pt_film_errors(const double* accum_dev, const float* reference_dev, uint32_t width, uint32_t height, const int32_t* window, float* out,
                     void* stream_) {
  using namespace b2pt;
  B2_CHECK(accum_dev && reference_dev && out, B2PT_ERR_INVALID, "b2pt_film_errors: null argument");
  B2_CHECK(width > 0 && height > 0, B2PT_ERR_INVALID, "b2pt_film_errors: empty image");
  int dev_count = 0;
  if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0) { cudaGetLastError(); B2_CHECK(false, B2PT_ERR_NO_DEVICE, "b2pt_film_errors: no CUDA device"); }
  cudaStream_t stream = (cudaStream_t)stream_;
  int x0 = 0, y0 = 0, x1 = int(width), y1 = int(height);
  if (window) {  // ImageView.cpp:103-104
    x0 = max(0, window[0] - window[2] + 1); x1 = min(window[0] + window[2], int(width));
    y0 = max(0, window[1] - window[2] + 1); y1 = min(window[1] + window[2], int(height));
  }
  double sums[3] = {0.0, 0.0, 0.0};
  if (x1 > x0 && y1 > y0) {
    double* sums_dev = nullptr;
    B2_CUDA(cudaMallocAsync(&sums_dev, sizeof(sums), stream));  // stream-ordered: served from the driver's pool, no device-wide sync per record
    cudaMemsetAsync(sums_dev, 0, sizeof(sums), stream);
    const size_t n = size_t(x1 - x0) * size_t(y1 - y0);
    const int block = 256, grid = int(std::min<size_t>((n + block - 1) / block, size_t(kNumSMs) * 4));
    film_errors_kernel<<<grid, block, 0, stream>>>(accum_dev, reference_dev, width, x0, y0, x1, y1, sums_dev);
    cudaError_t e = cudaMemcpyAsync(sums, sums_dev, sizeof(sums), cudaMemcpyDeviceToHost, stream);
    cudaFreeAsync(sums_dev, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    B2_CUDA(e);
  }
  // whole image: num = width * height * 3 (ImageView.cpp:81-84); window: num = pixels (ImageView.cpp:109-116); an empty window is 0 / 0
  const float num = window ? float(sums[2]) : float(size_t(width) * height * 3);
  out[0] = sqrtf(float(sums[0]) / num);
  out[1] = float(sums[1]) / num;
  return B2PT_OK;
}

}  // extern "C"
