// GpuTechnique.cpp — see GpuTechnique.hpp.  Build with the reference's include path and -fno-access-control
// (b2s_io.hpp reads the BSDF classes' private parameters); link against master_b200/libb2pt.so.
#include "GpuTechnique.hpp"

#include <stdexcept>

#include <ImageView.hpp>
#include <utility.hpp>

#include "b2s_io.hpp"

namespace b2pt_host {

using namespace haste;

static void check(int rc, const char* what) {
  // the reference reports failures as C++ exceptions (runtime_assert.cpp:7-11, Scene.cpp:55-57)
  if (rc != B2PT_OK) throw std::runtime_error(std::string(what) + ": " + b2pt_last_error());
}

#ifdef B2PT_HAVE_OPTIONS_HPP
// Options -> GpuOptions, the argument sets of make_bpt_technique / make_upg_technique / PT (make_technique.cpp:86-139)
GpuOptions GpuOptions::from(const haste::Options& o) {
  GpuOptions g;
  g.technique = o.technique == Options::PT ? B2PT_TECHNIQUE_PT : o.technique == Options::BPT ? B2PT_TECHNIQUE_BPT
              : o.technique == Options::VCM ? B2PT_TECHNIQUE_VCM : B2PT_TECHNIQUE_UPG;
  g.beta = float(o.beta); g.roulette = float(o.roulette); g.lights = o.lights;
  g.max_path = o.max_path == size_t(PTRDIFF_MAX) ? 0 : o.max_path;
  g.radius = float(o.radius); g.alpha = float(o.alpha); g.num_photons = o.num_photons;
  g.enable_vc = o.enable_vc; g.enable_vm = o.enable_vm; g.from_light = o.from_light;
  g.fixed_seed = o.enable_seed; g.seed = uint32_t(o.seed);
  return g;
}
#endif

GpuTechnique::GpuTechnique(const shared<const Scene>& scene, const GpuOptions& options)
    : Technique(scene, 1), _options(options) {
  FlatSceneArrays flat(*scene);
  B2ptSceneDesc desc = flat.desc();
  check(b2pt_scene_create(&desc, &_device_scene), "b2pt_scene_create");  // replaces Scene::buildAccelStructs
}

GpuTechnique::~GpuTechnique() {
  b2pt_film_destroy(_film);
  b2pt_scene_destroy(_device_scene);
}

void GpuTechnique::fill_params(B2ptRenderParams& p, size_t width, size_t height, size_t cameraId, uint32_t passes) {
  auto& cameras = _scene->cameras();
  std::memset(&p, 0, sizeof p);
  p.technique = _options.technique;
  p.beta = _options.beta; p.roulette = _options.roulette; p.lights = _options.lights;
  p.max_path = _options.max_path > 0xFFFFFFFFull ? 0u : uint32_t(_options.max_path);
  p.radius = _options.radius; p.alpha = _options.alpha; p.num_photons = uint32_t(_options.num_photons);
  p.unbiased = _options.technique == B2PT_TECHNIQUE_UPG;
  p.enable_vc = _options.enable_vc; p.enable_vm = _options.enable_vm; p.from_light = _options.from_light;
  p.width = uint32_t(width); p.height = uint32_t(height);
  // render_context_t, Technique.cpp:37-45
  const mat3 v2w = cameras.view_to_world_mat3(cameraId), w2v = cameras.world_to_view_mat3(cameraId);
  const vec3 pos = cameras.position(cameraId);
  std::memcpy(p.camera.view_to_world, &v2w, 36);
  std::memcpy(p.camera.world_to_view, &w2v, 36);
  std::memcpy(p.camera.position, &pos, 12);
  p.camera.focal_length_y = cameras.focal_length_y(cameraId, float(width) / float(height));
  p.seed = _seed;
  p.first_sample = _options.first_sample + uint32_t(_statistics.num_samples) * _options.sample_stride;
  p.sample_stride = _options.sample_stride;
  p.num_samples = passes;
  std::memcpy(p.sky_horizon, &_sky_horizon, 12);
  std::memcpy(p.sky_zenith, &_sky_zenith, 12);
}

void GpuTechnique::account(double start_time, uint32_t passes) {
  // statistics, Technique.cpp:52-76
  const double current_time = high_resolution_time();
  _statistics.num_samples += passes;
  _statistics.num_basic_rays += _last.num_basic_rays;
  _statistics.num_shadow_rays += _last.num_shadow_rays;
  _statistics.total_time = current_time - _start_time;
  statistics_t::record_t record;
  record.sample_index = _statistics.num_samples - 1;
  record.rms_error = 0.0f;
  record.abs_error = 0.0f;
  record.clock_time = float(_statistics.total_time);
  record.frame_duration = float(current_time - start_time);
  record.numeric_errors = _last.numeric_errors;
  _statistics.records.push_back(record);
}

void GpuTechnique::render_resident(size_t width, size_t height, size_t cameraId, uint32_t passes) {
  if (!_film || _film_width != width || _film_height != height) {
    b2pt_film_destroy(_film); _film = nullptr;
    check(b2pt_film_create(uint32_t(width), uint32_t(height), &_film), "b2pt_film_create");
    _film_width = width; _film_height = height;
  }
  if (!std::isfinite(_start_time)) _start_time = high_resolution_time();
  if (!_seeded) { _seed = _options.seed; _seeded = true; }  // ranks of a multi-GPU run share the seed: fixed, never drawn per rank
  const double start_time = high_resolution_time();
  B2ptRenderParams p;
  fill_params(p, width, height, cameraId, passes);
  check(b2pt_render(_device_scene, &p, b2pt_film_data(_film), nullptr, nullptr, &_last), "b2pt_render");
  account(start_time, passes);
}

void GpuTechnique::reduce_film(b2pt_comm_t comm, int root) {
  runtime_assert(_film != nullptr);
  check(b2pt_reduce_accum(comm, b2pt_film_data(_film), _film_width * _film_height * 4, root, nullptr), "b2pt_reduce_accum");
}

void GpuTechnique::read_film(subimage_view_t& view) {
  runtime_assert(_film != nullptr && view.width() == _film_width && view.height() == _film_height);
  check(b2pt_film_read(_film, (double*)view.data(), 1, nullptr), "b2pt_film_read");
}

void GpuTechnique::render(subimage_view_t& view, RandomEngine& engine, size_t cameraId, const vector<vec3>& reference,
                          const vector<ivec3>& trace_points) {
  runtime_assert(view.xWindow() == view.width() && view.yWindow() == view.height(),
                 "GpuTechnique renders the whole image (batch mode window)");
  auto& cameras = _scene->cameras();
  if (!std::isfinite(_start_time)) {  // Technique.cpp:24-31
    double time_offset = _statistics.records.empty() ? 0.0f : _statistics.records.back().clock_time;
    _start_time = high_resolution_time() - time_offset;
  }
  if (!_seeded) {
    _seed = _options.fixed_seed ? _options.seed : uint32_t(engine());
    _seeded = true;
  }
  const double start_time = high_resolution_time();

  B2ptRenderParams p;
  fill_params(p, view.width(), view.height(), cameraId, _options.passes_per_render);

  check(b2pt_render_host(_device_scene, &p, (double*)view.data(), &_last), "b2pt_render_host");

  // statistics, Technique.cpp:52-76
  const double current_time = high_resolution_time();
  _statistics.num_samples += _options.passes_per_render;
  _statistics.num_basic_rays += _last.num_basic_rays;
  _statistics.num_shadow_rays += _last.num_shadow_rays;
  _statistics.total_time = current_time - _start_time;
  statistics_t::record_t record;
  record.sample_index = _statistics.num_samples - 1;
  record.rms_error = 0.0f;
  record.abs_error = 0.0f;
  record.clock_time = float(_statistics.total_time);
  record.frame_duration = float(current_time - start_time);
  record.numeric_errors = _last.numeric_errors;
  if (!reference.empty()) {
    auto a = image_view_t<dvec4>(view);
    auto b = image_view_t<vec3>(reference, view.width(), view.height());
    rms_abs_errors(record.rms_error, record.abs_error, a, b);
    _make_measurements(trace_points, a, b);
  }
  _statistics.records.push_back(record);
}

}  // namespace b2pt_host
