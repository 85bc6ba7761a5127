// GpuTechnique.hpp — the drop-in: a haste::Technique whose render() runs on a B200 through the C ABI.
//
// Compiled against the reference's own headers (Technique.hpp:29-114).  It follows the precedent of
// haste::Viewer (Viewer.hpp:13-18), which overrides Technique::render wholesale; _traceEye / _preprocess
// are never called.  Construction mirrors makeTechnique's argument sets (make_technique.cpp:86-170):
//
//   case Options::BPT:  // with --gpu
//     result = std::make_shared<b2pt_host::GpuTechnique>(scene, b2pt_host::GpuOptions::from(options));
//
// One pass of one sample per pixel is added to `view` per render() call, exactly like Technique::render
// (Technique.cpp:15-77); the random streams are per (pixel, sample) with a base seed drawn from the
// caller's RandomEngine on the first call, so a seeded run is reproducible on any number of GPUs.
#pragma once
#include <Technique.hpp>
#ifdef B2PT_HAVE_OPTIONS_HPP
#include <Options.hpp>
#endif

#include "../../include/b2pt.h"

namespace b2pt_host {

struct GpuOptions {
  uint32_t technique = B2PT_TECHNIQUE_PT;  // Options::technique (Options.hpp:20)
  float beta = 1.0f, roulette = 0.9f, lights = 1.0f;
  size_t max_path = 0;                     // PT only; 0 = unlimited (Options::max_path = PTRDIFF_MAX)
  float radius = 0.01f, alpha = 0.75f;
  size_t num_photons = 0;
  bool enable_vc = true, enable_vm = true, from_light = true;
  uint32_t passes_per_render = 1;          // > 1 amortises the device->host copy of the film
  uint32_t first_sample = 0, sample_stride = 1;  // rank g of G: first = g, stride = G (multi-GPU sample split)
  bool fixed_seed = false; uint32_t seed = 1234;

#ifdef B2PT_HAVE_OPTIONS_HPP
  static GpuOptions from(const haste::Options& o);
#endif
};

class GpuTechnique : public haste::Technique {
 public:
  GpuTechnique(const haste::shared<const haste::Scene>& scene, const GpuOptions& options);
  ~GpuTechnique() override;

  void render(haste::subimage_view_t& view, haste::RandomEngine& engine, size_t cameraId,
              const haste::vector<haste::vec3>& reference, const haste::vector<haste::ivec3>& trace_points) override;

  const B2ptStats& last_stats() const { return _last; }

  // ---- the film kept on the device ("copies ... only when the host needs it"), and the multi-GPU reduce.
  // render_resident adds `passes` passes of the next sample indices into a dvec4 film that stays in HBM (b2pt_film_*);
  // reduce_film sums the films of all ranks of `comm` (b2pt_reduce_accum = ncclAllReduce / ncclReduce: merge_exr's a + b,
  // Options.cpp:1340-1409); read_film adds the device film into a host view.  Statistics are kept like render() keeps them.
  void render_resident(size_t width, size_t height, size_t cameraId, uint32_t passes);
  void reduce_film(b2pt_comm_t comm, int root);
  void read_film(haste::subimage_view_t& view);

 private:
  void fill_params(B2ptRenderParams& p, size_t width, size_t height, size_t cameraId, uint32_t passes);
  void account(double start_time, uint32_t passes);
  b2pt_film_t _film = nullptr;
  size_t _film_width = 0, _film_height = 0;
  GpuOptions _options;
  b2pt_scene_t _device_scene = nullptr;
  B2ptStats _last{};
  bool _seeded = false;
  uint32_t _seed = 0;
};

}  // namespace b2pt_host
