// host_main.cpp — headless `master` with the GPU technique plugged in.
//
//   b2pt_master <scene.b2s> --BPT --beta=2 --resolution=1024x1024 --num-samples=64 --output=image.exr [--seed= --camera= ...]
//   b2pt_master continue <image.exr> [<scene.b2s>] --num-samples=128        master continue (main.cpp:108-…)
//   b2pt_master merge <out.exr> <fst.exr> <snd.exr>                         master merge    (Options.cpp:1340-1409)
//   b2pt_master import <scene.b2i> <scene.b2s>                              the loader's own math without Assimp (raw_import.hpp)
//   b2pt_master <scene.b2s> --BPT ... --gpus=N                              N GPUs of this node: one host thread per device, the
//                                                                           samples split g, g + N, ..., films summed with ONE
//                                                                           b2pt_reduce_accum (NCCL) — merge's a + b on the device
//
// Rendering is Framework::runBatch (framework.cpp:426-437) around GpuTechnique::render, the way Application::render drives
// any Technique (Application.cpp:41-79): one pass (or passes_per_render of them) per call until num_samples / num_seconds.
// Scenes come from a .b2s (no Assimp on this path).  The command line is the reference's (cli.hpp); `key=value` arguments
// are accepted too.  Output: the reference's EXR checkpoint (exr_io.hpp) and/or the dvec4 accumulator as raw float64 (out=).
// Without --seed the seed is 1234 (the reference would ask std::random_device; a headless driver stays reproducible).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <map>
#include <thread>

#include <xmmintrin.h>
#include <pmmintrin.h>

#include "GpuTechnique.hpp"
#include "b2s_io.hpp"
#include "cli.hpp"
#include "exr_io.hpp"
#include "raw_import.hpp"

using namespace haste;
using namespace b2pt_host;

namespace {

typedef std::map<std::string, std::string> Meta;

// Options::to_dict (Options.cpp:1185-1232)
Meta options_to_dict(const CliOptions& o, const std::string& scene_path) {
  using std::to_string;
  Meta m;
  m["options.input0"] = scene_path; m["options.input1"] = ""; m["options.output"] = o.output; m["options.reference"] = o.reference;
  m["options.technique"] = o.technique; m["options.action"] = "Render";
  m["options.num_photons"] = to_string(o.num_photons); m["options.radius"] = to_string(o.radius);
  m["options.max_path"] = to_string(size_t(PTRDIFF_MAX)); m["options.alpha"] = to_string(o.alpha); m["options.beta"] = to_string(o.beta);
  m["options.roulette"] = to_string(o.roulette); m["options.batch"] = "1"; m["options.quiet"] = to_string(o.quiet);
  m["options.enable_vc"] = to_string(o.enable_vc); m["options.enable_vm"] = to_string(o.enable_vm);
  m["options.from_light"] = to_string(o.from_light); m["options.lights"] = to_string(float(o.lights));
  m["options.num_samples"] = to_string(o.num_samples); m["options.num_seconds"] = to_string(o.num_seconds);
  m["options.num_threads"] = "1"; m["options.reload"] = "0"; m["options.enable_seed"] = "1"; m["options.enable_ui"] = "0";
  m["options.seed"] = to_string(o.seed); m["options.snapshot"] = "0"; m["options.camera_id"] = to_string(o.camera_id);
  m["options.width"] = to_string(o.width); m["options.height"] = to_string(o.height);
  auto x3 = [](const float* v) { return std::to_string(v[0]) + "x" + std::to_string(v[1]) + "x" + std::to_string(v[2]); };
  m["options.sky_horizon"] = x3(o.sky_horizon); m["options.sky_zenith"] = x3(o.sky_zenith);
  return m;
}

// the inverse for `continue`: Options(const map<string, string>&) (Options.cpp:1107-1183) for the fields the path reads
void options_from_dict(const Meta& m, CliOptions& o, std::string& scene_path) {
  auto get = [&](const char* k) { auto i = m.find(k); return i == m.end() ? std::string() : i->second; };
  auto num = [&](const char* k, double d) { const std::string v = get(k); return v.empty() ? d : atof(v.c_str()); };
  scene_path = get("options.input0");
  if (!get("options.technique").empty()) o.technique = get("options.technique");
  o.num_photons = size_t(num("options.num_photons", 0)); o.radius = num("options.radius", o.radius); o.alpha = num("options.alpha", o.alpha);
  o.beta = num("options.beta", o.beta); o.roulette = num("options.roulette", o.roulette); o.lights = num("options.lights", o.lights);
  o.enable_vc = num("options.enable_vc", 1) != 0; o.enable_vm = num("options.enable_vm", 1) != 0; o.from_light = num("options.from_light", 1) != 0;
  o.seed = size_t(num("options.seed", 1234)); o.enable_seed = true; o.camera_id = size_t(num("options.camera_id", 0));
  o.width = size_t(num("options.width", 512)); o.height = size_t(num("options.height", 512));
  const std::string h = get("options.sky_horizon"), z = get("options.sky_zenith");
  if (!h.empty()) sscanf(h.c_str(), "%fx%fx%f", o.sky_horizon, o.sky_horizon + 1, o.sky_horizon + 2);
  if (!z.empty()) sscanf(z.c_str(), "%fx%fx%f", o.sky_zenith, o.sky_zenith + 1, o.sky_zenith + 2);
}

// ---- `master merge` (main.cpp:90-92): the CPU analogue of the multi-GPU film reduce; the other image tools of main.cpp:80-107
// (average, errors, time, strip, bake, gnuplot) are post-processing and stay with the reference's own binary
int tool_merge(const std::string& dst, const std::string& fst, const std::string& snd) {  // merge_exr (Options.cpp:1340-1409)
  ExrImage a = read_exr(fst), b = read_exr(snd);
  if (a.width != b.width || a.height != b.height) throw std::runtime_error("Sizes of '" + fst + "' and '" + snd + "' doesn't match.");
  if (a.metadata["options.technique"] != b.metadata["options.technique"]) {
    std::cerr << "Cannot merge images rendered using different techniques." << std::endl;
    return 0;
  }
  for (size_t i = 0; i < a.rgbd.size(); ++i) a.rgbd[i] += b.rgbd[i];
  statistics_t s1(a.metadata), s2(b.metadata);
  double duration = 0;
  for (auto& r : s1.records) duration += r.frame_duration;
  for (auto& r : s2.records) duration += r.frame_duration;
  if (!s1.records.empty()) s1.records.erase(s1.records.begin(), s1.records.end() - 1);
  s1.num_samples += s2.num_samples; s1.num_basic_rays += s2.num_basic_rays; s1.num_shadow_rays += s2.num_shadow_rays;
  s1.num_tentative_rays += s2.num_tentative_rays; s1.num_photons += s2.num_photons; s1.num_scattered += s2.num_scattered;
  s1.total_time += s2.total_time; s1.scatter_time += s2.scatter_time; s1.build_time += s2.build_time; s1.gather_time += s2.gather_time;
  s1.merge_time += s2.merge_time; s1.density_time += s2.density_time; s1.intersect_time += s2.intersect_time;
  s1.trace_eye_time += s2.trace_eye_time; s1.trace_light_time += s2.trace_light_time;
  if (!s1.records.empty()) s1.records.back().frame_duration = float(duration);
  s1.measurements.clear();
  Meta meta = s1.to_dict();
  for (auto& kv : a.metadata) if (kv.first.rfind("options.", 0) == 0) meta[kv.first] = kv.second;
  meta["options.output"] = dst;
  save_exr(dst, meta, a.width, a.height, a.rgbd.data(), 4);
  return 0;
}


// --gpus=N: one host thread per device.  Thread g builds its own device scene (replicated BVH), renders the sample indices
// g, g + N, ... into a film that stays on its GPU, then all films are added into rank 0's with one b2pt_reduce_accum and rank 0
// brings the sum to the host once.  No other communication.
int render_multi_gpu(const CliOptions& o, const GpuOptions& base, const shared<Scene>& scene, const std::string& scene_path) {
  const int N = int(o.gpus);
  if (b2pt_device_count() < N) throw std::runtime_error("--gpus=" + std::to_string(N) + ": only " + std::to_string(b2pt_device_count()) + " CUDA devices are visible");
  if (o.num_samples == 0) throw std::runtime_error("--gpus needs --num-samples (the sample budget that is split over the GPUs)");
  const size_t W = o.width, H = o.height;
  char id[B2PT_COMM_ID_BYTES];
  std::memset(id, 0, sizeof id);
  if (N > 1 && b2pt_comm_unique_id(id) != B2PT_OK) throw std::runtime_error(std::string("b2pt_comm_unique_id: ") + b2pt_last_error());
  std::vector<dvec4> accum(W * H, dvec4(0.0));
  std::vector<std::string> errors(N);
  std::vector<statistics_t> stats(N);
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> threads;
  for (int r = 0; r < N; ++r) threads.emplace_back([&, r] {
    try {
      if (b2pt_set_device(r) != B2PT_OK) throw std::runtime_error(std::string("b2pt_set_device: ") + b2pt_last_error());
      GpuOptions g = base;
      g.first_sample = uint32_t(r); g.sample_stride = uint32_t(N);
      GpuTechnique technique(scene, g);
      technique.set_sky_gradient(vec3(o.sky_horizon[0], o.sky_horizon[1], o.sky_horizon[2]), vec3(o.sky_zenith[0], o.sky_zenith[1], o.sky_zenith[2]));
      b2pt_comm_t comm = nullptr;
      if (b2pt_comm_init(N, r, id, &comm) != B2PT_OK) throw std::runtime_error(std::string("b2pt_comm_init: ") + b2pt_last_error());
      const size_t mine = (o.num_samples - size_t(r) + size_t(N) - 1) / size_t(N);  // passes r, r + N, ... below num_samples
      const size_t chunk = o.passes_per_render > 1 ? o.passes_per_render : 16;
      for (size_t done = 0; done < mine; done += chunk)
        technique.render_resident(W, H, o.camera_id, uint32_t(std::min(chunk, mine - done)));
      if (mine == 0) technique.render_resident(W, H, o.camera_id, 0);  // creates the (empty) film this rank contributes
      technique.reduce_film(comm, 0);
      if (r == 0) { subimage_view_t view(accum.data(), W, H); technique.read_film(view); }
      else b2pt_device_synchronize();
      stats[r] = technique.statistics();
      b2pt_comm_destroy(comm);
    } catch (const std::exception& e) { errors[r] = e.what(); }
  });
  for (auto& t : threads) t.join();
  for (int r = 0; r < N; ++r) if (!errors[r].empty()) throw std::runtime_error("rank " + std::to_string(r) + ": " + errors[r]);
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  statistics_t total = stats[0];
  for (int r = 1; r < N; ++r) { total.num_samples += stats[r].num_samples; total.num_basic_rays += stats[r].num_basic_rays; total.num_shadow_rays += stats[r].num_shadow_rays; }
  total.total_time = sec;
  if (!o.raw_out.empty()) {
    std::ofstream f(o.raw_out, std::ios::binary);
    f.write((const char*)accum.data(), accum.size() * sizeof(dvec4));
  }
  if (!o.output.empty()) {
    Meta meta = total.to_dict();
    const Meta opts = options_to_dict(o, scene_path);
    meta.insert(opts.begin(), opts.end());
    std::vector<float> rgbd(W * H * 4);
    for (size_t i = 0; i < W * H; ++i) { rgbd[i * 4] = float(accum[i].x); rgbd[i * 4 + 1] = float(accum[i].y); rgbd[i * 4 + 2] = float(accum[i].z); rgbd[i * 4 + 3] = float(accum[i].w); }
    save_exr(o.output, meta, W, H, rgbd.data(), 4);
  }
  if (!o.quiet)
    printf("{\"width\": %zu, \"height\": %zu, \"num_samples\": %zu, \"gpus\": %d, \"seconds\": %.6f, \"basic_rays\": %zu, \"shadow_rays\": %zu, "
           "\"msamples_per_s\": %.6f, \"mrays_per_s\": %.6f}\n",
           W, H, size_t(total.num_samples), N, sec, size_t(total.num_basic_rays), size_t(total.num_shadow_rays),
           W * H * double(total.num_samples) / sec * 1e-6, double(total.num_basic_rays + total.num_shadow_rays) / sec * 1e-6);
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);  // main.cpp:70-71
  _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
  try {
    if (argc < 2) { fprintf(stderr, "usage: b2pt_master <scene.b2s> [--PT|--BPT|--VCM|--UPG] [options]   (the reference's `master` command line)\n"); return 2; }
    CliOptions o = parse_cli(argc, argv);
    if (o.action == "merge") return tool_merge(o.inputs[0], o.inputs[1], o.inputs[2]);
    if (o.action == "import") {  // .b2i (what an importer hands over) -> .b2s (what the device scene is built from): loader.cpp:293-456 on the host
      shared<Scene> imported = scene_from_import(read_b2i(o.inputs[0]));
      write_b2s(*imported, o.inputs[1]);
      size_t tris = 0;
      for (auto& m : imported->meshes) tris += m.indices.size() / 3;
      printf("{\"meshes\": %zu, \"triangles\": %zu, \"materials\": %zu, \"lights\": %zu, \"cameras\": %zu}\n", imported->meshes.size(), tris,
             imported->materials.bsdfs.size(), imported->lights.num_lights(), imported->cameras().numCameras());
      return 0;
    }

    std::string scene_path = o.inputs[0];
    std::vector<dvec4> accum;
    statistics_t restored;
    bool resumed = false;
    if (o.action == "continue") {
      // master continue: options, statistics and accumulators come from the checkpoint; the command line may raise the budget
      Meta meta; size_t w0 = 0, h0 = 0; std::vector<float> rgbd;
      load_exr(o.inputs[0], meta, w0, h0, rgbd);
      const size_t budget = o.num_samples; const double seconds = o.num_seconds; const std::string out = o.output, raw = o.raw_out;
      const size_t ppr = o.passes_per_render;
      options_from_dict(meta, o, scene_path);
      if (o.inputs.size() == 2) scene_path = o.inputs[1];
      if (o.width != w0 || o.height != h0) throw std::runtime_error("continue: the checkpoint's options do not match its size");
      o.num_samples = budget; o.num_seconds = seconds; o.raw_out = raw; o.passes_per_render = ppr;
      o.output = out.empty() ? o.inputs[0] : out;
      accum.resize(w0 * h0);
      for (size_t i = 0; i < w0 * h0; ++i) accum[i] = dvec4(rgbd[i * 4], rgbd[i * 4 + 1], rgbd[i * 4 + 2], rgbd[i * 4 + 3]);
      restored = statistics_t(meta);
      resumed = true;
    }
    if (!o.enable_seed) o.seed = 1234;
    shared<Scene> scene = read_b2s(scene_path);
    GpuOptions g;
    g.technique = o.technique == "PT" ? B2PT_TECHNIQUE_PT : o.technique == "BPT" ? B2PT_TECHNIQUE_BPT : o.technique == "VCM" ? B2PT_TECHNIQUE_VCM
                  : o.technique == "UPG" ? B2PT_TECHNIQUE_UPG : throw std::runtime_error("Unsupported technique " + o.technique + ".");
    g.beta = float(o.beta); g.roulette = float(o.roulette); g.lights = float(o.lights);
    g.radius = float(o.radius); g.alpha = float(o.alpha); g.num_photons = o.num_photons;
    g.enable_vc = o.enable_vc; g.enable_vm = o.enable_vm; g.from_light = o.from_light;
    g.passes_per_render = uint32_t(o.passes_per_render ? o.passes_per_render : 1);
    g.max_path = o.max_path;
    g.fixed_seed = true; g.seed = uint32_t(o.seed);
    const size_t W = o.width, H = o.height;
    // B2PT_FORCE_MULTI=1 takes the multi-GPU path with one rank too (tests of the resident film + reduce on a single-GPU box)
    if (o.gpus > 1 || (getenv("B2PT_FORCE_MULTI") && o.action == "render")) return render_multi_gpu(o, g, scene, scene_path);
    shared<Technique> technique = std::make_shared<GpuTechnique>(scene, g);  // makeTechnique's --gpu case
    technique->set_sky_gradient(vec3(o.sky_horizon[0], o.sky_horizon[1], o.sky_horizon[2]), vec3(o.sky_zenith[0], o.sky_zenith[1], o.sky_zenith[2]));
    if (resumed) technique->set_statistics(restored);
    else accum.assign(W * H, dvec4(0.0));
    std::vector<vec3> reference;
    if (!o.reference.empty()) {  // --reference: per-pass rms / abs errors in the records (Technique.cpp:67-74)
      ExrImage r = read_exr(o.reference);
      if (r.width != W || r.height != H) throw std::runtime_error("The reference image has a different size.");
      reference.resize(W * H);
      for (size_t i = 0; i < W * H; ++i) reference[i] = vec3(r.rgbd[i * 4], r.rgbd[i * 4 + 1], r.rgbd[i * 4 + 2]) / r.rgbd[i * 4 + 3];
    }
    subimage_view_t view(accum.data(), W, H);
    random_generator_t generator(size_t(g.seed));
    if (o.num_samples == 0 && o.num_seconds == 0.0) o.num_samples = 1;
    const auto t0 = std::chrono::steady_clock::now();
    auto elapsed = [&] { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    // Framework::runBatch + Application::updateScene's budget test (framework.cpp:426-437, Application.cpp:81-96)
    while ((o.num_samples == 0 || technique->statistics().num_samples < o.num_samples) && (o.num_seconds == 0.0 || elapsed() < o.num_seconds))
      technique->render(view, generator, o.camera_id, reference, {});
    const double sec = elapsed();
    const statistics_t& s = technique->statistics();
    if (!o.raw_out.empty()) {
      std::ofstream f(o.raw_out, std::ios::binary);
      f.write((const char*)accum.data(), accum.size() * sizeof(dvec4));
    }
    if (!o.output.empty()) {
      Meta meta = s.to_dict();  // fuse_metadata (Options.cpp:1254-1259)
      const Meta opts = options_to_dict(o, scene_path);
      meta.insert(opts.begin(), opts.end());
      std::vector<float> rgbd(W * H * 4);
      for (size_t i = 0; i < W * H; ++i) { rgbd[i * 4] = float(accum[i].x); rgbd[i * 4 + 1] = float(accum[i].y); rgbd[i * 4 + 2] = float(accum[i].z); rgbd[i * 4 + 3] = float(accum[i].w); }
      save_exr(o.output, meta, W, H, rgbd.data(), 4);  // save_exr(options, statistics, const dvec4*), Options.cpp:1296-1308
    }
    if (!o.quiet)
      printf("{\"width\": %zu, \"height\": %zu, \"num_samples\": %zu, \"seconds\": %.6f, \"basic_rays\": %zu, \"shadow_rays\": %zu, "
             "\"msamples_per_s\": %.6f, \"mrays_per_s\": %.6f}\n",
             W, H, size_t(s.num_samples), sec, size_t(s.num_basic_rays), size_t(s.num_shadow_rays),
             W * H * double(s.num_samples - (resumed ? restored.num_samples : 0)) / sec * 1e-6,
             double(s.num_basic_rays + s.num_shadow_rays - (resumed ? restored.num_basic_rays + restored.num_shadow_rays : 0)) / sec * 1e-6);
    return 0;
  } catch (const std::exception& e) {
    fprintf(stderr, "b2pt_master: %s\n", e.what());
    return 1;
  }
}
