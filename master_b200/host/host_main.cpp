// host_main.cpp — headless `master --batch` with the GPU technique plugged in: Framework::runBatch
// (framework.cpp:426-437) around GpuTechnique::render, the way Application::render drives any Technique
// (Application.cpp:41-79).  Scenes come from a .b2s (no Assimp on this path) or, when built with the
// loader, a .blend.  Output: the dvec4 accumulator as raw float64 (out=) and/or the reference's EXR checkpoint (exr=:
// R, G, B, denom + options.* / statistics.* / records[i].* attributes, exr_io.hpp); continue=<file.exr> resumes from such a
// checkpoint the way `master continue` does (Application.cpp / Options.cpp:1411-…: accumulators and statistics are restored,
// the sample index goes on from statistics.num_samples).
//
//   b2pt_master <scene.b2s> technique=BPT width=512 height=512 num_samples=16 [beta= roulette= seed= out= exr= continue= passes_per_render=]
#include <chrono>
#include <cstdio>
#include <fstream>
#include <map>

#include <xmmintrin.h>
#include <pmmintrin.h>

#include "GpuTechnique.hpp"
#include "b2s_io.hpp"
#include "exr_io.hpp"

using namespace haste;
using namespace b2pt_host;

int main(int argc, char** argv) {
  _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);  // main.cpp:70-71
  _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
  try {
    if (argc < 2) { fprintf(stderr, "usage: b2pt_master <scene.b2s> key=value ...\n"); return 2; }
    std::map<std::string, std::string> a;
    for (int i = 2; i < argc; ++i) {
      std::string s = argv[i];
      size_t e = s.find('=');
      if (e == std::string::npos) throw std::runtime_error("expected key=value, got " + s);
      a[s.substr(0, e)] = s.substr(e + 1);
    }
    auto num = [&](const char* k, double d) { auto i = a.find(k); return i == a.end() ? d : atof(i->second.c_str()); };
    auto str = [&](const char* k, const char* d) { auto i = a.find(k); return i == a.end() ? std::string(d) : i->second; };
    shared<Scene> scene = read_b2s(argv[1]);
    GpuOptions o;
    const std::string t = str("technique", "PT");
    o.technique = t == "PT" ? B2PT_TECHNIQUE_PT : t == "BPT" ? B2PT_TECHNIQUE_BPT : t == "VCM" ? B2PT_TECHNIQUE_VCM : B2PT_TECHNIQUE_UPG;
    o.beta = float(num("beta", t == "PT" ? 1.0 : 2.0));
    o.roulette = float(num("roulette", 0.9)); o.lights = float(num("lights", 1.0));
    o.radius = float(num("radius", 0.01)); o.alpha = float(num("alpha", 0.75));
    o.passes_per_render = uint32_t(num("passes_per_render", 1));
    o.fixed_seed = true; o.seed = uint32_t(num("seed", 1234));
    const size_t W = size_t(num("width", 512)), H = size_t(num("height", 512)), spp = size_t(num("num_samples", 1));
    shared<Technique> technique = std::make_shared<GpuTechnique>(scene, o);  // makeTechnique's --gpu case
    std::vector<dvec4> accum(W * H, dvec4(0.0));
    const std::string resume = str("continue", "");
    if (!resume.empty()) {
      std::map<std::string, std::string> meta; size_t w0 = 0, h0 = 0; std::vector<float> rgbd;
      load_exr(resume, meta, w0, h0, rgbd);
      if (w0 != W || h0 != H) throw std::runtime_error("continue: the checkpoint is " + std::to_string(w0) + "x" + std::to_string(h0));
      for (size_t i = 0; i < W * H; ++i) accum[i] = dvec4(rgbd[i * 4], rgbd[i * 4 + 1], rgbd[i * 4 + 2], rgbd[i * 4 + 3]);
      technique->set_statistics(statistics_t(meta));
    }
    subimage_view_t view(accum.data(), W, H);
    random_generator_t generator(size_t(o.seed));
    auto t0 = std::chrono::steady_clock::now();
    while (technique->statistics().num_samples < spp) technique->render(view, generator, size_t(num("camera", 0)), {}, {});
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const std::string out = str("out", "");
    if (!out.empty()) {
      std::ofstream f(out, std::ios::binary);
      f.write((const char*)accum.data(), accum.size() * sizeof(dvec4));
    }
    const statistics_t& s = technique->statistics();
    const std::string exr = str("exr", "");
    if (!exr.empty()) {
      // fuse_metadata (Options.cpp:1254-1259): statistics_t::to_dict() + Options::to_dict()
      std::map<std::string, std::string> meta = s.to_dict();
      using std::to_string;
      meta["options.input0"] = argv[1]; meta["options.input1"] = ""; meta["options.output"] = exr; meta["options.reference"] = "";
      meta["options.technique"] = t; meta["options.action"] = "Render";
      meta["options.num_photons"] = to_string(o.num_photons); meta["options.radius"] = to_string(o.radius);
      meta["options.max_path"] = to_string(o.max_path); meta["options.alpha"] = to_string(o.alpha); meta["options.beta"] = to_string(o.beta);
      meta["options.roulette"] = to_string(o.roulette); meta["options.batch"] = "1"; meta["options.quiet"] = "0";
      meta["options.enable_vc"] = to_string(o.enable_vc); meta["options.enable_vm"] = to_string(o.enable_vm);
      meta["options.from_light"] = to_string(o.from_light); meta["options.lights"] = to_string(o.lights);
      meta["options.num_samples"] = to_string(spp); meta["options.num_seconds"] = to_string(0.0); meta["options.num_threads"] = "0";
      meta["options.reload"] = "0"; meta["options.enable_seed"] = "1"; meta["options.enable_ui"] = "0"; meta["options.seed"] = to_string(o.seed);
      meta["options.snapshot"] = "0"; meta["options.camera_id"] = to_string(size_t(num("camera", 0)));
      meta["options.width"] = to_string(W); meta["options.height"] = to_string(H);
      meta["options.sky_horizon"] = "0.000000x0.000000x0.000000"; meta["options.sky_zenith"] = "0.000000x0.000000x0.000000";
      std::vector<float> rgbd(W * H * 4);
      for (size_t i = 0; i < W * H; ++i) { rgbd[i * 4] = float(accum[i].x); rgbd[i * 4 + 1] = float(accum[i].y); rgbd[i * 4 + 2] = float(accum[i].z); rgbd[i * 4 + 3] = float(accum[i].w); }
      save_exr(exr, meta, W, H, rgbd.data(), 4);  // save_exr(options, statistics, const dvec4*), Options.cpp:1296-1308
    }
    printf("{\"width\": %zu, \"height\": %zu, \"num_samples\": %zu, \"seconds\": %.6f, \"basic_rays\": %zu, \"shadow_rays\": %zu, "
           "\"msamples_per_s\": %.6f, \"mrays_per_s\": %.6f}\n",
           W, H, size_t(s.num_samples), sec, size_t(s.num_basic_rays), size_t(s.num_shadow_rays),
           W * H * s.num_samples / sec * 1e-6, (s.num_basic_rays + s.num_shadow_rays) / sec * 1e-6);
    return 0;
  } catch (const std::exception& e) {
    fprintf(stderr, "b2pt_master: %s\n", e.what());
    return 1;
  }
}
