// cli.hpp — the reference's command line for the headless GPU driver.
//
// parseArgs (Options.cpp:483-1033) restated for the options the sampling path reads: the same flags, the same defaults
// (Options.hpp:26-55), the same sub-commands for the image tools that only need the checkpoint files
// (main.cpp:90-92,108: merge, continue — the reference's other image tools are post-processing and out of this repo's scope).  `key=value` arguments (the form the tests
// use) are accepted next to `--flag[=value]`.  Display options (--no-ui, --no-reload, --snapshot) are parsed and ignored:
// there is no window on this box.
#pragma once
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace b2pt_host {

struct CliOptions {
  std::string action = "render";  // render | continue | merge | import
  std::vector<std::string> inputs;
  std::string technique = "PT";  // Options::technique = PT
  double beta = 1.0, roulette = 0.9, radius = 0.01, alpha = 0.75, lights = 1.0;
  size_t num_photons = 0, num_samples = 0, seed = 0, camera_id = 0, width = 512, height = 512, passes_per_render = 1;
  size_t gpus = 1, max_path = 0;  // --gpus=N: devices of this node (sample split + one film reduce); --max-path (PT only, PT.cpp:23-40)
  double num_seconds = 0.0;
  bool enable_vc = true, enable_vm = true, from_light = true, enable_seed = false, quiet = false;
  std::string output, reference, raw_out;
  float sky_horizon[3] = {0, 0, 0}, sky_zenith[3] = {0, 0, 0};
};

namespace cli_detail {
inline void parse3(const std::string& v, float* out, const char* what) {
  if (sscanf(v.c_str(), "%fx%fx%f", out, out + 1, out + 2) != 3) throw std::runtime_error(std::string("Invalid value for ") + what + ".");
}
inline bool is_number(const std::string& s) { char* e = nullptr; strtod(s.c_str(), &e); return !s.empty() && e && *e == 0; }
}  // namespace cli_detail

inline CliOptions parse_cli(int argc, char** argv) {
  using namespace cli_detail;
  CliOptions o;
  std::vector<std::string> positional;
  int num_techniques = 0, num_from = 0;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    std::string key, value;
    bool has_value = false;
    if (a.rfind("--", 0) == 0) {
      const size_t e = a.find('=');
      key = a.substr(2, e == std::string::npos ? std::string::npos : e - 2);
      if (e != std::string::npos) { value = a.substr(e + 1); has_value = true; }
    } else if (a.find('=') != std::string::npos && a.find_first_not_of("abcdefghijklmnopqrstuvwxyz_") == a.find('=')) {
      const size_t e = a.find('=');
      key = a.substr(0, e); value = a.substr(e + 1); has_value = true;
      for (auto& c : key) if (c == '_') c = '-';
    } else {
      positional.push_back(a);
      continue;
    }
    auto need = [&](bool numeric) {
      if (!has_value || (numeric && !is_number(value))) throw std::runtime_error("Invalid value for --" + key + ".");
    };
    if (key == "PT" || key == "BPT" || key == "VCM" || key == "UPG") { o.technique = key; ++num_techniques; }
    else if (key == "technique") { need(false); o.technique = value; ++num_techniques; }
    else if (key == "num-photons") { need(true); o.num_photons = size_t(atof(value.c_str())); }
    else if (key == "radius") { need(true); o.radius = atof(value.c_str()); }
    else if (key == "roulette") { need(true); o.roulette = atof(value.c_str()); }
    else if (key == "beta") { need(true); o.beta = atof(value.c_str()); }
    else if (key == "alpha") { need(true); o.alpha = atof(value.c_str()); }
    else if (key == "lights") { need(true); o.lights = atof(value.c_str()); }
    else if (key == "no-lights") o.lights = 0.0;
    else if (key == "no-vc") o.enable_vc = false;
    else if (key == "no-vm") o.enable_vm = false;
    else if (key == "from-camera") { o.from_light = false; ++num_from; }
    else if (key == "from-light") { if (has_value) o.from_light = atoi(value.c_str()) != 0; else o.from_light = true; ++num_from; }
    else if (key == "num-samples") { need(true); o.num_samples = size_t(atof(value.c_str())); }
    else if (key == "num-seconds") { need(true); o.num_seconds = atof(value.c_str()); }
    else if (key == "num-minutes") { need(true); o.num_seconds = atof(value.c_str()) * 60.0; }
    else if (key == "output" || key == "exr") { need(false); o.output = value; }
    else if (key == "out") { need(false); o.raw_out = value; }
    else if (key == "continue") { need(false); o.action = "continue"; positional.insert(positional.begin(), value); }
    else if (key == "reference") { need(false); o.reference = value; }
    else if (key == "seed") { need(true); o.seed = size_t(atof(value.c_str())); o.enable_seed = true; }
    else if (key == "camera") { need(true); o.camera_id = size_t(atof(value.c_str())); }
    else if (key == "width") { need(true); o.width = size_t(atof(value.c_str())); }
    else if (key == "height") { need(true); o.height = size_t(atof(value.c_str())); }
    else if (key == "passes-per-render") { need(true); o.passes_per_render = size_t(atof(value.c_str())); }
    else if (key == "gpus") { need(true); o.gpus = size_t(atof(value.c_str())); if (o.gpus < 1) throw std::runtime_error("Invalid value for --gpus."); }
    else if (key == "max-path") { need(true); o.max_path = size_t(atof(value.c_str())); }
    else if (key == "resolution") {
      unsigned w = 0, h = 0;
      if (!has_value || sscanf(value.c_str(), "%ux%u", &w, &h) != 2 || !w || !h) throw std::runtime_error("Invalid value for --resolution.");
      o.width = w; o.height = h;
    }
    else if (key == "sky-horizon") { need(false); parse3(value, o.sky_horizon, "--sky-horizon"); }
    else if (key == "sky-zenith") { need(false); parse3(value, o.sky_zenith, "--sky-zenith"); }
    else if (key == "blue-sky") { need(true); o.sky_horizon[0] = o.sky_horizon[1] = o.sky_horizon[2] = 0; o.sky_zenith[0] = o.sky_zenith[1] = 0; o.sky_zenith[2] = float(atof(value.c_str())); }
    else if (key == "quiet") o.quiet = true;
    else if (key == "batch" || key == "parallel" || key == "no-ui" || key == "no-reload" || key == "snapshot" || key == "trace") {}
    else throw std::runtime_error("Unsupported option --" + key + ".");
  }
  // Options.cpp:600-612, 672-680
  if (num_techniques > 1) throw std::runtime_error("More than one technique specified.");
  if (num_from > 1) throw std::runtime_error("--from-light and --from-camera cannot be specified at the same time.");
  if (positional.empty()) throw std::runtime_error("Input file is required.");
  static const char* actions[] = {"merge", "continue", "import"};
  for (const char* act : actions)
    if (positional[0] == act && o.action == "render") { o.action = act; positional.erase(positional.begin()); break; }
  o.inputs = positional;
  const size_t want = o.action == "merge" ? 3 : o.action == "import" ? 2 : 1;
  // `continue` takes the checkpoint; the scene comes from its options.input0 unless a second path overrides it
  if (o.inputs.size() != want && !(o.action == "continue" && o.inputs.size() == 2)) throw std::runtime_error("Input file is required.");
  return o;
}

}  // namespace b2pt_host
