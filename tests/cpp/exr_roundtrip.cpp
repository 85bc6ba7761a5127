// Test driver for master_b200/host/exr_io.hpp (tests/test_exr.py builds it with g++ -lz):
//   exr_roundtrip write <out.exr> <width> <height>   deterministic vec4 image + three attributes
//   exr_roundtrip read  <in.exr>                      prints width, height, has_denom, a checksum and the attributes as JSON
#include <cstdio>
#include <cstdlib>

#include "../../master_b200/host/exr_io.hpp"

static float value(size_t x, size_t y, int c) { return float((x * 31 + y * 17 + c * 7) % 1013) * 0.125f + (c == 3 ? 1.0f : 0.0f); }

int main(int argc, char** argv) {
  try {
    if (argc >= 5 && std::string(argv[1]) == "write") {
      const size_t w = size_t(atoi(argv[3])), h = size_t(atoi(argv[4]));
      std::vector<float> d(w * h * 4);
      for (size_t y = 0; y < h; ++y)
        for (size_t x = 0; x < w; ++x)
          for (int c = 0; c < 4; ++c) d[(y * w + x) * 4 + c] = value(x, y, c);
      std::map<std::string, std::string> m;
      m["options.technique"] = "BPT";
      m["statistics.num_samples"] = "16";
      m["measurements[100x512x512].rms_error"] = "0.250000";  // a name longer than 31 characters: needs the long-names flag
      b2pt_host::save_exr(argv[2], m, w, h, d.data(), 4);
      return 0;
    }
    if (argc >= 3 && std::string(argv[1]) == "read") {
      b2pt_host::ExrImage img = b2pt_host::read_exr(argv[2]);
      double sum[4] = {0, 0, 0, 0};
      for (size_t i = 0; i < img.width * img.height; ++i)
        for (int c = 0; c < 4; ++c) sum[c] += double(img.rgbd[i * 4 + c]) * double(1 + (i % 7));
      printf("{\"width\": %zu, \"height\": %zu, \"has_denom\": %d, \"sum\": [%.6f, %.6f, %.6f, %.6f], \"first\": [%.6f, %.6f, %.6f, %.6f], \"metadata\": {",
             img.width, img.height, int(img.has_denom), sum[0], sum[1], sum[2], sum[3], img.rgbd[0], img.rgbd[1], img.rgbd[2], img.rgbd[3]);
      bool first = true;
      for (auto& kv : img.metadata) { printf("%s\"%s\": \"%s\"", first ? "" : ", ", kv.first.c_str(), kv.second.c_str()); first = false; }
      printf("}}\n");
      return 0;
    }
    fprintf(stderr, "usage: exr_roundtrip write <out> <w> <h> | read <in>\n");
    return 2;
  } catch (const std::exception& e) {
    fprintf(stderr, "exr_roundtrip: %s\n", e.what());
    return 1;
  }
}
