// exr_io.hpp — the reference's checkpoint image format without the OpenEXR library.
//
// haste::save_exr / load_exr (exr.cpp:68-230, exr.hpp:15-66) keep a render as a single-part scanline OpenEXR file:
// FLOAT channels R, G, B and — for vec4 data — `denom` (the per-pixel sample count: the accumulators are stored
// UN-normalised so that `master continue` / `master merge` can add to them, exr.cpp:183-230), rows flipped so that the
// file is top-down while the buffers are y-up (exr.cpp:201-208), and every options.* / statistics.* / records[i].* entry
// as a string attribute of the header (exr.cpp:194-196, Options.cpp:1185-1232, statistics.cpp:118-175).
//
// This header writes that file (uncompressed scanline blocks, which every OpenEXR reader accepts) and reads back both
// uncompressed and ZIP / ZIPS blocks (what the reference's OpenEXR writer produces by default), so GPU renders can be
// consumed by the author's tools and the author's files by this code.  Same names and argument meaning as the reference.
#pragma once
#include <zlib.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace b2pt_host {

namespace exr_detail {

inline void put(std::vector<unsigned char>& o, const void* p, size_t n) { o.insert(o.end(), (const unsigned char*)p, (const unsigned char*)p + n); }
inline void put_str0(std::vector<unsigned char>& o, const std::string& s) { put(o, s.c_str(), s.size() + 1); }
inline void put_i32(std::vector<unsigned char>& o, int32_t v) { put(o, &v, 4); }
inline void put_f32(std::vector<unsigned char>& o, float v) { put(o, &v, 4); }
inline void put_attr(std::vector<unsigned char>& o, const std::string& name, const char* type, const void* value, size_t size) {
  put_str0(o, name); put_str0(o, type); put_i32(o, int32_t(size)); put(o, value, size);
}

struct Reader {
  const unsigned char* p; const unsigned char* end;
  void need(size_t n) const { if (size_t(end - p) < n) throw std::runtime_error("load_exr: truncated file"); }
  std::string str0() { const unsigned char* q = p; while (q < end && *q) ++q; if (q == end) throw std::runtime_error("load_exr: truncated header");
                       std::string s((const char*)p, q - p); p = q + 1; return s; }
  int32_t i32() { need(4); int32_t v; memcpy(&v, p, 4); p += 4; return v; }
  uint64_t u64() { need(8); uint64_t v; memcpy(&v, p, 8); p += 8; return v; }
};

}  // namespace exr_detail

// haste::save_exr(path, metadata, width, height, const vec4* data) (exr.cpp:183-230): `data` is width * height * `comps`
// floats, y-up; comps = 4 writes R, G, B, denom, comps = 3 writes R, G, B (exr.cpp:68-110).
inline void save_exr(const std::string& path, const std::map<std::string, std::string>& metadata, size_t width, size_t height,
                     const float* data, int comps = 4) {
  using namespace exr_detail;
  if (comps != 3 && comps != 4) throw std::runtime_error("save_exr: 3 or 4 components");
  if (width == 0 || height == 0 || width >= 0x7FFFFFFFu || height >= 0x7FFFFFFFu) throw std::runtime_error("save_exr: bad image size");
  // channels in the order the file stores them (sorted by name) and the component each one reads
  const char* names[4] = {"B", "G", "R", "denom"};
  const int comp_of[4] = {2, 1, 0, 3};
  const int nch = comps;
  std::vector<unsigned char> h;
  size_t longest = 5;
  for (auto& kv : metadata) longest = std::max(longest, kv.first.size());
  const unsigned char magic[4] = {0x76, 0x2f, 0x31, 0x01};
  put(h, magic, 4);
  put_i32(h, 2 | (longest > 31 ? 0x400 : 0));  // version 2; bit 10: attribute names longer than 31 characters
  {
    std::vector<unsigned char> cl;
    for (int c = 0; c < nch; ++c) {
      put_str0(cl, names[c]);
      put_i32(cl, 2);  // FLOAT
      const unsigned char lin[4] = {0, 0, 0, 0};
      put(cl, lin, 4);
      put_i32(cl, 1); put_i32(cl, 1);
    }
    cl.push_back(0);
    put_attr(h, "channels", "chlist", cl.data(), cl.size());
  }
  { const unsigned char none = 0; put_attr(h, "compression", "compression", &none, 1); }
  const int32_t window[4] = {0, 0, int32_t(width) - 1, int32_t(height) - 1};
  put_attr(h, "dataWindow", "box2i", window, 16);
  put_attr(h, "displayWindow", "box2i", window, 16);
  { const unsigned char inc = 0; put_attr(h, "lineOrder", "lineOrder", &inc, 1); }
  { const float one = 1.0f; put_attr(h, "pixelAspectRatio", "float", &one, 4); }
  { const float c[2] = {0.f, 0.f}; put_attr(h, "screenWindowCenter", "v2f", c, 8); }
  { const float one = 1.0f; put_attr(h, "screenWindowWidth", "float", &one, 4); }
  for (auto& kv : metadata) put_attr(h, kv.first, "string", kv.second.data(), kv.second.size());  // StringAttribute, exr.cpp:194-196
  h.push_back(0);
  const size_t row_bytes = size_t(nch) * width * 4;
  const uint64_t table_at = h.size();
  std::vector<uint64_t> offsets(height);
  for (size_t y = 0; y < height; ++y) offsets[y] = table_at + height * 8 + y * (8 + row_bytes);
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) throw std::runtime_error("save_exr: cannot open " + path);
  bool ok = fwrite(h.data(), 1, h.size(), f) == h.size() && fwrite(offsets.data(), 8, height, f) == height;
  std::vector<float> row(size_t(nch) * width);
  for (size_t y = 0; ok && y < height; ++y) {
    const float* src = data + (height - y - 1) * width * comps;  // file row y = buffer row height - y - 1 (exr.cpp:201-208)
    for (int c = 0; c < nch; ++c)
      for (size_t x = 0; x < width; ++x) row[size_t(c) * width + x] = src[x * comps + comp_of[c]];
    const int32_t head[2] = {int32_t(y), int32_t(row_bytes)};
    ok = fwrite(head, 4, 2, f) == 2 && fwrite(row.data(), 1, row_bytes, f) == row_bytes;
  }
  ok = (fclose(f) == 0) && ok;
  if (!ok) throw std::runtime_error("save_exr: write failed for " + path);
}

// The raw channels of a file: R, G, B, denom as stored (y-up again), plus every string attribute.
struct ExrImage {
  size_t width = 0, height = 0;
  bool has_denom = false;
  std::vector<float> rgbd;  // width * height * 4, y-up; denom = 1 when the file has no such channel
  std::map<std::string, std::string> metadata;
};

inline ExrImage read_exr(const std::string& path) {
  using namespace exr_detail;
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) throw std::runtime_error("load_exr: cannot open " + path);
  std::vector<unsigned char> buf;
  { unsigned char tmp[1 << 16]; size_t n; while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + n); }
  fclose(f);
  Reader r{buf.data(), buf.data() + buf.size()};
  r.need(8);
  if (!(r.p[0] == 0x76 && r.p[1] == 0x2f && r.p[2] == 0x31 && r.p[3] == 0x01)) throw std::runtime_error("load_exr: not an OpenEXR file: " + path);
  r.p += 4;
  const int32_t version = r.i32();
  if (version & 0x1A00) throw std::runtime_error("load_exr: tiled, deep and multi-part files are not what the renderer writes");
  struct Ch { std::string name; int type; };
  std::vector<Ch> chans;
  int compression = 0;
  int32_t dw[4] = {0, 0, -1, -1};
  ExrImage img;
  for (;;) {
    const std::string name = r.str0();
    if (name.empty()) break;
    const std::string type = r.str0();
    const int32_t size = r.i32();
    if (size < 0) throw std::runtime_error("load_exr: bad attribute size");
    r.need(size_t(size));
    const unsigned char* v = r.p;
    r.p += size;
    if (name == "channels") {
      Reader c{v, v + size};
      for (;;) {
        const std::string cn = c.str0();
        if (cn.empty()) break;
        Ch ch; ch.name = cn; ch.type = c.i32();
        c.need(12); c.p += 4;
        const int32_t xs = c.i32(), ys = c.i32();
        if (xs != 1 || ys != 1) throw std::runtime_error("load_exr: subsampled channels");
        chans.push_back(ch);
      }
    } else if (name == "compression") {
      compression = v[0];
    } else if (name == "dataWindow") {
      memcpy(dw, v, 16);
    } else if (type == "string") {
      img.metadata[name] = std::string((const char*)v, size_t(size));
    }
  }
  if (compression != 0 && compression != 2 && compression != 3) throw std::runtime_error("load_exr: only NONE / ZIPS / ZIP compression is read");
  const size_t W = size_t(dw[2] - dw[0] + 1), H = size_t(dw[3] - dw[1] + 1);
  if (dw[2] < dw[0] || dw[3] < dw[1]) throw std::runtime_error("load_exr: empty data window");
  img.width = W; img.height = H;
  img.rgbd.assign(W * H * 4, 0.0f);
  for (size_t i = 0; i < W * H; ++i) img.rgbd[i * 4 + 3] = 1.0f;
  size_t line_bytes = 0;
  for (auto& c : chans) line_bytes += W * (c.type == 1 ? 2 : 4);
  const size_t lines_per_block = compression == 3 ? 16 : 1;
  const size_t blocks = (H + lines_per_block - 1) / lines_per_block;
  std::vector<uint64_t> offs(blocks);
  for (auto& o : offs) o = r.u64();
  std::vector<unsigned char> raw, tmp;
  for (size_t b = 0; b < blocks; ++b) {
    if (offs[b] + 8 > buf.size()) throw std::runtime_error("load_exr: bad block offset");
    Reader br{buf.data() + offs[b], buf.data() + buf.size()};
    const int32_t y0 = br.i32() - dw[1];
    const int32_t csize = br.i32();
    if (y0 < 0 || size_t(y0) >= H || csize < 0) throw std::runtime_error("load_exr: bad block");
    br.need(size_t(csize));
    const size_t lines = std::min(lines_per_block, H - size_t(y0));
    const size_t want = lines * line_bytes;
    raw.resize(want);
    if (compression == 0 || size_t(csize) == want) {
      if (size_t(csize) != want) throw std::runtime_error("load_exr: bad block size");
      memcpy(raw.data(), br.p, want);
    } else {
      tmp.resize(want);
      uLongf got = uLongf(want);
      if (uncompress(tmp.data(), &got, br.p, uLong(csize)) != Z_OK || got != want) throw std::runtime_error("load_exr: zlib failed");
      for (size_t i = 1; i < want; ++i) tmp[i] = (unsigned char)(tmp[i - 1] + tmp[i] - 128);  // undo the byte predictor
      const size_t half = (want + 1) / 2;                                                      // undo the even / odd byte split
      for (size_t i = 0; i < want; ++i) raw[i] = (i & 1) ? tmp[half + i / 2] : tmp[i / 2];
    }
    const unsigned char* q = raw.data();
    for (size_t l = 0; l < lines; ++l) {
      const size_t row = H - 1 - (size_t(y0) + l);  // back to y-up (exr.cpp:150-165)
      for (auto& c : chans) {
        int comp = -1;
        if (c.name == "R") comp = 0; else if (c.name == "G") comp = 1; else if (c.name == "B") comp = 2;
        else if (c.name == "denom") { comp = 3; img.has_denom = true; }
        const size_t bytes = W * (c.type == 1 ? 2 : 4);
        if (comp >= 0) {
          if (c.type != 2) throw std::runtime_error("load_exr: channel " + c.name + " is not FLOAT");
          for (size_t x = 0; x < W; ++x) memcpy(&img.rgbd[(row * W + x) * 4 + comp], q + x * 4, 4);
        }
        q += bytes;
      }
    }
  }
  return img;
}

// haste::load_exr(path, metadata, width, height, vector<vec4>& data) (exr.cpp:232-…): the accumulators as stored
inline void load_exr(const std::string& path, std::map<std::string, std::string>& metadata, size_t& width, size_t& height,
                     std::vector<float>& rgbd) {
  ExrImage img = read_exr(path);
  metadata = img.metadata; width = img.width; height = img.height; rgbd.swap(img.rgbd);
}

}  // namespace b2pt_host
