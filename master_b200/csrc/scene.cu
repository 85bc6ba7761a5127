// scene.cu — b2pt_scene_create: upload + GPU BVH build.
//
// Replaces Scene::buildAccelStructs / updateRTCScene / makeRTCMesh (Scene.cpp:20-73), i.e. Embree's
// rtcNewTriangleMesh + rtcCommit (SAH builder with spatial splits, kernels/common/scene.cpp:100-124),
// with a build that runs entirely on the device:
//   1. per-triangle bounds + centroid, scene bounds by float atomics              (bounds_kernel)
//   2. 30-bit Morton code of the centroid, made unique with the triangle index    (morton_kernel)
//   3. 64-bit radix sort of the keys                                              (cub::DeviceRadixSort)
//   4. binary radix tree over the sorted keys, Karras 2012                        (radix_tree_kernel)
//   5. bottom-up box fitting with one atomic counter per inner node               (fit_kernel)
//   6. top-down collapse of the binary tree into 4-wide nodes, opening the child
//      with the largest surface area first; subtrees of <= 4 triangles become leaves (collapse_kernel, per level)
//   7. triangle + shading records written in leaf order                           (pack_kernel)
// The result is a different tree from Embree's; parity is defined on what the tree returns
// (closest triangle id, t, u, v), not on the tree.
#include <cub/cub.cuh>
#include <algorithm>
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "scene.cuh"

namespace b2pt {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
const char* get_error() { return g_error.c_str(); }
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof buf, "%s failed: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  g_error = buf;
  return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? B2PT_ERR_NO_DEVICE : B2PT_ERR_CUDA;
}

// ---------------------------------------------------------------- build kernels
struct BuildInput {
  uint32_t num_meshes, num_tris;
  const uint32_t* mesh_material_id;
  const uint32_t* mesh_vertex_offset;
  const uint32_t* mesh_index_offset;
  const float* vertices;
  const float* tangents;
  const int32_t* indices;
};

__device__ __forceinline__ uint32_t find_mesh(const BuildInput& in, uint32_t tri) {
  // last mesh m with index_offset[m] <= 3*tri
  uint32_t lo = 0, hi = in.num_meshes;
  uint32_t key = tri * 3u;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (in.mesh_index_offset[mid] <= key) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ float3 load_vertex(const BuildInput& in, uint32_t mesh, uint32_t corner_index) {
  uint32_t v = in.mesh_vertex_offset[mesh] + uint32_t(in.indices[corner_index]);
  return f3(in.vertices[v * 3 + 0], in.vertices[v * 3 + 1], in.vertices[v * 3 + 2]);
}

// monotone float <-> uint mapping so that atomicMin/atomicMax order floats
__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_to_float(uint32_t u) {
  u = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}

__global__ void bounds_kernel(BuildInput in, float4* tri_lo, float4* tri_hi, uint32_t* scene_bounds) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  float3 lo = f3(FLT_MAX), hi = f3(-FLT_MAX);
  if (t < in.num_tris) {
    uint32_t m = find_mesh(in, t);
    float3 a = load_vertex(in, m, t * 3 + 0), b = load_vertex(in, m, t * 3 + 1), c = load_vertex(in, m, t * 3 + 2);
    lo = f3(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)));
    hi = f3(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)));
    tri_lo[t] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(m));
    tri_hi[t] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(t));  // w: the source triangle of this reference (split_refs_kernel)
  }
  // warp reduce, then one atomic per warp and component
  for (int o = 16; o > 0; o >>= 1) {
    lo.x = fminf(lo.x, __shfl_xor_sync(0xFFFFFFFFu, lo.x, o));
    lo.y = fminf(lo.y, __shfl_xor_sync(0xFFFFFFFFu, lo.y, o));
    lo.z = fminf(lo.z, __shfl_xor_sync(0xFFFFFFFFu, lo.z, o));
    hi.x = fmaxf(hi.x, __shfl_xor_sync(0xFFFFFFFFu, hi.x, o));
    hi.y = fmaxf(hi.y, __shfl_xor_sync(0xFFFFFFFFu, hi.y, o));
    hi.z = fmaxf(hi.z, __shfl_xor_sync(0xFFFFFFFFu, hi.z, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(&scene_bounds[0], float_to_ordered(lo.x));
    atomicMin(&scene_bounds[1], float_to_ordered(lo.y));
    atomicMin(&scene_bounds[2], float_to_ordered(lo.z));
    atomicMax(&scene_bounds[3], float_to_ordered(hi.x));
    atomicMax(&scene_bounds[4], float_to_ordered(hi.y));
    atomicMax(&scene_bounds[5], float_to_ordered(hi.z));
  }
}


// ---------------------------------------------------------------- triangle pre-splitting (the builder's stand-in for Embree's spatial splits)
// Embree builds this scene with RTC_SCENE_HIGH_QUALITY = SAH + spatial splits (kernels/common/scene.cpp:100-124): a large or slanted
// triangle is referenced from several leaves, each with the part of its box that lies in that leaf.  A bottom-up clustering build
// cannot split while it builds, so the triangles are split BEFORE it, after Karras & Aila, "Fast Parallel Construction of High-Quality
// Bounding Volume Hierarchies" (HPG 2013, section 4.1), restated from the paper: every triangle gets a priority
//   (2^-level * (area(box) - area_ideal))^(1/3)     level = that of the coarsest scene-grid plane that cuts the box
// a budget of `budget * triangles` extra references is handed out in proportion to it, and a triangle's box is cut recursively at the
// coarsest grid plane that crosses it, each side shrunk to the bounds of the clipped triangle.  A reference is (box, triangle): the
// leaf test still tests the WHOLE triangle with Embree's arithmetic, so what a ray hits, and its t, u, v, do not change — only how many
// boxes and triangles it has to look at.
constexpr int kSplitGridBits = 20;
constexpr int kMaxRefsPerTri = 64;

struct SceneGrid { float3 lo, ext; };

__device__ __forceinline__ float comp(float3 v, int a) { return a == 0 ? v.x : a == 1 ? v.y : v.z; }
__device__ __forceinline__ void set_comp(float3& v, int a, float x) { if (a == 0) v.x = x; else if (a == 1) v.y = x; else v.z = x; }

// coarsest grid plane strictly inside (lo, hi) along the best axis; returns the bit of that plane (-1: the box is not cut by any plane)
__device__ __forceinline__ int split_plane(const SceneGrid& g, float3 lo, float3 hi, int& axis, float& pos) {
  int best_bit = -1; float best_ext = -1.f;
  axis = 0; pos = 0.f;
  const float scale = float(1u << kSplitGridBits);
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float l = comp(lo, a), h = comp(hi, a), o = comp(g.lo, a), e = comp(g.ext, a);
    const uint32_t il = uint32_t(fminf(fmaxf((l - o) / e * scale, 0.f), scale - 1.f));
    const uint32_t ih = uint32_t(fminf(fmaxf((h - o) / e * scale, 0.f), scale - 1.f));
    if (il == ih) continue;
    const int bit = 31 - __clz(int(il ^ ih));
    const float p = o + e * (float((ih >> bit) << bit) / scale);
    if (!(p > l && p < h)) continue;  // rounding put the plane on the border: nothing to gain
    if (bit > best_bit || (bit == best_bit && h - l > best_ext)) { best_bit = bit; best_ext = h - l; axis = a; pos = p; }
  }
  return best_bit;
}

__device__ __forceinline__ float box_half_area(float3 lo, float3 hi) {
  const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
  return dx * dy + dy * dz + dz * dx;
}

__global__ void split_priority_kernel(BuildInput in, SceneGrid g, const float4* tri_lo, const float4* tri_hi, float* priority, double* sum) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  float p = 0.f;
  if (t < in.num_tris) {
    const float4 lo4 = tri_lo[t], hi4 = tri_hi[t];
    const float3 lo = f3(lo4.x, lo4.y, lo4.z), hi = f3(hi4.x, hi4.y, hi4.z);
    int axis; float pos;
    const int bit = split_plane(g, lo, hi, axis, pos);
    if (bit >= 0) {
      const uint32_t m = __float_as_uint(lo4.w);
      const float3 a = load_vertex(in, m, t * 3 + 0), b = load_vertex(in, m, t * 3 + 1), c = load_vertex(in, m, t * 3 + 2);
      const float3 n = cross(b - a, c - a);
      const float ideal = fabsf(n.x) + fabsf(n.y) + fabsf(n.z);        // twice the sum of the three projected areas
      const float excess = 2.0f * box_half_area(lo, hi) - ideal;
      const float root = 2.0f * box_half_area(f3(0.f), g.ext);
      const int level = kSplitGridBits - bit;                           // 1 = the plane that halves the scene
      if (excess > 0.f && root > 0.f) p = cbrtf(exp2f(-float(level)) * excess / root);
    }
    priority[t] = p;
  }
  for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xFFFFFFFFu, p, o);
  if ((threadIdx.x & 31) == 0 && p > 0.f) atomicAdd(sum, double(p));
}

// total[0] += sum over triangles of min(floor(scale * priority), kMaxRefsPerTri - 1): the splits a scale would hand out
__global__ void split_budget_kernel(uint32_t n, const float* priority, float scale, unsigned long long* total) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned c = t < n ? unsigned(fminf(floorf(scale * priority[t]), float(kMaxRefsPerTri - 1))) : 0u;
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(total, (unsigned long long)c);
}

// One thread per triangle walks its split tree depth first.  kEmit = false: counts[t] = references produced;
// kEmit = true: writes them at offsets[t].  Both passes run the same arithmetic, so the counts are exact.
template <bool kEmit>
__global__ void split_refs_kernel(BuildInput in, SceneGrid g, const float4* tri_lo, const float4* tri_hi, const float* priority, float scale,
                                  uint32_t* counts, const uint32_t* offsets, float4* ref_lo, float4* ref_hi) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= in.num_tris) return;
  const float4 lo4 = tri_lo[t], hi4 = tri_hi[t];
  const int splits = int(fminf(floorf(scale * priority[t]), float(kMaxRefsPerTri - 1)));
  uint32_t produced = 0;
  const uint32_t base = kEmit ? offsets[t] : 0u;
  auto emit = [&](float3 lo, float3 hi) {
    if (kEmit) {
      ref_lo[base + produced] = make_float4(lo.x, lo.y, lo.z, lo4.w);
      ref_hi[base + produced] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(t));
    }
    ++produced;
  };
  if (splits <= 0) {
    emit(f3(lo4.x, lo4.y, lo4.z), f3(hi4.x, hi4.y, hi4.z));
  } else {
    const uint32_t m = __float_as_uint(lo4.w);
    const float3 v[3] = {load_vertex(in, m, t * 3 + 0), load_vertex(in, m, t * 3 + 1), load_vertex(in, m, t * 3 + 2)};
    struct Task { float3 lo, hi; int splits; };
    Task stack[kMaxRefsPerTri];
    int sp = 0;
    stack[sp++] = Task{f3(lo4.x, lo4.y, lo4.z), f3(hi4.x, hi4.y, hi4.z), splits};
    while (sp > 0) {
      const Task k = stack[--sp];
      int axis; float pos;
      if (k.splits <= 0 || split_plane(g, k.lo, k.hi, axis, pos) < 0) { emit(k.lo, k.hi); continue; }
      // bounds of the triangle on either side of the plane (Sutherland-Hodgman against one plane, kept as boxes)
      float3 llo = f3(INFINITY), lhi = f3(-INFINITY), rlo = f3(INFINITY), rhi = f3(-INFINITY);
      auto grow = [](float3& lo, float3& hi, float3 p) {
        lo = f3(fminf(lo.x, p.x), fminf(lo.y, p.y), fminf(lo.z, p.z)); hi = f3(fmaxf(hi.x, p.x), fmaxf(hi.y, p.y), fmaxf(hi.z, p.z));
      };
#pragma unroll
      for (int e = 0; e < 3; ++e) {
        const float3 p0 = v[e], p1 = v[(e + 1) % 3];
        const float c0 = comp(p0, axis), c1 = comp(p1, axis);
        if (c0 <= pos) grow(llo, lhi, p0);
        if (c0 >= pos) grow(rlo, rhi, p0);
        if ((c0 < pos && c1 > pos) || (c0 > pos && c1 < pos)) {
          const float w = fminf(fmaxf((pos - c0) / (c1 - c0), 0.f), 1.f);
          float3 q = p0 + (p1 - p0) * w;
          set_comp(q, axis, pos);
          grow(llo, lhi, q); grow(rlo, rhi, q);
        }
      }
      // each side stays inside the parent box and on its side of the plane
      float3 plo = k.lo, phi = k.hi;
      float3 cl_hi = phi; set_comp(cl_hi, axis, pos);
      float3 cr_lo = plo; set_comp(cr_lo, axis, pos);
      llo = f3(fmaxf(llo.x, plo.x), fmaxf(llo.y, plo.y), fmaxf(llo.z, plo.z)); lhi = f3(fminf(lhi.x, cl_hi.x), fminf(lhi.y, cl_hi.y), fminf(lhi.z, cl_hi.z));
      rlo = f3(fmaxf(rlo.x, cr_lo.x), fmaxf(rlo.y, cr_lo.y), fmaxf(rlo.z, cr_lo.z)); rhi = f3(fminf(rhi.x, phi.x), fminf(rhi.y, phi.y), fminf(rhi.z, phi.z));
      const bool l_ok = llo.x <= lhi.x && llo.y <= lhi.y && llo.z <= lhi.z, r_ok = rlo.x <= rhi.x && rlo.y <= rhi.y && rlo.z <= rhi.z;
      if (!l_ok || !r_ok) {  // the triangle's part in this box lies on one side only: keep the box that is not empty, with its splits
        if (l_ok) stack[sp++] = Task{llo, lhi, k.splits - 1};
        else if (r_ok) stack[sp++] = Task{rlo, rhi, k.splits - 1};
        else emit(k.lo, k.hi);
        continue;
      }
      const float wl = fmaxf(fmaxf(lhi.x - llo.x, lhi.y - llo.y), lhi.z - llo.z), wr = fmaxf(fmaxf(rhi.x - rlo.x, rhi.y - rlo.y), rhi.z - rlo.z);
      const int rest = k.splits - 1;
      int sl = wl + wr > 0.f ? int(float(rest) * wl / (wl + wr) + 0.5f) : rest / 2;
      sl = sl < 0 ? 0 : sl > rest ? rest : sl;
      stack[sp++] = Task{rlo, rhi, rest - sl};   // sp stays below kMaxRefsPerTri: a task with s splits yields at most s + 1 references
      stack[sp++] = Task{llo, lhi, sl};
    }
  }
  if (!kEmit) counts[t] = produced;
}

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

__global__ void morton_kernel(uint32_t n, const float4* tri_lo, const float4* tri_hi, const uint32_t* scene_bounds,
                              uint64_t* keys) {
  uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  float3 slo = f3(ordered_to_float(scene_bounds[0]), ordered_to_float(scene_bounds[1]), ordered_to_float(scene_bounds[2]));
  float3 shi = f3(ordered_to_float(scene_bounds[3]), ordered_to_float(scene_bounds[4]), ordered_to_float(scene_bounds[5]));
  float4 lo = tri_lo[t], hi = tri_hi[t];
  float3 c = f3(0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z));
  float3 ext = f3(fmaxf(shi.x - slo.x, 1e-30f), fmaxf(shi.y - slo.y, 1e-30f), fmaxf(shi.z - slo.z, 1e-30f));
  uint32_t x = uint32_t(fminf(fmaxf((c.x - slo.x) / ext.x * 1024.f, 0.f), 1023.f));
  uint32_t y = uint32_t(fminf(fmaxf((c.y - slo.y) / ext.y * 1024.f, 0.f), 1023.f));
  uint32_t z = uint32_t(fminf(fmaxf((c.z - slo.z) / ext.z * 1024.f, 0.f), 1023.f));
  uint32_t code = (expand_bits10(x) << 2) | (expand_bits10(y) << 1) | expand_bits10(z);
  keys[t] = (uint64_t(code) << 32) | t;
}

// binary radix tree node ids: inner nodes 0..n-2, leaves encoded as (i | 0x80000000)
struct BinNode {
  uint32_t left, right;   // child ids (bit 31 = leaf = sorted triangle position)
  uint32_t first, last;   // range of sorted triangle positions covered
};

__device__ __forceinline__ int delta(const uint64_t* keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  return __clzll(keys[i] ^ keys[j]);
}

__global__ void radix_tree_kernel(int n, const uint64_t* keys, BinNode* nodes, uint32_t* parent_inner, uint32_t* parent_leaf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  int dmin = delta(keys, n, i, i - d);
  int lmax = 2;
  while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  int j = i + l * d;
  int dnode = delta(keys, n, i, j);
  int s = 0;
  int t = l;
  do {
    t = (t + 1) >> 1;
    if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  int gamma = i + s * d + min(d, 0);
  int first = min(i, j), last = max(i, j);
  BinNode nd;
  nd.first = first; nd.last = last;
  if (first == gamma) { nd.left = uint32_t(gamma) | 0x80000000u; parent_leaf[gamma] = i; }
  else { nd.left = gamma; parent_inner[gamma] = i; }
  if (last == gamma + 1) { nd.right = uint32_t(gamma + 1) | 0x80000000u; parent_leaf[gamma + 1] = i; }
  else { nd.right = gamma + 1; parent_inner[gamma + 1] = i; }
  nodes[i] = nd;
  if (i == 0) parent_inner[0] = kInvalid;
}

__global__ void fit_kernel(int n, const uint64_t* keys, const float4* tri_lo, const float4* tri_hi, const BinNode* nodes,
                           const uint32_t* parent_inner, const uint32_t* parent_leaf, float4* box_lo, float4* box_hi,
                           uint32_t* visit) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t p = parent_leaf[i];
  while (p != kInvalid) {
    __threadfence();
    if (atomicAdd(&visit[p], 1u) == 0u) return;  // first arrival waits for the sibling
    BinNode nd = nodes[p];
    float4 llo, lhi, rlo, rhi;
    if (nd.left & 0x80000000u) { uint32_t t = uint32_t(keys[nd.left & 0x7FFFFFFFu]); llo = tri_lo[t]; lhi = tri_hi[t]; }
    else { llo = __ldcg(&box_lo[nd.left]); lhi = __ldcg(&box_hi[nd.left]); }  // L2: written by another SM
    if (nd.right & 0x80000000u) { uint32_t t = uint32_t(keys[nd.right & 0x7FFFFFFFu]); rlo = tri_lo[t]; rhi = tri_hi[t]; }
    else { rlo = __ldcg(&box_lo[nd.right]); rhi = __ldcg(&box_hi[nd.right]); }
    box_lo[p] = make_float4(fminf(llo.x, rlo.x), fminf(llo.y, rlo.y), fminf(llo.z, rlo.z), 0.f);
    box_hi[p] = make_float4(fmaxf(lhi.x, rhi.x), fmaxf(lhi.y, rhi.y), fmaxf(lhi.z, rhi.z), 0.f);
    p = parent_inner[p];
  }
}

// ---------------------------------------------------------------- PLOC: parallel locally-ordered clustering
// Bottom-up agglomerative build over the Morton-sorted triangles (Meister & Bittner 2017): every cluster looks
// kPlocRadius neighbours to either side for the partner that minimises the surface area of the union; mutual
// nearest neighbours merge; the survivors are compacted and the search repeats until one cluster is left.
// Cluster ids: [0, n) are the sorted triangles, [n, 2n-1) inner nodes in order of creation.
constexpr int kPlocRadiusDefault = 16;  // B2PT_PLOC_RADIUS overrides (A/B runs)

__global__ void ploc_init_kernel(uint32_t n, const uint64_t* keys, const float4* tri_lo, const float4* tri_hi, uint32_t* cid,
                                 float4* clo, float4* chi, float4* node_lo, float4* node_hi, uint32_t* size, uint32_t* parent) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t t = uint32_t(keys[i]);
  cid[i] = i;
  clo[i] = tri_lo[t]; chi[i] = tri_hi[t];
  node_lo[i] = tri_lo[t]; node_hi[i] = tri_hi[t];
  size[i] = 1u;
  parent[i] = kInvalid;
}

__global__ void ploc_nearest_kernel(uint32_t m, int radius, const float4* clo, const float4* chi, uint32_t* nn) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const float4 lo = clo[i], hi = chi[i];
  float best = INFINITY; uint32_t bj = kInvalid;
  const int j0 = int(i) - radius < 0 ? 0 : int(i) - radius;
  const int j1 = int(i) + radius >= int(m) ? int(m) - 1 : int(i) + radius;
  for (int j = j0; j <= j1; ++j) {
    if (j == int(i)) continue;
    const float4 l2 = clo[j], h2 = chi[j];
    const float dx = fmaxf(hi.x, h2.x) - fminf(lo.x, l2.x), dy = fmaxf(hi.y, h2.y) - fminf(lo.y, l2.y),
                dz = fmaxf(hi.z, h2.z) - fminf(lo.z, l2.z);
    const float a = dx * dy + dy * dz + dz * dx;
    if (a < best) { best = a; bj = uint32_t(j); }  // ties keep the smaller index: symmetric and deterministic
  }
  nn[i] = bj;
}

__global__ void ploc_merge_kernel(uint32_t m, uint32_t n, const uint32_t* nn, const uint32_t* cid, const float4* clo,
                                  const float4* chi, uint32_t* out_cid, float4* out_lo, float4* out_hi, uint32_t* valid,
                                  uint32_t* next_node, uint32_t* left, uint32_t* right, float4* node_lo, float4* node_hi,
                                  uint32_t* size, uint32_t* parent) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const uint32_t j = nn[i];
  const bool mutual = j != kInvalid && nn[j] == i;
  if (mutual && i > j) { valid[i] = 0u; return; }  // absorbed by the partner
  valid[i] = 1u;
  if (!mutual) { out_cid[i] = cid[i]; out_lo[i] = clo[i]; out_hi[i] = chi[i]; return; }
  const uint32_t id = atomicAdd(next_node, 1u);
  const uint32_t a = cid[i], b = cid[j];
  const float4 l1 = clo[i], h1 = chi[i], l2 = clo[j], h2 = chi[j];
  const float4 lo = make_float4(fminf(l1.x, l2.x), fminf(l1.y, l2.y), fminf(l1.z, l2.z), 0.f);
  const float4 hi = make_float4(fmaxf(h1.x, h2.x), fmaxf(h1.y, h2.y), fmaxf(h1.z, h2.z), 0.f);
  left[id - n] = a; right[id - n] = b;
  node_lo[id] = lo; node_hi[id] = hi;
  size[id] = size[a] + size[b];
  parent[a] = id; parent[b] = id; parent[id] = kInvalid;
  out_cid[i] = id; out_lo[i] = lo; out_hi[i] = hi;
}

__global__ void ploc_compact_kernel(uint32_t m, const uint32_t* valid, const uint32_t* offset, const uint32_t* in_cid,
                                    const float4* in_lo, const float4* in_hi, uint32_t* cid, float4* clo, float4* chi) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m || !valid[i]) return;
  const uint32_t o = offset[i];
  cid[o] = in_cid[i]; clo[o] = in_lo[i]; chi[o] = in_hi[i];
}

// position of a cluster's first triangle in depth-first (left to right) leaf order
__global__ void ploc_first_kernel(uint32_t total, uint32_t n, const uint32_t* parent, const uint32_t* left, const uint32_t* size,
                                  uint32_t* first) {
  uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= total) return;
  uint32_t pos = 0, k = c;
  for (uint32_t p = parent[k]; p != kInvalid; k = p, p = parent[k]) {
    const uint32_t l = left[p - n];
    if (l != k) pos += size[l];
  }
  first[c] = pos;
}

// PLOC tree -> the BinNode form collapse_kernel reads; the leaf order becomes the new triangle order
__global__ void ploc_finish_kernel(uint32_t n, const uint64_t* keys, const uint32_t* left, const uint32_t* right, const uint32_t* size,
                                   const uint32_t* first, const float4* node_lo, const float4* node_hi, BinNode* bin, float4* box_lo,
                                   float4* box_hi, uint64_t* order) {
  uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= 2u * n - 1u) return;
  if (c < n) { order[first[c]] = keys[c] & 0xFFFFFFFFull; return; }
  const uint32_t k = c - n, l = left[k], r = right[k];
  BinNode nd;
  nd.left = l < n ? (first[l] | 0x80000000u) : l - n;
  nd.right = r < n ? (first[r] | 0x80000000u) : r - n;
  nd.first = first[c]; nd.last = first[c] + size[c] - 1u;
  bin[k] = nd;
  box_lo[k] = node_lo[c]; box_hi[k] = node_hi[c];
}

struct CollapseItem { uint32_t wide, bin; };

__device__ __forceinline__ float half_area(float4 lo, float4 hi) {
  float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
  return dx * dy + dy * dz + dz * dx;
}

// One thread per wide node of the current level.
__global__ void collapse_kernel(uint32_t count, const CollapseItem* items, const uint64_t* keys, const float4* tri_lo,
                                const float4* tri_hi, const BinNode* bin, const float4* box_lo, const float4* box_hi,
                                float4* nodes, uint32_t* num_wide, CollapseItem* next_items, uint32_t* next_count,
                                float pad) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  CollapseItem it = items[i];
  uint32_t ch[4];
  int nch = 2;
  ch[0] = bin[it.bin].left; ch[1] = bin[it.bin].right;
  auto is_final = [&](uint32_t c) {  // a child that will not be opened further
    if (c & 0x80000000u) return true;
    return bin[c].last - bin[c].first + 1u <= uint32_t(kMaxLeafTris);
  };
  while (nch < 4) {
    int best = -1; float best_area = -1.f;
    for (int k = 0; k < nch; ++k) {
      if (is_final(ch[k])) continue;
      float a = half_area(box_lo[ch[k]], box_hi[ch[k]]);
      if (a > best_area) { best_area = a; best = k; }
    }
    if (best < 0) break;
    uint32_t c = ch[best];
    ch[best] = bin[c].left;
    ch[nch++] = bin[c].right;
  }
  float lo[3][4], hi[3][4];
  uint32_t link[4];
  for (int k = 0; k < 4; ++k) {
    if (k >= nch) {
      for (int a = 0; a < 3; ++a) { lo[a][k] = INFINITY; hi[a][k] = -INFINITY; }
      link[k] = kInvalid;
      continue;
    }
    uint32_t c = ch[k];
    float4 blo, bhi;
    if (c & 0x80000000u) {
      uint32_t pos = c & 0x7FFFFFFFu;
      uint32_t t = uint32_t(keys[pos]);
      blo = tri_lo[t]; bhi = tri_hi[t];
      link[k] = make_leaf(pos, 1);
    } else {
      blo = box_lo[c]; bhi = box_hi[c];
      uint32_t cnt = bin[c].last - bin[c].first + 1u;
      if (cnt <= uint32_t(kMaxLeafTris)) {
        link[k] = make_leaf(bin[c].first, cnt);
      } else {
        uint32_t w = atomicAdd(num_wide, 1u);
        uint32_t q = atomicAdd(next_count, 1u);
        next_items[q] = CollapseItem{w, c};
        link[k] = w;
      }
    }
    lo[0][k] = blo.x - pad; lo[1][k] = blo.y - pad; lo[2][k] = blo.z - pad;
    hi[0][k] = bhi.x + pad; hi[1][k] = bhi.y + pad; hi[2][k] = bhi.z + pad;
  }
  float4* out = nodes + size_t(it.wide) * kNodeFloat4s;
  for (int a = 0; a < 3; ++a) {
    out[a] = make_float4(lo[a][0], lo[a][1], lo[a][2], lo[a][3]);
    out[3 + a] = make_float4(hi[a][0], hi[a][1], hi[a][2], hi[a][3]);
  }
  out[6] = make_float4(__uint_as_float(link[0]), __uint_as_float(link[1]), __uint_as_float(link[2]), __uint_as_float(link[3]));
  out[7] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ---------------------------------------------------------------- collapse to the compressed 8-wide tree (bvh8.cuh)
// One thread per wide node of the current level: open the child with the largest surface area until eight children
// (subtrees of <= 3 triangles stay closed and become leaves), quantise the child boxes on the node's power-of-two
// grid, give every child the slot whose octant direction its centroid points along most (greedy assignment), make
// the inner children and the leaf triangles of the node contiguous.
__global__ void collapse8_kernel(uint32_t count, const CollapseItem* items, const uint64_t* keys, const float4* tri_lo, const float4* tri_hi,
                                 const BinNode* bin, const float4* box_lo, const float4* box_hi, float4* nodes8, uint32_t* num_wide,
                                 CollapseItem* next_items, uint32_t* next_count, uint32_t* tri_counter, uint64_t* tri_map, float pad,
                                 uint32_t single_leaf_count, uint32_t leaf_max) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const CollapseItem it = items[i];
  uint32_t ref[8], cnt[8];
  float lo[8][3], hi[8][3];
  int n = 0;
  auto fetch = [&](uint32_t c, int k) {
    ref[k] = c;
    float4 a, b;
    if (c & 0x80000000u) { const uint32_t t = uint32_t(keys[c & 0x7FFFFFFFu]); a = tri_lo[t]; b = tri_hi[t]; cnt[k] = 1u; }
    else { a = box_lo[c]; b = box_hi[c]; cnt[k] = bin[c].last - bin[c].first + 1u; }
    lo[k][0] = a.x; lo[k][1] = a.y; lo[k][2] = a.z; hi[k][0] = b.x; hi[k][1] = b.y; hi[k][2] = b.z;
  };
  if (single_leaf_count) {  // a scene of <= 3 triangles: the root has one leaf child holding all of them
    ref[0] = 0x80000000u; cnt[0] = single_leaf_count;
    for (int a = 0; a < 3; ++a) { lo[0][a] = INFINITY; hi[0][a] = -INFINITY; }
    for (uint32_t j = 0; j < single_leaf_count; ++j) {
      const uint32_t t = uint32_t(keys[j]);
      const float4 a = tri_lo[t], b = tri_hi[t];
      lo[0][0] = fminf(lo[0][0], a.x); lo[0][1] = fminf(lo[0][1], a.y); lo[0][2] = fminf(lo[0][2], a.z);
      hi[0][0] = fmaxf(hi[0][0], b.x); hi[0][1] = fmaxf(hi[0][1], b.y); hi[0][2] = fmaxf(hi[0][2], b.z);
    }
    n = 1;
  } else {
    fetch(bin[it.bin].left, 0); fetch(bin[it.bin].right, 1);
    n = 2;
    while (n < 8) {
      int best = -1; float best_area = -1.f;
      for (int k = 0; k < n; ++k) {
        if ((ref[k] & 0x80000000u) || cnt[k] <= leaf_max) continue;
        const float dx = hi[k][0] - lo[k][0], dy = hi[k][1] - lo[k][1], dz = hi[k][2] - lo[k][2];
        const float a = dx * dy + dy * dz + dz * dx;
        if (a > best_area) { best_area = a; best = k; }
      }
      if (best < 0) break;
      const uint32_t c = ref[best];
      fetch(bin[c].left, best);
      fetch(bin[c].right, n++);
    }
  }
  // ---- the node's grid
  float p[3], ext[3], scale[3], inv[3];
  uint32_t ebits[3];
  for (int a = 0; a < 3; ++a) {
    float mn = INFINITY, mx = -INFINITY;
    for (int k = 0; k < n; ++k) { mn = fminf(mn, lo[k][a]); mx = fmaxf(mx, hi[k][a]); }
    p[a] = mn - pad;
    ext[a] = (mx + pad) - p[a];
    int e = 0;
    frexpf(ext[a] / 255.0f, &e);  // ext / 255 = m * 2^e with m in [0.5, 1): 2^e covers it
    e = max(-100, min(100, e));
    ebits[a] = uint32_t(e + 127);
    scale[a] = __uint_as_float(ebits[a] << 23);
    inv[a] = __uint_as_float(uint32_t(127 - e) << 23);
  }
  // ---- slots: greedy assignment of children to octant directions
  int slot_of[8];
  int child_in[8];
  for (int k = 0; k < 8; ++k) { slot_of[k] = -1; child_in[k] = -1; }
  float cen[8][3];
  for (int k = 0; k < n; ++k)
    for (int a = 0; a < 3; ++a) cen[k][a] = 0.5f * (lo[k][a] + hi[k][a]) - (p[a] + 0.5f * ext[a]);
  for (int round = 0; round < n; ++round) {
    float best = -INFINITY; int bk = -1, bs = -1;
    for (int k = 0; k < n; ++k) {
      if (slot_of[k] >= 0) continue;
      for (int sl = 0; sl < 8; ++sl) {
        if (child_in[sl] >= 0) continue;
        const float c = ((sl & 1) ? cen[k][0] : -cen[k][0]) + ((sl & 2) ? cen[k][1] : -cen[k][1]) + ((sl & 4) ? cen[k][2] : -cen[k][2]);
        if (c > best) { best = c; bk = k; bs = sl; }
      }
    }
    slot_of[bk] = bs; child_in[bs] = bk;
  }
  // ---- children: inner nodes contiguous, leaf triangles contiguous
  uint32_t n_inner = 0, n_tris = 0;
  for (int k = 0; k < n; ++k) {
    const bool leaf = single_leaf_count || (ref[k] & 0x80000000u) || cnt[k] <= leaf_max;
    if (leaf) n_tris += cnt[k]; else ++n_inner;
  }
  const uint32_t child_base = n_inner ? atomicAdd(num_wide, n_inner) : 0u;
  const uint32_t tri_base = n_tris ? atomicAdd(tri_counter, n_tris) : 0u;
  uint32_t imask = 0u, rank = 0u, offset = 0u;
  uint32_t meta[8], q[6][8];
  for (int sl = 0; sl < 8; ++sl) {
    const int k = child_in[sl];
    meta[sl] = 0u;
    for (int a = 0; a < 3; ++a) { q[a][sl] = 255u; q[3 + a][sl] = 0u; }  // empty slot: inverted box
    if (k < 0) continue;
    for (int a = 0; a < 3; ++a) {
      const float l = floorf(((lo[k][a] - pad) - p[a]) * inv[a]), h = ceilf(((hi[k][a] + pad) - p[a]) * inv[a]);
      q[a][sl] = uint32_t(fminf(fmaxf(l, 0.f), 255.f));
      q[3 + a][sl] = uint32_t(fminf(fmaxf(h, 0.f), 255.f));
    }
    const bool leaf = single_leaf_count || (ref[k] & 0x80000000u) || cnt[k] <= leaf_max;
    if (!leaf) {
      imask |= 1u << sl;
      const uint32_t w = child_base + rank++;
      const uint32_t qi = atomicAdd(next_count, 1u);
      next_items[qi] = CollapseItem{w, ref[k]};
      meta[sl] = 0x20u | (24u + uint32_t(sl));  // one hit bit at position 24 + slot (the traversal XORs the ray octant into the slot)
    } else {
      const uint32_t first = single_leaf_count ? 0u : ((ref[k] & 0x80000000u) ? (ref[k] & 0x7FFFFFFFu) : bin[ref[k]].first);
      meta[sl] = (((1u << cnt[k]) - 1u) << 5) | offset;
      for (uint32_t j = 0; j < cnt[k]; ++j) tri_map[tri_base + offset + j] = keys[first + j] & 0xFFFFFFFFull;
      offset += cnt[k];
    }
  }
  auto pack4 = [](const uint32_t* b) { return b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24); };
  float4* out = nodes8 + size_t(it.wide) * 5;
  out[0] = make_float4(p[0], p[1], p[2], __uint_as_float(ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (imask << 24)));
  out[1] = make_float4(__uint_as_float(child_base), __uint_as_float(tri_base), __uint_as_float(pack4(meta)), __uint_as_float(pack4(meta + 4)));
  out[2] = make_float4(__uint_as_float(pack4(q[0])), __uint_as_float(pack4(q[0] + 4)), __uint_as_float(pack4(q[1])), __uint_as_float(pack4(q[1] + 4)));
  out[3] = make_float4(__uint_as_float(pack4(q[2])), __uint_as_float(pack4(q[2] + 4)), __uint_as_float(pack4(q[3])), __uint_as_float(pack4(q[3] + 4)));
  out[4] = make_float4(__uint_as_float(pack4(q[4])), __uint_as_float(pack4(q[4] + 4)), __uint_as_float(pack4(q[5])), __uint_as_float(pack4(q[5] + 4)));
}

// Triangle + shading records in leaf (sorted) order, one per REFERENCE (a split triangle is stored once per leaf that holds a
// part of it: the leaf test reads the record next to its siblings instead of chasing an index).
__global__ void pack_kernel(BuildInput in, uint32_t num_refs, const uint64_t* keys, const float4* ref_lo, const float4* ref_hi, float4* tris, float4* frames) {
  uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= num_refs) return;
  const uint32_t r = uint32_t(keys[pos]);
  const uint32_t t = __float_as_uint(ref_hi[r].w);
  uint32_t m = __float_as_uint(ref_lo[r].w);
  uint32_t prim = t - in.mesh_index_offset[m] / 3u;
  float3 a = load_vertex(in, m, t * 3 + 0), b = load_vertex(in, m, t * 3 + 1), c = load_vertex(in, m, t * 3 + 2);
  float3 e1 = a - b, e2 = c - a;  // Embree: e1 = v0 - v1, e2 = v2 - v0
  uint32_t material = in.mesh_material_id[m];
  uint32_t mask = 1u << (material & 3u);  // rtcSetMask, Scene.cpp:47
  float4* tr = tris + size_t(pos) * kTriFloat4s;
  tr[0] = make_float4(a.x, a.y, a.z, __uint_as_float(m));
  tr[1] = make_float4(e1.x, e1.y, e1.z, __uint_as_float(prim));
  tr[2] = make_float4(e2.x, e2.y, e2.z, __uint_as_float(mask));
  float fr[28];
  for (int k = 0; k < 3; ++k) {
    uint32_t v = in.mesh_vertex_offset[m] + uint32_t(in.indices[t * 3 + k]);
    for (int j = 0; j < 9; ++j) fr[k * 9 + j] = in.tangents[size_t(v) * 9 + j];
  }
  fr[27] = __uint_as_float(material);
  float4* out = frames + size_t(pos) * kFrameFloat4s;
  for (int k = 0; k < kFrameFloat4s; ++k) out[k] = make_float4(fr[k * 4], fr[k * 4 + 1], fr[k * 4 + 2], fr[k * 4 + 3]);
}

// ---------------------------------------------------------------- host side
template <class T>
static int dev_alloc(b2pt_scene_s* s, T** p, size_t count, bool keep) {
  void* q = nullptr;
  B2_CUDA(cudaMalloc(&q, count * sizeof(T) + 16));
  *p = (T*)q;
  if (keep) s->allocations[s->num_allocations++] = q;
  return B2PT_OK;
}

static int build_scene(const B2ptSceneDesc* d, b2pt_scene_s* s) {
  const uint32_t M = d->num_meshes;
  const uint32_t nv = d->mesh_vertex_offset[M], ni = d->mesh_index_offset[M];
  const uint32_t nt = ni / 3;  // source triangles
  uint32_t n = nt;             // references = leaves of the tree (> nt once triangles are pre-split)
  B2_CHECK(nt >= 1, B2PT_ERR_INVALID, "scene has no triangles");
  B2_CHECK(nt < (1u << 27), B2PT_ERR_CAPACITY, "more than 2^27 triangles");
  for (uint32_t m = 0; m < M; ++m)
    B2_CHECK(d->mesh_index_offset[m] % 3 == 0, B2PT_ERR_INVALID, "mesh index count not a multiple of 3");

  std::vector<void*> temps;
  auto tmp = [&](size_t bytes) -> void* { void* q = nullptr; if (cudaMalloc(&q, bytes + 16) != cudaSuccess) return nullptr; temps.push_back(q); return q; };
  struct Cleanup { std::vector<void*>& t; ~Cleanup() { for (void* q : t) cudaFree(q); } } cleanup{temps};

  cudaEvent_t ev0, ev1;
  B2_CUDA(cudaEventCreate(&ev0)); B2_CUDA(cudaEventCreate(&ev1));

  // ---- upload (inputs are temporaries except the small tables)
  BuildInput in;
  in.num_meshes = M; in.num_tris = nt;
  uint32_t* d_mat = (uint32_t*)tmp(M * 4); uint32_t* d_voff = (uint32_t*)tmp((M + 1) * 4); uint32_t* d_ioff = (uint32_t*)tmp((M + 1) * 4);
  float* d_vert = (float*)tmp(size_t(nv) * 12); float* d_tan = (float*)tmp(size_t(nv) * 36); int32_t* d_idx = (int32_t*)tmp(size_t(ni) * 4);
  B2_CHECK(d_mat && d_voff && d_ioff && d_vert && d_tan && d_idx, B2PT_ERR_CUDA, "cudaMalloc failed for scene upload");
  B2_CUDA(cudaMemcpy(d_mat, d->mesh_material_id, M * 4, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(d_voff, d->mesh_vertex_offset, (M + 1) * 4, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(d_ioff, d->mesh_index_offset, (M + 1) * 4, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(d_vert, d->vertices, size_t(nv) * 12, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(d_tan, d->tangents, size_t(nv) * 36, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(d_idx, d->indices, size_t(ni) * 4, cudaMemcpyHostToDevice));
  in.mesh_material_id = d_mat; in.mesh_vertex_offset = d_voff; in.mesh_index_offset = d_ioff;
  in.vertices = d_vert; in.tangents = d_tan; in.indices = d_idx;

  B2_CUDA(cudaEventRecord(ev0));
  const int B = 256;
  uint32_t G = (nt + B - 1) / B;
  float4* tri_lo = (float4*)tmp(size_t(nt) * 16); float4* tri_hi = (float4*)tmp(size_t(nt) * 16);
  uint32_t* bounds = (uint32_t*)tmp(6 * 4);
  B2_CHECK(tri_lo && tri_hi && bounds, B2PT_ERR_CUDA, "cudaMalloc failed for BVH build");
  uint32_t init_bounds[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
  B2_CUDA(cudaMemcpy(bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice));
  bounds_kernel<<<G, B>>>(in, tri_lo, tri_hi, bounds);
  uint32_t h_bounds[6];
  B2_CUDA(cudaMemcpy(h_bounds, bounds, sizeof h_bounds, cudaMemcpyDeviceToHost));
  float ext = 0.f;
  SceneGrid grid;
  {
    float lo[3], ex[3];
    for (int a = 0; a < 3; ++a) { lo[a] = ordered_to_float(h_bounds[a]); ex[a] = fmaxf(ordered_to_float(h_bounds[3 + a]) - lo[a], 1e-30f); ext = fmaxf(ext, ex[a]); }
    grid.lo = f3(lo[0], lo[1], lo[2]);
    grid.ext = f3(ext, ext, ext);  // one cubic grid: its planes are what the Morton order of the build splits at
    for (int a = 0; a < 3; ++a) { s->dev.bounds_lo[a] = lo[a]; s->dev.bounds_inv_ext[a] = 1.0f / ex[a]; }
  }

  // ---- triangle pre-splitting: budget * triangles extra references (B2PT_SPLIT_BUDGET, default kSplitBudgetDefault; 0 = off)
  float split_budget = kSplitBudgetDefault;
  { const char* e = getenv("B2PT_SPLIT_BUDGET"); if (e) split_budget = float(atof(e)); }
  if (split_budget > 0.f && nt > 3) {
    float* priority = (float*)tmp(size_t(nt) * 4);
    double* psum = (double*)tmp(16);
    unsigned long long* total = (unsigned long long*)tmp(16);
    uint32_t* counts = (uint32_t*)tmp(size_t(nt) * 4); uint32_t* offsets = (uint32_t*)tmp(size_t(nt) * 4);
    B2_CHECK(priority && psum && total && counts && offsets, B2PT_ERR_CUDA, "cudaMalloc failed for triangle splitting");
    B2_CUDA(cudaMemset(psum, 0, 8));
    split_priority_kernel<<<G, B>>>(in, grid, tri_lo, tri_hi, priority, psum);
    double h_psum = 0.0;
    B2_CUDA(cudaMemcpy(&h_psum, psum, 8, cudaMemcpyDeviceToHost));
    if (h_psum > 0.0) {
      // the scale that hands out the budget: floor() and the per-triangle cap lose some, so bisect upwards from budget / sum
      const double want = double(split_budget) * nt;
      double lo_s = want / h_psum, hi_s = lo_s * 16.0;
      float scale = float(lo_s);
      for (int it = 0; it < 8; ++it) {
        const double mid = 0.5 * (lo_s + hi_s);
        B2_CUDA(cudaMemset(total, 0, 8));
        split_budget_kernel<<<G, B>>>(nt, priority, float(mid), total);
        unsigned long long h_total = 0;
        B2_CUDA(cudaMemcpy(&h_total, total, 8, cudaMemcpyDeviceToHost));
        if (double(h_total) > want) hi_s = mid; else { lo_s = mid; scale = float(mid); }
      }
      split_refs_kernel<false><<<G, B>>>(in, grid, tri_lo, tri_hi, priority, scale, counts, nullptr, nullptr, nullptr);
      size_t scan_bytes = 0;
      B2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, counts, offsets, int(nt)));
      void* scan_tmp = tmp(scan_bytes);
      B2_CHECK(scan_tmp, B2PT_ERR_CUDA, "cudaMalloc failed for scan");
      B2_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, counts, offsets, int(nt)));
      uint32_t tail[2];
      B2_CUDA(cudaMemcpy(&tail[0], offsets + (nt - 1), 4, cudaMemcpyDeviceToHost));
      B2_CUDA(cudaMemcpy(&tail[1], counts + (nt - 1), 4, cudaMemcpyDeviceToHost));
      const uint32_t refs = tail[0] + tail[1];
      B2_CHECK(refs >= nt && refs < (1u << 28), B2PT_ERR_CAPACITY, "more than 2^28 triangle references after splitting");
      float4* ref_lo = (float4*)tmp(size_t(refs) * 16); float4* ref_hi = (float4*)tmp(size_t(refs) * 16);
      B2_CHECK(ref_lo && ref_hi, B2PT_ERR_CUDA, "cudaMalloc failed for triangle splitting");
      split_refs_kernel<true><<<G, B>>>(in, grid, tri_lo, tri_hi, priority, scale, counts, offsets, ref_lo, ref_hi);
      tri_lo = ref_lo; tri_hi = ref_hi;
      n = refs;
      G = (n + B - 1) / B;
    }
  }

  uint64_t* keys_in = (uint64_t*)tmp(size_t(n) * 8); uint64_t* keys = (uint64_t*)tmp(size_t(n) * 8);
  B2_CHECK(keys_in && keys, B2PT_ERR_CUDA, "cudaMalloc failed for BVH build");
  morton_kernel<<<G, B>>>(n, tri_lo, tri_hi, bounds, keys_in);
  size_t sort_bytes = 0;
  B2_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, sort_bytes, keys_in, keys, int(n)));
  void* sort_tmp = tmp(sort_bytes);
  B2_CHECK(sort_tmp, B2PT_ERR_CUDA, "cudaMalloc failed for sort");
  B2_CUDA(cub::DeviceRadixSort::SortKeys(sort_tmp, sort_bytes, keys_in, keys, int(n)));

  float4* nodes = nullptr;
  uint32_t num_wide = 1;
  const float pad = ext * 1e-6f;  // conservative slack on every child box (see trace.cuh)

  const char* bvh_env = getenv("B2PT_BVH");
#ifdef B2PT_EXPERIMENTS
  const bool wide8 = !(bvh_env && bvh_env[0] == '4');  // A/B builds: B2PT_BVH=4 keeps the 4-wide uncompressed tree
#else
  const bool wide8 = true; (void)bvh_env;
#endif
  float4* nodes8 = nullptr;
  uint64_t* tri_map = nullptr;
  if (n == 1 && !wide8) {
    // a single leaf under the root
    if (int rc = dev_alloc(s, &nodes, kNodeFloat4s, true)) return rc;
    float4 lo, hi;
    B2_CUDA(cudaMemcpy(&lo, tri_lo, 16, cudaMemcpyDeviceToHost));
    B2_CUDA(cudaMemcpy(&hi, tri_hi, 16, cudaMemcpyDeviceToHost));
    float4 h[kNodeFloat4s];
    const float inf = INFINITY;
    h[0] = make_float4(lo.x - pad, inf, inf, inf); h[1] = make_float4(lo.y - pad, inf, inf, inf); h[2] = make_float4(lo.z - pad, inf, inf, inf);
    h[3] = make_float4(hi.x + pad, -inf, -inf, -inf); h[4] = make_float4(hi.y + pad, -inf, -inf, -inf); h[5] = make_float4(hi.z + pad, -inf, -inf, -inf);
    uint32_t l0 = make_leaf(0, 1), inv = kInvalid;
    float f0, fi; memcpy(&f0, &l0, 4); memcpy(&fi, &inv, 4);
    h[6] = make_float4(f0, fi, fi, fi); h[7] = make_float4(0, 0, 0, 0);
    B2_CUDA(cudaMemcpy(nodes, h, sizeof h, cudaMemcpyHostToDevice));
  } else if (n > 1) {
    BinNode* bin = (BinNode*)tmp(size_t(n - 1) * sizeof(BinNode));
    uint32_t* parent_inner = (uint32_t*)tmp(size_t(n) * 4); uint32_t* parent_leaf = (uint32_t*)tmp(size_t(n) * 4);
    float4* box_lo = (float4*)tmp(size_t(n) * 16); float4* box_hi = (float4*)tmp(size_t(n) * 16);
    uint32_t* visit = (uint32_t*)tmp(size_t(n) * 4);
    CollapseItem* items_a = (CollapseItem*)tmp(size_t(n) * sizeof(CollapseItem));
    CollapseItem* items_b = (CollapseItem*)tmp(size_t(n) * sizeof(CollapseItem));
    uint32_t* counters = (uint32_t*)tmp(16);  // [0] num_wide, [1] next_count
    B2_CHECK(bin && parent_inner && parent_leaf && box_lo && box_hi && visit && items_a && items_b && counters, B2PT_ERR_CUDA,
             "cudaMalloc failed for BVH build");
    uint32_t root_bin = 0;
    int ploc_radius = kPlocRadiusDefault;
    { const char* e = getenv("B2PT_PLOC_RADIUS"); if (e) { const int v = atoi(e); if (v >= 1 && v <= 256) ploc_radius = v; } }
    const char* builder = getenv("B2PT_BUILDER");
    if (builder && !strcmp(builder, "lbvh")) {
      B2_CUDA(cudaMemset(visit, 0, size_t(n) * 4));
      radix_tree_kernel<<<G, B>>>(int(n), keys, bin, parent_inner, parent_leaf);
      fit_kernel<<<G, B>>>(int(n), keys, tri_lo, tri_hi, bin, parent_inner, parent_leaf, box_lo, box_hi, visit);
    } else {
      // PLOC over the Morton order (default): near-SAH quality, still entirely on the device
      const uint32_t total = 2u * n - 1u;
      uint32_t* cid_a = (uint32_t*)tmp(size_t(n) * 4); uint32_t* cid_b = (uint32_t*)tmp(size_t(n) * 4);
      float4* clo_a = (float4*)tmp(size_t(n) * 16); float4* chi_a = (float4*)tmp(size_t(n) * 16);
      float4* clo_b = (float4*)tmp(size_t(n) * 16); float4* chi_b = (float4*)tmp(size_t(n) * 16);
      float4* node_lo = (float4*)tmp(size_t(total) * 16); float4* node_hi = (float4*)tmp(size_t(total) * 16);
      uint32_t* size = (uint32_t*)tmp(size_t(total) * 4); uint32_t* parent = (uint32_t*)tmp(size_t(total) * 4);
      uint32_t* first = (uint32_t*)tmp(size_t(total) * 4);
      uint32_t* left = (uint32_t*)tmp(size_t(n) * 4); uint32_t* right = (uint32_t*)tmp(size_t(n) * 4);
      uint32_t* nn = visit;  // reuse: nearest-neighbour index per cluster
      uint32_t* valid = parent_inner; uint32_t* offset = parent_leaf;
      uint32_t* next_node = (uint32_t*)tmp(16);
      uint64_t* order = (uint64_t*)tmp(size_t(n) * 8);
      B2_CHECK(cid_a && cid_b && clo_a && chi_a && clo_b && chi_b && node_lo && node_hi && size && parent && first && left && right &&
                   next_node && order, B2PT_ERR_CUDA, "cudaMalloc failed for the PLOC build");
      size_t scan_bytes = 0;
      B2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, valid, offset, int(n)));
      void* scan_tmp = tmp(scan_bytes);
      B2_CHECK(scan_tmp, B2PT_ERR_CUDA, "cudaMalloc failed for scan");
      B2_CUDA(cudaMemcpy(next_node, &n, 4, cudaMemcpyHostToDevice));
      ploc_init_kernel<<<G, B>>>(n, keys, tri_lo, tri_hi, cid_a, clo_a, chi_a, node_lo, node_hi, size, parent);
      uint32_t m = n;
      int guard = 0;
      while (m > 1) {
        const uint32_t g = (m + B - 1) / B;
        ploc_nearest_kernel<<<g, B>>>(m, ploc_radius, clo_a, chi_a, nn);
        ploc_merge_kernel<<<g, B>>>(m, n, nn, cid_a, clo_a, chi_a, cid_b, clo_b, chi_b, valid, next_node, left, right, node_lo, node_hi,
                                    size, parent);
        B2_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, valid, offset, int(m)));
        ploc_compact_kernel<<<g, B>>>(m, valid, offset, cid_b, clo_b, chi_b, cid_a, clo_a, chi_a);
        uint32_t tail[2];
        B2_CUDA(cudaMemcpy(&tail[0], offset + (m - 1), 4, cudaMemcpyDeviceToHost));
        B2_CUDA(cudaMemcpy(&tail[1], valid + (m - 1), 4, cudaMemcpyDeviceToHost));
        const uint32_t m2 = tail[0] + tail[1];
        B2_CHECK(m2 < m && ++guard < 4096, B2PT_ERR_CUDA, "PLOC build made no progress");
        m = m2;
      }
      ploc_first_kernel<<<(total + B - 1) / B, B>>>(total, n, parent, left, size, first);
      ploc_finish_kernel<<<(total + B - 1) / B, B>>>(n, keys, left, right, size, first, node_lo, node_hi, bin, box_lo, box_hi, order);
      keys = order;  // triangles are now in depth-first leaf order of the PLOC tree
      root_bin = n - 2u;
    }
    // triangles per leaf: a subtree of at most leaf_max triangles stays closed (B2PT_LEAF_MAX = 1..3 for A/B runs)
    uint32_t leaf_max = 3u;
    { const char* e = getenv("B2PT_LEAF_MAX"); if (e && e[0] >= '1' && e[0] <= '3') leaf_max = uint32_t(e[0] - '0'); }
    if (wide8 && n > 3) {
      // compressed 8-wide tree (bvh8.cuh); a wide inner node has at least two children, so n nodes is a safe bound
      nodes8 = (float4*)tmp(size_t(n) * 5 * 16);  // built here, then copied at its real size into the BVH block below
      B2_CHECK(nodes8, B2PT_ERR_CUDA, "cudaMalloc failed for BVH build");
      tri_map = (uint64_t*)tmp(size_t(n) * 8);
      uint32_t* counters8 = (uint32_t*)tmp(16);  // [0] num_wide, [1] next_count, [2] triangles placed
      B2_CHECK(tri_map && counters8, B2PT_ERR_CUDA, "cudaMalloc failed for BVH build");
      CollapseItem root{0u, root_bin};
      B2_CUDA(cudaMemcpy(items_a, &root, sizeof root, cudaMemcpyHostToDevice));
      uint32_t h_counters[3] = {1u, 0u, 0u};
      B2_CUDA(cudaMemcpy(counters8, h_counters, 12, cudaMemcpyHostToDevice));
      uint32_t count = 1;
      while (count > 0) {
        collapse8_kernel<<<(count + B - 1) / B, B>>>(count, items_a, keys, tri_lo, tri_hi, bin, box_lo, box_hi, nodes8, &counters8[0], items_b,
                                                    &counters8[1], &counters8[2], tri_map, pad, 0u, leaf_max);
        B2_CUDA(cudaMemcpy(h_counters, counters8, 12, cudaMemcpyDeviceToHost));
        count = h_counters[1];
        num_wide = h_counters[0];
        B2_CUDA(cudaMemset(&counters8[1], 0, 4));
        std::swap(items_a, items_b);
      }
      B2_CHECK(h_counters[2] == n, B2PT_ERR_CUDA, "BVH8 collapse lost triangles");
      keys = tri_map;  // final triangle order: the leaves of one node are contiguous
    } else if (!wide8) {
      // worst case one wide node per binary inner node
      if (int rc = dev_alloc(s, &nodes, size_t(n) * kNodeFloat4s, true)) return rc;
      CollapseItem root{0u, root_bin};
      B2_CUDA(cudaMemcpy(items_a, &root, sizeof root, cudaMemcpyHostToDevice));
      uint32_t h_counters[2] = {1u, 0u};
      B2_CUDA(cudaMemcpy(counters, h_counters, 8, cudaMemcpyHostToDevice));
      uint32_t count = 1;
      while (count > 0) {
        collapse_kernel<<<(count + B - 1) / B, B>>>(count, items_a, keys, tri_lo, tri_hi, bin, box_lo, box_hi, nodes, &counters[0],
                                                   items_b, &counters[1], pad);
        B2_CUDA(cudaMemcpy(h_counters, counters, 8, cudaMemcpyDeviceToHost));
        count = h_counters[1];
        num_wide = h_counters[0];
        B2_CUDA(cudaMemset(&counters[1], 0, 4));
        std::swap(items_a, items_b);
      }
    }
  }
  if (wide8 && n <= 3) {
    // a root with one leaf child that holds every triangle
    nodes8 = (float4*)tmp(5 * 16);
    B2_CHECK(nodes8, B2PT_ERR_CUDA, "cudaMalloc failed for BVH build");
    tri_map = (uint64_t*)tmp(size_t(n) * 8);
    CollapseItem* item = (CollapseItem*)tmp(sizeof(CollapseItem));
    uint32_t* counters8 = (uint32_t*)tmp(16);
    B2_CHECK(tri_map && item && counters8, B2PT_ERR_CUDA, "cudaMalloc failed for BVH build");
    CollapseItem root{0u, 0u};
    B2_CUDA(cudaMemcpy(item, &root, sizeof root, cudaMemcpyHostToDevice));
    uint32_t h_counters[3] = {1u, 0u, 0u};
    B2_CUDA(cudaMemcpy(counters8, h_counters, 12, cudaMemcpyHostToDevice));
    collapse8_kernel<<<1, 32>>>(1u, item, keys, tri_lo, tri_hi, nullptr, nullptr, nullptr, nodes8, &counters8[0], item, &counters8[1],
                               &counters8[2], tri_map, pad, n, 3u);
    keys = tri_map;
    num_wide = 1;
  }

  // ---- the BVH block: [8-wide nodes | triangle records], ONE allocation so that a single access-policy window covers what every
  // ray reads again and again (b2pt_scene_l2_window below); the shading frames are read once per hit and stay outside
  float4 *tris = nullptr, *frames = nullptr, *bvh_block = nullptr;
  const size_t node_f4 = nodes8 ? size_t(num_wide) * 5 : 0;  // kNode8Float4s (bvh8.cuh)
  if (int rc = dev_alloc(s, &bvh_block, node_f4 + size_t(n) * kTriFloat4s, true)) return rc;
  if (nodes8) {
    B2_CUDA(cudaMemcpy(bvh_block, nodes8, node_f4 * 16, cudaMemcpyDeviceToDevice));
    nodes8 = bvh_block;
  }
  tris = bvh_block + node_f4;
  s->bvh_block = bvh_block; s->bvh_block_bytes = (node_f4 + size_t(n) * kTriFloat4s) * 16;
  if (int rc = dev_alloc(s, &frames, size_t(n) * kFrameFloat4s, true)) return rc;
  pack_kernel<<<G, B>>>(in, n, keys, tri_lo, tri_hi, tris, frames);
  B2_CUDA(cudaEventRecord(ev1));
  B2_CUDA(cudaEventSynchronize(ev1));
  B2_CUDA(cudaGetLastError());
  B2_CUDA(cudaEventElapsedTime(&s->build_ms, ev0, ev1));
  cudaEventDestroy(ev0); cudaEventDestroy(ev1);

  // ---- small tables
  B2ptMaterial* d_materials; B2ptLight* d_lights = nullptr;
  if (int rc = dev_alloc(s, &d_materials, d->num_materials ? d->num_materials : 1, true)) return rc;
  B2_CUDA(cudaMemcpy(d_materials, d->materials, d->num_materials * sizeof(B2ptMaterial), cudaMemcpyHostToDevice));
  const uint32_t L = d->num_lights;
  double *d_cp = nullptr, *d_den = nullptr; float* d_int = nullptr;
  if (L) {
    if (int rc = dev_alloc(s, &d_lights, L, true)) return rc;
    B2_CUDA(cudaMemcpy(d_lights, d->lights, L * sizeof(B2ptLight), cudaMemcpyHostToDevice));
    // std::piecewise_constant_distribution<float>(L, 0, 1, weights) as libstdc++ initialises it
    // (Sample.inl:235-247; bits/random.tcc param_type(nw, xmin, xmax, fw) + _M_initialize)
    std::vector<float> ints(L + 1);
    std::vector<double> den(L), cp(L);
    const float delta = (1.0f - 0.0f) / float(L);
    for (uint32_t k = 0; k <= L; ++k) ints[k] = 0.0f + float(k) * delta;
    for (uint32_t k = 0; k < L; ++k) den[k] = double(d->lights[k].weight);
    double sum = 0.0;
    for (uint32_t k = 0; k < L; ++k) sum += den[k];
    for (uint32_t k = 0; k < L; ++k) den[k] /= sum;
    double acc = 0.0;
    for (uint32_t k = 0; k < L; ++k) { acc += den[k]; cp[k] = acc; }
    cp[L - 1] = 1.0;
    for (uint32_t k = 0; k < L; ++k) den[k] /= double(ints[k + 1] - ints[k]);
    if (int rc = dev_alloc(s, &d_cp, L, true)) return rc;
    if (int rc = dev_alloc(s, &d_den, L, true)) return rc;
    if (int rc = dev_alloc(s, &d_int, L + 1, true)) return rc;
    B2_CUDA(cudaMemcpy(d_cp, cp.data(), L * 8, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(d_den, den.data(), L * 8, cudaMemcpyHostToDevice));
    B2_CUDA(cudaMemcpy(d_int, ints.data(), (L + 1) * 4, cudaMemcpyHostToDevice));
  }
  uint32_t* d_flags;
  if (int rc = dev_alloc(s, &d_flags, 4, true)) return rc;
  B2_CUDA(cudaMemset(d_flags, 0, 16));

  DeviceScene& ds = s->dev;
  ds.num_tris = n; ds.num_source_tris = nt; ds.num_nodes = num_wide; ds.num_materials = d->num_materials; ds.num_lights = L;
  ds.tris = tris; ds.tri_frames = frames; ds.nodes = nodes; ds.nodes8 = nodes8;
  ds.materials = d_materials; ds.lights = d_lights; ds.light_cp = d_cp; ds.light_den = d_den; ds.light_int = d_int;
  ds.error_flags = d_flags;
  { const char* e = getenv("B2PT_DYN_FETCH"); if (e) ds.dyn_fetch_lanes = atoi(e); }
  { const char* e = getenv("B2PT_DYN_TRI"); if (e) ds.dyn_tri_lanes = atoi(e); }
  { const char* e = getenv("B2PT_QUEUE_DRAIN"); if (e) ds.queue_drain_lanes = atoi(e); }
  { const char* e = getenv("B2PT_QUEUE_CHUNK"); if (e && atoi(e) >= 32) ds.queue_chunk = atoi(e); }
  { const char* e = getenv("B2PT_QUEUE_FETCH"); if (e && atoi(e) >= 1 && atoi(e) <= 32) ds.queue_fetch_lanes = atoi(e); }
  for (uint32_t k = 0; k < L; ++k) if (d->lights[k].diffuse == 0.0f) ds.has_directional = 1u;
  s->bvh_bytes = uint64_t(num_wide) * (nodes8 ? 5 : kNodeFloat4s) * 16 + uint64_t(n) * kTriFloat4s * 16;
  // ---- L2 residency (north_star: "top BVH levels staged in shared memory / L2"): set aside as much persisting L2 as the device
  // allows for the BVH block; render streams attach the window (scene_l2_window).  OFF by default — measured on B200: with
  // the window the closest-hit stage is unchanged (the BVH block is L2-resident anyway) and the shading stages, which lose the set-aside
  // part of the L2 for their pools, slow down by 9 %: 77.7 Msamples/s with it, 79.4 without.  B2PT_L2_PERSIST=1 turns it on.
  {
    const char* e = getenv("B2PT_L2_PERSIST");
    int dev = 0; cudaDeviceProp prop{};
    if ((e && e[0] == '1') && cudaGetDevice(&dev) == cudaSuccess && cudaGetDeviceProperties(&prop, dev) == cudaSuccess &&
        prop.persistingL2CacheMaxSize > 0 && prop.accessPolicyMaxWindowSize > 0) {
      const size_t want = std::min<size_t>(s->bvh_block_bytes, size_t(prop.persistingL2CacheMaxSize));
      if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
        s->l2_window_bytes = std::min<size_t>(s->bvh_block_bytes, size_t(prop.accessPolicyMaxWindowSize));
        s->l2_hit_ratio = float(std::min(1.0, double(want) / double(s->l2_window_bytes)));
      } else {
        cudaGetLastError();
      }
    }
  }
  return B2PT_OK;
}

// Attaches the scene's persisting-L2 window to a stream: loads of the BVH block (nodes + triangle records) are kept in the set-aside
// part of the L2 with probability hit_ratio, everything else the stream touches (path pools, queues) is streamed through the rest.
void scene_l2_window(const b2pt_scene_s* s, cudaStream_t stream) {
  if (!s->l2_window_bytes) return;
  cudaStreamAttrValue attr{};
  attr.accessPolicyWindow.base_ptr = s->bvh_block;
  attr.accessPolicyWindow.num_bytes = s->l2_window_bytes;
  attr.accessPolicyWindow.hitRatio = s->l2_hit_ratio;
  attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  if (cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &attr) != cudaSuccess) cudaGetLastError();
}

}  // namespace b2pt

extern "C" {

const char* b2pt_last_error(void) { return b2pt::get_error(); }
int b2pt_abi_version(void) { return B2PT_ABI_VERSION; }

int b2pt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int b2pt_set_device(int device) {
  B2_CHECK(device >= 0 && device < b2pt_device_count(), B2PT_ERR_INVALID, "b2pt_set_device: no such device");
  B2_CUDA(cudaSetDevice(device));
  return B2PT_OK;
}

int b2pt_device_synchronize(void) {
  B2_CUDA(cudaDeviceSynchronize());
  return B2PT_OK;
}

int b2pt_scene_create(const B2ptSceneDesc* desc, b2pt_scene_t* out) {
  B2_CHECK(desc && out, B2PT_ERR_INVALID, "b2pt_scene_create: null argument");
  B2_CHECK(desc->num_meshes > 0 && desc->mesh_material_id && desc->mesh_vertex_offset && desc->mesh_index_offset && desc->vertices &&
               desc->tangents && desc->indices && desc->materials,
           B2PT_ERR_INVALID, "b2pt_scene_create: incomplete scene description");
  B2_CHECK(b2pt_device_count() > 0, B2PT_ERR_NO_DEVICE, "no CUDA device: the B200 path has no CPU fallback");
  for (uint32_t m = 0; m < desc->num_meshes; ++m)
    B2_CHECK((desc->mesh_material_id[m] >> 2) < desc->num_materials, B2PT_ERR_INVALID, "mesh material index out of range");
  b2pt_scene_s* s = new b2pt_scene_s();
  int rc = b2pt::build_scene(desc, s);
  if (rc != B2PT_OK) { b2pt_scene_destroy(s); return rc; }
  *out = s;
  return B2PT_OK;
}

void b2pt_scene_destroy(b2pt_scene_t s) {
  if (!s) return;
  if (s->pipeline && s->pipeline_free) s->pipeline_free(s->pipeline);  // joins the worker before the pools go away
  if (s->workspace && s->workspace_free) s->workspace_free(s->workspace);
  for (int i = 0; i < s->num_allocations; ++i) cudaFree(s->allocations[i]);
  for (cudaEvent_t e : s->ev) if (e) cudaEventDestroy(e);
  delete s;
}

int b2pt_scene_info(b2pt_scene_t s, uint64_t* num_triangles, uint64_t* num_nodes, uint64_t* bvh_bytes, float* build_ms) {
  B2_CHECK(s, B2PT_ERR_INVALID, "b2pt_scene_info: null scene");
  if (num_triangles) *num_triangles = s->dev.num_tris;
  if (num_nodes) *num_nodes = s->dev.num_nodes;
  if (bvh_bytes) *bvh_bytes = s->bvh_bytes;
  if (build_ms) *build_ms = s->build_ms;
  return B2PT_OK;
}

}  // extern "C"
