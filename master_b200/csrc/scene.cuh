// scene.cuh — the scene as it lives in HBM.
//
// Replaces what haste::Scene + Embree hold on the host (Scene.hpp:28-76, Scene.cpp:20-73):
//   * triangle references in BVH-leaf order, one 48-byte record each (three 128-bit loads from two adjacent 32-byte sectors;
//     the up-to-three triangles of a leaf are contiguous):
//       (v0.xyz, bits geomID)   (v0-v1, bits primID)   (v2-v0, bits mask)
//     (a 64-byte record that also holds Embree's precomputed Ng = cross(e1, e2) measured neutral: 79.0 vs 79.5 Msamples/s, 33 % more bytes)
//     — Embree's Triangle4 convention e1 = v0-v1, e2 = v2-v0 (kernels/geometry/triangle.h:54), so the
//     Möller–Trumbore arithmetic can follow triangle_intersector_moeller.h:78-111 operation by operation.  A triangle the builder
//     split (scene.cu: split_refs_kernel) has one record per leaf that holds a part of it;
//   * the per-triangle shading record in the same order: the three vertex tangent frames
//     (Mesh::tangents, AreaLights.hpp:14-20) + material id, 7 float4 = 112 B, so querySurface
//     (Scene.cpp:80-126) is one contiguous read instead of index + 3 gathers;
//   * a 4-wide BVH, 128 B per node (8 float4), child boxes SoA so one node test is 6 float4 loads
//     + 1 for the child links;
//   * small tables: materials (64 B), lights (80 B), the light-pick CDF in double.
#pragma once
#include "common.cuh"

namespace b2pt {

constexpr uint32_t kInvalid = 0xFFFFFFFFu;
constexpr int kNodeFloat4s = 8;    // 128-byte BVH4 node
constexpr int kFrameFloat4s = 7;   // 112-byte shading record per triangle
constexpr int kTriFloat4s = 3;     // 48-byte triangle record
constexpr float kSplitBudgetDefault = 0.0f;  // extra triangle references the builder may create, as a fraction of the triangles (0 = no pre-splitting; measured in DESIGN.md §7)
constexpr int kMaxLeafTris = 4;

// Child link encoding in node float4 #6:
//   0xFFFFFFFF                      empty slot (box is +inf/-inf, never hit)
//   bit 31 set                      leaf: bits 0..27 first triangle (leaf order), bits 28..30 count-1
//   otherwise                       index of an inner node
__host__ __device__ __forceinline__ bool child_is_leaf(uint32_t c) { return (c & 0x80000000u) != 0; }
__host__ __device__ __forceinline__ uint32_t leaf_first(uint32_t c) { return c & 0x0FFFFFFFu; }
__host__ __device__ __forceinline__ uint32_t leaf_count(uint32_t c) { return ((c >> 28) & 7u) + 1u; }
__host__ __device__ __forceinline__ uint32_t make_leaf(uint32_t first, uint32_t count) {
  return 0x80000000u | ((count - 1u) << 28) | first;
}

struct DeviceScene {
  uint32_t num_tris = 0;         // triangle REFERENCES = records in `tris` / `tri_frames`
  uint32_t num_source_tris = 0;  // triangles of the scene
  uint32_t num_nodes = 0;
  uint32_t num_materials = 0;
  uint32_t num_lights = 0;
  uint32_t has_directional = 0;  // some light is a sun (AreaLight::diffuse == false): the integrators run the directional stage
  const float4* tris = nullptr;        // [num_tris * kTriFloat4s]
  const float4* tri_frames = nullptr;  // [num_tris * kFrameFloat4s]
  float bounds_lo[3] = {0.f, 0.f, 0.f}, bounds_inv_ext[3] = {1.f, 1.f, 1.f};  // scene box: the cells of the ray-sorting key (bpt.cu)
  const float4* nodes = nullptr;       // 4-wide tree: [num_nodes * kNodeFloat4s], root = node 0 (null when the 8-wide tree is built)
  const float4* nodes8 = nullptr;      // compressed 8-wide tree (bvh8.cuh): [num_nodes * 5], root = node 0
  const B2ptMaterial* materials = nullptr;
  const B2ptLight* lights = nullptr;
  // std::piecewise_constant_distribution<float> state (Sample.inl:233-247), in double as libstdc++ keeps it
  const double* light_cp = nullptr;    // cumulative probabilities [num_lights] (empty semantics when num_lights == 1)
  const double* light_den = nullptr;   // densities [num_lights]
  const float* light_int = nullptr;    // interval boundaries [num_lights + 1]
  uint32_t* error_flags = nullptr;     // bit 0: traversal stack overflow
  // scheduling thresholds of trace_dyn8 (trace.cuh); B2PT_DYN_FETCH / B2PT_DYN_TRI override them for A/B runs
  int dyn_fetch_lanes = 20;
  int dyn_tri_lanes = 8;
  int queue_fetch_lanes = 8;   // trace_queue8: idle lanes of a warp that trigger a refill (B2PT_QUEUE_FETCH)
  int queue_chunk = 128;       // trace_queue8: tickets a warp reserves per atomic (B2PT_QUEUE_CHUNK)
  uint32_t prmt_base = 0x47000000u;  // bits of 2^15: the constant half of bvh8.cuh's byte -> float permute, kept out of the compiler's sight on purpose
  int queue_drain_lanes = 12;  // trace_queue8: a partial chunk of the triangle ring is tested when fewer lanes than this advanced a node (B2PT_QUEUE_DRAIN)
};

}  // namespace b2pt

// host-side owner of the device arrays; defined in scene.cu
struct b2pt_scene_s {
  b2pt::DeviceScene dev;
  uint64_t bvh_bytes = 0;
  float build_ms = 0.f;
  void* allocations[24] = {nullptr};
  int num_allocations = 0;
  // per-scene render workspace (path pool, queues), created lazily by render.cu
  void* workspace = nullptr;
  void (*workspace_free)(void*) = nullptr;
  // b2pt_render_host's pipeline (host_pipeline.cu): per-pass films rendered ahead of the caller, staging buffers, worker threads
  void* pipeline = nullptr;
  void (*pipeline_free)(void*) = nullptr;
  void (*pipeline_quiesce)(b2pt_scene_s*) = nullptr;  // waits for passes still being rendered ahead
  cudaEvent_t ev[2] = {nullptr, nullptr};  // timing pair of b2pt_render
  // nodes + triangle records in one allocation, and the persisting-L2 window over it (scene.cu: scene_l2_window)
  void* bvh_block = nullptr;
  size_t bvh_block_bytes = 0, l2_window_bytes = 0;
  float l2_hit_ratio = 0.f;
};
namespace b2pt { void scene_l2_window(const b2pt_scene_s* s, cudaStream_t stream); }
