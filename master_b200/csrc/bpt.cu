// bpt.cu — wavefront bidirectional path tracer: BPTBase<Beta>::_traceEye / _traceLight / _connect* (BPT.cpp:14-324).
//
// The reference traces, per pixel-sample and on one thread, a light subpath into a 1024-slot stack array
// (BPT.hpp:29-30), then walks the eye subpath and connects every eye vertex to every stored light vertex.
// Here a BATCH of pixel-samples (whole passes, <= 2^21 paths) moves through four stages per iteration:
//
//   logic    (bpt_logic_kernel)    per path: a state machine over the reference's control flow.  Light phase:
//                                  sample the emitter, roulette, BSDF sample, consume the closest mesh hit, push /
//                                  overwrite / pop light vertices (BPT.cpp:122-192).  Eye phase: camera vertex,
//                                  consume the closest hit, MIS partial sums c / C, emission through lights
//                                  (_connect_light), roulette, next-event sample, BSDF sample (BPT.cpp:14-101).
//                                  Every new eye vertex EXPANDS into (eye vertex, light vertex) pair descriptors.
//   connect  (bpt_connect_kernel)  one thread per pair: two BSDF queries, the edge, the MIS weight Ap + Cp + 1
//                                  (BPT.cpp:194-223); camera pairs also project to the pixel and apply the camera
//                                  coefficient (BPT.cpp:294-319, Technique.cpp:246-305).  Non-zero contributions
//                                  become shadow-queue entries.
//   trace    (bpt_trace_kernel)    closest hit for every extension ray (mask: meshes only for light subpaths,
//                                  meshes + lights for eye subpaths).
//   shadow   (bpt_shadow_kernel)   any-hit for every queued connection; visible -> add to the path's radiance
//                                  (float atomics) or splat into the pass's light image (double atomics).
//
// Light vertices live in ONE pool in HBM (112 B records, bump-allocated with a warp-aggregated atomic, chained
// per path through a link array) instead of a 100 KB array per thread, so the stated cap is on the pool, not
// on the path: 12 vertices per path on average (the expectation with roulette 0.9 is <= 8.1) and
// B2ptRenderParams::max_subpath (default 1024, the reference's cap) per path.
// Every path owns one counter-based RNG stream; light phase and eye phase consume it in the reference's order.
#include <algorithm>
#include <memory>

#include <cub/cub.cuh>

#include "render_common.cuh"

#ifndef B2PT_LOGIC_MIN_BLOCKS
// resident CTAs per SM the shading stages are compiled for (register cap = 65536 / (256 * this)).  Measured with the two-ended live queue:
// 1: 72.6, 2: 75.3, 3: 73.4, 4: 71.6, 5: 68.0 Msamples/s — spills cost more than occupancy gives.  (Prefetching the next item's record and
// shading frame across grid strides was measured too: 72.8 vs 74.7 — with two batches in flight the other batch's kernels already fill the stalls.)
#define B2PT_LOGIC_MIN_BLOCKS 2
#endif
#ifndef B2PT_LOGIC_THREADS
#define B2PT_LOGIC_THREADS 256  // largest CTA the shading stages are compiled for; with B2PT_LOGIC_MIN_BLOCKS it sets the register cap
#endif
// The connect stage fits 96 registers with 36 bytes of spills: 128-thread CTAs, five per SM (20 warps instead of 16) measured
// 102.15 / 102.12 against 101.75 / 101.71 Msamples/s (256 x 2) and 101.74 (128 x 4, the same registers); the logic stage spills
// 418 bytes at that cap and stays at 256 x 2.
#ifndef B2PT_CONNECT_THREADS
#define B2PT_CONNECT_THREADS 128
#endif
#ifndef B2PT_CONNECT_MIN_BLOCKS
#define B2PT_CONNECT_MIN_BLOCKS 5
#endif
#ifndef B2PT_LOGIC_SPLIT_MIN_BLOCKS
#define B2PT_LOGIC_SPLIT_MIN_BLOCKS 2
#endif
#ifndef B2PT_TRACE_MIN_BLOCKS
#define B2PT_TRACE_MIN_BLOCKS 4  // traversal stages: 64 registers per thread
#endif

namespace b2pt {

namespace {

enum : uint32_t { kNew = 0, kLightWait = 1, kLightLoop = 2, kLightFinal = 3, kEyeStart = 4, kEyeWait = 5, kDone = 6 };
constexpr uint32_t kPairCamera = 0x80000000u;  // descriptor.x bit 31: the eye vertex is the camera (light-image splat)
constexpr uint32_t kNeeVertex = 0xFF000000u;   // descriptor.y >= this: the freshly sampled emitter point instead of a stored vertex; low 24 bits = light id
                                               // (so the connect stage asks for the light's record together with the vertices, not after the path record's word)
constexpr uint32_t kShadowSplat = 1u;
constexpr int kVertexFloat4s = 7;
constexpr int kRecFloat4s = 16;
enum { kRecThr = 7, kRecOmega = 8, kRecRayO = 9, kRecHit = 10, kRecRad = 11, kRecCtl = 12, kRecNee = 14, kRecLadj = 15 };
// Expanding a path's connections into pair descriptors by walking its vertex chain is a string of dependent loads
// (one L2 / HBM round trip per vertex); the first kLpTable vertex indices are therefore also kept by position, so the
// expansion reads them independently.  Longer paths (3 % at roulette 0.9) walk the chain for the rest.
constexpr uint32_t kLpTable = 32;

struct BptCounters {
  unsigned long long basic_rays, shadow_rays, numeric_errors, capacity_events;
  unsigned int done;
  unsigned int lv_count;
  unsigned int trace_count[2], pair_count[2], shadow_count[2];
  unsigned int light_count[2];  // the live queue has two ends: eye-phase paths from the front (trace_count), light-phase paths from the back
  unsigned int dir_count[2];
  unsigned int ticket[2];  // dynamic ray fetch of the traversal stages (trace, shadow), zeroed by the logic stage
  unsigned int oct_count[2][8];  // B2PT_SORT_RAYS=2: rays per direction octant, [iteration parity][octant]
};

// LightVertex / EyeVertex (BPT.hpp:13-27): w0 = a or c, w1 = A or C
struct Vertex {
  Surface s;
  float3 omega, thr;
  float w0, w1;
  uint32_t finite;
  float ladj;  // light_lobe_adjust of the surface if it is a point on an area emitter (computed once, when the vertex is made), else -1
};

struct BptPool {
  uint32_t size;         // slots in the batch (paths rounded up to the block size; the padding is born dead)
  uint32_t valid;        // paths in the batch = passes * pixels
  uint32_t lv_capacity;  // light vertex pool
  uint32_t q_capacity;   // pair / shadow queues
  // One 256-byte record per path slot (AoS): the stages reach slots through queues of live slots, i.e. as gathers, and a
  // record that is contiguous is fetched as whole 32-byte sectors that are used in full; the same words kept as SoA streams
  // cost 2.5x their size in HBM traffic once most neighbours of a live slot are dead (ncu, profiles/).
  //   float4 0..6  current vertex (pack_vertex)            7  pending BSDF sample: throughput.xyz, density_rev
  //   8  pending BSDF sample: omega.xyz, density           9  extension ray origin, w = bits(bsdf.finite | ray mask selector << 1)
  //   10 hit: t, u, v, bits(triangle) — only with ray sorting; otherwise hits live in hit_q, in queue order
  //   11 eye-path radiance (float atomics from the shadow and logic stages)
  //   12 bits(rng key, rng counter, light path head, light path count)
  //   14 next-event sample: position, bits(light id)       15 its emitter's lobe restriction (light_lobe_adjust), -, -, -
  // (14 and 15 share one 32-byte sector and are written together.)  The state of a live path is not stored: it waits for the hit of
  // its extension ray, and the end of the live queue it sits at says for which subpath.
  float4* rec;
  double* splat;     // light image of the pass the path belongs to: 3 doubles per slot (= per pixel of that pass)
  float4* lv;        // light vertex pool, AoS: 7 float4 per vertex
  uint32_t* lv_link; // previous vertex of the same path
  uint32_t* lp_idx;  // the first kLpTable vertices of every light path by position: lp_idx[slot * kLpTable + k], one 128-byte line per path
  float4* hit_q;         // closest hits in QUEUE order: hit_q[position in trace_q] (t, u, v, bits(triangle)), written by the closest-hit stage
  uint32_t* trace_q[2];  // live slots = the slots with an extension ray in flight; written by the logic stage of iteration i into [i & 1]
  // ray sorting (B2PT_SORT_RAYS): the logic stage writes one key per entry of the live queue and counts the keys; a scan and a scatter
  // then lay the queue out in key order in sort_q, each end on its own, for the TRAVERSAL stage only — the next logic stage keeps
  // walking trace_q, whose order follows the path pool (slots that are neighbours in the queue are neighbours in HBM)
  uint32_t* sort_q; uint32_t* key_q; uint32_t* hist; uint32_t* offs;
  // B2PT_SORT_RAYS=2: one queue per direction octant (8 x size entries), appended by the logic stage next to trace_q; the traversal
  // stage walks them one after the other — direction coherence without a sorting pass
  uint32_t* oct_q;
  uint2* pair_q;
  float4* shadow_q;  // 3 float4 per entry
  float4* dir_q;     // connections to sun lights, 4 float4 per entry (render_common.cuh: directional_kernel)
};

__device__ __forceinline__ void pack_vertex(const Vertex& v, float4* q) {
  q[0] = make_float4(v.s.pos.x, v.s.pos.y, v.s.pos.z, __uint_as_float(v.s.material_id));
  q[1] = make_float4(v.s.gnormal.x, v.s.gnormal.y, v.s.gnormal.z, __uint_as_float(v.finite));
  q[2] = make_float4(v.s.tan.c0.x, v.s.tan.c0.y, v.s.tan.c0.z, v.w0);
  q[3] = make_float4(v.s.tan.c1.x, v.s.tan.c1.y, v.s.tan.c1.z, v.w1);
  q[4] = make_float4(v.s.tan.c2.x, v.s.tan.c2.y, v.s.tan.c2.z, v.ladj);
  q[5] = make_float4(v.omega.x, v.omega.y, v.omega.z, 0.f);
  q[6] = make_float4(v.thr.x, v.thr.y, v.thr.z, 0.f);
}
__device__ __forceinline__ Vertex unpack_vertex(const float4* q) {
  Vertex v;
  v.s.pos = f3(q[0].x, q[0].y, q[0].z); v.s.material_id = __float_as_uint(q[0].w);
  v.s.gnormal = f3(q[1].x, q[1].y, q[1].z); v.finite = __float_as_uint(q[1].w);
  v.s.tan.c0 = f3(q[2].x, q[2].y, q[2].z); v.w0 = q[2].w;
  v.s.tan.c1 = f3(q[3].x, q[3].y, q[3].z); v.w1 = q[3].w;
  v.s.tan.c2 = f3(q[4].x, q[4].y, q[4].z); v.ladj = q[4].w;
  v.omega = f3(q[5].x, q[5].y, q[5].z);
  v.thr = f3(q[6].x, q[6].y, q[6].z);
  return v;
}
__device__ __forceinline__ void store_slot_vertex(const BptPool& p, uint32_t slot, const Vertex& v) {
  float4 q[kVertexFloat4s];
  pack_vertex(v, q);
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) p.rec[size_t(slot) * kRecFloat4s + k] = q[k];
}
__device__ __forceinline__ Vertex load_slot_vertex(const BptPool& p, uint32_t slot) {
  float4 q[kVertexFloat4s];
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) q[k] = p.rec[size_t(slot) * kRecFloat4s + k];
  return unpack_vertex(q);
}
__device__ __forceinline__ void store_pool_vertex(const BptPool& p, uint32_t idx, const Vertex& v) {
  float4 q[kVertexFloat4s];
  pack_vertex(v, q);
  float4* o = p.lv + size_t(idx) * kVertexFloat4s;
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) o[k] = q[k];
}
__device__ __forceinline__ Vertex load_pool_vertex(const BptPool& p, uint32_t idx) {
  float4 q[kVertexFloat4s];
  const float4* o = p.lv + size_t(idx) * kVertexFloat4s;
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) q[k] = o[k];
  return unpack_vertex(q);
}

// BPTBase::_sample_to_vertex (BPT.cpp:103-114)
__device__ __forceinline__ Vertex sample_to_vertex(const LightSample& sample, float roulette_inv, float beta, float ladj) {
  Vertex v;
  v.ladj = ladj;
  v.s = sample.surface;
  v.omega = v.s.normal();
  v.thr = sample.radiance / sample.combined_density() * roulette_inv;
  v.w0 = sample.directional ? 0.0f : 1.0f / beta_fn(sample.combined_density(), beta);
  v.w1 = 0.0f;
  v.finite = 1;
  return v;
}

// the LightSample AreaLights::sample would return for light `id` at position `pos` (AreaLights.cpp:121-140)
__device__ __forceinline__ LightSample light_sample_at(const DeviceScene& sc, uint32_t id, float3 pos) {
  const B2ptLight& l = sc.lights[id];
  LightSample r;
  r.surface.pos = pos;
  r.surface.tan.c0 = f3(l.tangent[0], l.tangent[1], l.tangent[2]);
  r.surface.tan.c1 = f3(l.tangent[3], l.tangent[4], l.tangent[5]);
  r.surface.tan.c2 = f3(l.tangent[6], l.tangent[7], l.tangent[8]);
  r.surface.gnormal = r.surface.tan.c1;
  r.surface.material_id = l.material_id;
  r.radiance = f3(l.exitance[0], l.exitance[1], l.exitance[2]) * kOneOverPi;
  r.area_density = 1.0f / (l.size[0] * l.size[1]);
  r.light_density = l.weight;
  r.directional = l.diffuse == 0.0f;
  return r;
}

// BPTBase::_connect(light, eye) without the visibility factor (BPT.cpp:194-223)
__device__ __forceinline__ float3 connect_vertices(const DeviceScene& sc, const Vertex& light, const Vertex& eye, float beta) {
  const float3 omega = normalize(eye.s.pos - light.s.pos);
  const BsdfQuery lb = bsdf_query(sc, light.s, light.omega, omega, light.ladj);
  const BsdfQuery eb = bsdf_query(sc, eye.s, -omega, eye.omega);
  const Edge edge(light.s.pos, light.s.normal(), eye.s.pos, eye.s.normal(), omega);
  const float Ap = (light.w1 * beta_fn(lb.density_rev, beta) + light.w0 * float(light.finite)) * beta_fn(edge.b_geometry * eb.density_rev, beta);
  const float Cp = (eye.w1 * beta_fn(eb.density, beta) + eye.w0 * float(eye.finite)) * beta_fn(edge.f_geometry * lb.density, beta);
  const float weight_inv = Ap + Cp + 1.0f;
  const float3 result = light.thr * lb.throughput * eye.thr * eb.throughput * edge.b_cos * edge.f_geometry;
  return l1norm(result) < kFltEpsilon ? f3(0.0f) : result / weight_inv;
}

// BPTBase::_connect_light (BPT.cpp:225-243)
__device__ __forceinline__ float3 connect_light(const DeviceScene& sc, const Vertex& eye, float beta) {
  const BsdfQuery b = bsdf_query_emitter(sc, eye.s, f3(0.0f), eye.omega);  // the caller found eye.s.is_light()
  if (l1norm(b.throughput) < kFltMin) return f3(0.0f);
  float3 lr; float ld;
  query_lsdf(sc, eye.s, eye.omega, lr, ld);
  const float Cp = (eye.w1 * beta_fn(b.density, beta) + eye.w0 * float(eye.finite)) * beta_fn(ld, beta);
  return lr * eye.thr / (Cp + 1.0f);
}

// Technique::_camera_coefficient (Technique.cpp:246-274)
__device__ __forceinline__ float camera_coefficient(float3 light_omega, float3 light_gnormal, float3 light_normal, float3 eye_omega,
                                                    float3 eye_normal) {
  const float normal_c = fabsf(dot(eye_omega, light_gnormal) * dot(light_omega, light_normal) /
                               (dot(eye_omega, light_normal) * dot(light_omega, light_gnormal)));
  const float c = fabsf(dot(eye_normal, eye_omega));
  return normal_c * (1.0f / cr_cubef(c));
}

// warp-aggregated reservation of `n` consecutive queue entries per lane (one atomic per warp)
__device__ __forceinline__ unsigned warp_reserve(unsigned* counter, unsigned n) {
  const unsigned lane = threadIdx.x & 31u;
  unsigned incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= unsigned(o)) incl += t;
  }
  const unsigned total = __shfl_sync(0xFFFFFFFFu, incl, 31);
  unsigned base = 0;
  if (lane == 31 && total > 0) base = atomicAdd(counter, total);
  base = __shfl_sync(0xFFFFFFFFu, base, 31);
  return base + incl - n;
}


// ---------------------------------------------------------------- ray sorting
// After two or three bounces the rays of neighbouring pixels start from nearby points but leave in every direction: a warp of the
// traversal stage then walks 32 different parts of the tree (11.7 of 32 lanes per instruction, profiles/r01_bpt_trace_full_v7).  Rays
// that start in the same cell of a 32^3 grid over the scene and travel into the same octant visit the same upper nodes in the same
// order (the 8-wide tree orders its children by octant), so the live queue is put in the order of
//   key = octant << 15 | morton(cell)                                              (18 bits, one counting sort per iteration)
// before the traversal stage reads it.  The paths do not change, only which lanes run beside which.
constexpr uint32_t kSortBins = 1u << 18;
__device__ __forceinline__ uint32_t spread5(uint32_t v) {  // 5 bits -> every third bit
  v = (v | (v << 8)) & 0x0000100Fu;
  v = (v | (v << 4)) & 0x000010C3u;
  v = (v | (v << 2)) & 0x00001249u;
  return v;
}
__device__ __forceinline__ uint32_t ray_sort_key(const DeviceScene& sc, float3 O, float3 D) {
  const uint32_t cx = uint32_t(fminf(fmaxf((O.x - sc.bounds_lo[0]) * sc.bounds_inv_ext[0] * 32.0f, 0.0f), 31.0f));
  const uint32_t cy = uint32_t(fminf(fmaxf((O.y - sc.bounds_lo[1]) * sc.bounds_inv_ext[1] * 32.0f, 0.0f), 31.0f));
  const uint32_t cz = uint32_t(fminf(fmaxf((O.z - sc.bounds_lo[2]) * sc.bounds_inv_ext[2] * 32.0f, 0.0f), 31.0f));
  const uint32_t oct = (D.x >= 0.0f ? 1u : 0u) | (D.y >= 0.0f ? 2u : 0u) | (D.z >= 0.0f ? 4u : 0u);
  return (oct << 15) | (spread5(cx) << 2) | (spread5(cy) << 1) | spread5(cz);
}

// trace_q[parity] (two-ended, pool order) -> sort_q (two-ended, each end in key order); offs = exclusive scan of the key counts,
// eye-end keys first.  Clears nothing: the host zeroes hist after the scan.
__global__ void __launch_bounds__(256) bpt_sort_scatter_kernel(BptPool pool, BptCounters* cnt, int parity) {
  const uint32_t n_eye = cnt->trace_count[parity], n_light = cnt->light_count[parity];
  const uint32_t* in = pool.trace_q[parity];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_eye + n_light; i += gridDim.x * blockDim.x) {
    const uint32_t idx = i < n_eye ? i : pool.size - 1u - (i - n_eye);
    const uint32_t key = pool.key_q[idx] + (i < n_eye ? 0u : kSortBins);
    const uint32_t pos = atomicAdd(&pool.offs[key], 1u);
    pool.sort_q[pos < n_eye ? pos : pool.size - 1u - (pos - n_eye)] = in[idx];
  }
}

// ---------------------------------------------------------------- staging of gathered records through shared memory (A/B: B2PT_STAGED=2)
// The shading stages reach their records through queues, i.e. as gathers: every item costs one or two dependent round trips to HBM
// before any arithmetic can start (56 % / 47 % of the warp-state samples of the logic / connect stage were long-scoreboard stalls,
// profiles/r02_bpt_logic_full_v4_summary.md), and asking for the next item's words early in registers spills (128 registers, measured
// slower).  cp.async (LDGSTS) moves 16-byte words from global to shared memory without a destination register: a thread asks for the
// NEXT item's words right after it has read the current item's out of its own staging column, and the copy lands while it computes.
// Every thread owns one column [word][threadIdx.x] (conflict-free 128-bit accesses) and reads only what it copied itself, so the only
// synchronisation is the thread's own cp.async.wait_all.
// MEASURED NEGATIVE (profiles/r02_ab_runs.md, s31 / s32): 100.3 -> 96.3 Msamples/s for the connect stage, 93.0 for the logic stage
// (double-buffered, 96 KB per CTA; that variant was removed again) — CTAs that hold this much shared memory cannot share an SM with
// the other lane's traversal CTAs, and that overlap is worth more than the round trips.  The connect stage keeps the switch.
__device__ __forceinline__ void cp_async16(float4* smem_dst, const float4* gmem_src) {
  const uint32_t d = uint32_t(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// The stage runs over the LIVE paths only: after every iteration a path is either done or waiting for the hit of its
// extension ray, so the trace queue the previous iteration wrote is exactly the list of live slots (`first` = iteration 0
// of a batch walks every slot instead).  A pool-order loop would leave a warp with as many busy lanes as the pool has
// live paths — 4 of 32 halfway through a batch (ncu, profiles/).
// kEnd: 0 = both ends of the live queue (and the first iteration of a batch), 1 = the eye end only, 2 = the light end only — the two
// halves of the stage as kernels of their own (B2PT_LOGIC_SPLIT): each holds one subpath's code and state.
//
// Dependent round trips per item: queue entry + hit (both by queue POSITION, coalesced: the closest-hit stage stores its hits in queue
// order, BptPool::hit_q) -> path record + shading record of the hit triangle, asked for together -> arithmetic.  (With the hit inside
// the path record the shading record was a third round trip: 13 % of the stage's warp-state samples.)
// kFirst: iteration 0 of a batch (every slot is born here; kEnd = 0) — a kernel of its own so that the steady-state kernel does not carry its code.
// kSorting: the launch may be asked to order the closest-hit queue (B2PT_SORT_RAYS); the default instantiation does not carry that code.
template <int kEnd, bool kFirst, bool kSorting = false>
__global__ void __launch_bounds__(B2PT_LOGIC_THREADS, kEnd == 0 ? B2PT_LOGIC_MIN_BLOCKS : B2PT_LOGIC_SPLIT_MIN_BLOCKS) bpt_logic_kernel(RenderConfig cfg, BptPool pool, BptCounters* cnt, int parity, uint32_t batch_pass0,
                                                                               int sort_arg) {
  constexpr bool first = kFirst;
  const int sort = kSorting ? sort_arg : 0;
  const DeviceScene& sc = cfg.scene;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // trace_count[parity ^ 1] is this launch's input; the closest-hit stage clears it afterwards
    cnt->pair_count[parity ^ 1] = 0;
    cnt->shadow_count[parity ^ 1] = 0;
    cnt->ticket[0] = 0; cnt->ticket[1] = 0;
    cnt->dir_count[parity ^ 1] = 0;
#pragma unroll
    for (int o = 0; o < 8; ++o) cnt->oct_count[parity ^ 1][o] = 0;
  }
  const float beta = cfg.params.beta;
  const float roulette = cfg.params.roulette;
  const float roulette_inv = 1.0f / roulette;
  const uint32_t max_subpath = cfg.params.max_subpath ? cfg.params.max_subpath : 1024u;
  // trip counts are rounded up to whole warps: every lane of a warp reaches the warp_* helpers below
  // the two ends of the live queue are walked one after the other, each rounded up to whole warps: a warp holds paths of one phase
  const uint32_t n_eye = first || kEnd == 2 ? 0u : cnt->trace_count[parity ^ 1];
  const uint32_t n_light = first ? pool.size : kEnd == 1 ? 0u : cnt->light_count[parity ^ 1];
  const uint32_t n_eye_round = (n_eye + 31u) & ~31u;
  const uint32_t n_round = n_eye_round + ((n_light + 31u) & ~31u);
  const uint32_t* live_q = pool.trace_q[parity ^ 1];
  uint32_t* next_q = pool.trace_q[parity];
  const uint32_t stride = gridDim.x * blockDim.x;
  const bool hit_by_position = sort == 0;  // with ray sorting the closest-hit stage walks another order and leaves its hits in the records
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    // entry i of the two-ended queue: eye end first, each end rounded up to whole warps
    const bool eye_end = i < n_eye_round;
    const uint32_t j = eye_end ? i : i - n_eye_round;
    const bool lane_on = j < (eye_end ? n_eye : n_light);
    const uint32_t qpos = eye_end ? j : pool.size - 1u - j;
    uint32_t slot = 0u;
    float4 hv = make_float4(0.f, 0.f, 0.f, __uint_as_float(kInvalid));
    if (lane_on) {
      if (first) {
        slot = j;
      } else {
        slot = live_q[qpos];
        if (hit_by_position) hv = pool.hit_q[qpos];
      }
    }
    float4* rec = pool.rec + size_t(slot) * kRecFloat4s;
    if (lane_on && !first && !hit_by_position) hv = rec[kRecHit];
    // the shading record of the hit triangle, asked for here, next to the path record below
    SurfaceRaw raw;
    const bool has_tri = lane_on && !first && __float_as_uint(hv.w) != kInvalid;
    if (has_tri) load_surface_early(sc, __float_as_uint(hv.w), raw);
    // first entry of the path's vertex table, asked for now: the pair expansion at the end finds the table's line in L1
    // (its loads used to start a round trip of their own after all the arithmetic: 5 % of the stage's warp-state samples)
    uint32_t lp_first = 0u;
    if (lane_on && !first) lp_first = pool.lp_idx[size_t(slot) * kLpTable];
    // iteration 0 of a batch never reads the records: every slot is born kNew.  Afterwards a live path waits for the hit of its
    // extension ray, and which end of the queue it sits at says for which subpath: the state is not stored.
    const uint32_t st0 = !lane_on ? uint32_t(kDone) : kEnd == 1 ? uint32_t(kEyeWait) : kEnd == 2 ? uint32_t(kLightWait)
                         : first ? uint32_t(kNew) : eye_end ? uint32_t(kEyeWait) : uint32_t(kLightWait);
    uint32_t st = st0;
    bool want_trace = false, finished = false;
    uint32_t emit_n = 0, emit_flags = 0;  // pair descriptors to write after the branches
    bool emit_nee = false, emit_skip_first = false;
    uint32_t nee_light = 0;
    uint32_t head = kInvalid, count = 0;
    uint32_t ray_mesh_only = 0;
    int sample_kind = 0;  // the BSDF sample this item needs, taken at ONE place below: 1 = light subpath goes on, 2 = its last vertex, 3 = eye subpath
    float3 O = f3(0.f), rad = f3(0.f);  // rad: what this iteration adds to the path's radiance (emission through a light, the sky)
    bool rad_add = false;
    Rng rng; rng.key = 0; rng.ctr = 0;
    Vertex cur;
    BsdfSample bs;
    bs.omega = f3(0.f); bs.throughput = f3(0.f); bs.density = 0.f; bs.density_rev = 0.f; bs.finite = 1;
    bool cur_dirty = false;

    if (st != kDone) {
      if (kFirst && st == kNew && slot >= pool.valid) {
        finished = true;  // padding slot
      } else if (kFirst && st == kNew) {
        // Technique::_for_each_ray (Technique.cpp:321-331): jitter, primary direction
        const uint32_t pixel = slot % cfg.pixels;
        const uint32_t pass = batch_pass0 + slot / cfg.pixels;
        const uint32_t sample = cfg.params.first_sample + pass * (cfg.params.sample_stride ? cfg.params.sample_stride : 1u);
        rng.key = stream_key(cfg.params.seed, kDomainPixel, pixel, sample);
        const float3 dir = shoot_primary(cfg, rng, pixel);
        (void)dir;  // recomputed from the stream at kEyeStart; the light phase comes first (BPT.cpp:23)
        rec[kRecRad] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (roulette < rng.sample()) {  // BPT.cpp:19-21
          finished = true;
        } else if (roulette < rng.sample()) {  // _traceLight, BPT.cpp:126-128: empty light path
          head = kInvalid; count = 0;
          st = kEyeStart;
        } else {
          const LightSample ls = sample_light(sc, rng);
          cur = sample_to_vertex(ls, roulette_inv, beta, light_lobe_adjust(sc, ls.surface));
          const unsigned idx = warp_append(&cnt->lv_count, true);
          if (idx >= pool.lv_capacity) {
            atomicAdd(&cnt->capacity_events, 1ull);
            head = kInvalid; count = 0;
            st = kEyeStart;
          } else {
            store_pool_vertex(pool, idx, cur);
            pool.lv_link[idx] = kInvalid;
            pool.lp_idx[size_t(slot) * kLpTable] = idx;
            lp_first = idx;
            head = idx; count = 1;
            cur_dirty = true;
            st = kLightLoop;
          }
        }
      } else {
        const float4 ctl = rec[kRecCtl];
        rng.key = __float_as_uint(ctl.x); rng.ctr = __float_as_uint(ctl.y);
        head = __float_as_uint(ctl.z); count = __float_as_uint(ctl.w);
        {
          float4 q[kVertexFloat4s];
#pragma unroll
          for (int k = 0; k < kVertexFloat4s; ++k) q[k] = rec[k];
          cur = unpack_vertex(q);
        }
        const float4 bo = rec[kRecOmega], bt = rec[kRecThr], ro = rec[kRecRayO];
        bs.omega = f3(bo.x, bo.y, bo.z); bs.density = bo.w;
        bs.throughput = f3(bt.x, bt.y, bt.z); bs.density_rev = bt.w;
        bs.finite = __float_as_uint(ro.w) & 1u;
        O = f3(ro.x, ro.y, ro.z);
      }

      // ---- the hit of the extension ray becomes the next vertex: the same arithmetic for both subpaths (BPT.cpp:139-157 and
      // BPT.cpp:45-66), written once — the light subpath carries the roulette factor in its throughput
      const uint32_t tri = __float_as_uint(hv.w);
      Vertex nv;
      nv.ladj = -1.0f;
      bool nv_ok = false;  // the hit is a vertex with a throughput that counts
      if (has_tri) {
        load_surface_rest(sc, tri, raw);
        nv.s = query_surface_raw(raw, hv.y, hv.z, hv.x, O, bs.omega);
        nv.omega = -bs.omega;
        const Edge edge(cur.s.pos, cur.s.normal(), nv.s.pos, nv.s.normal(), nv.omega);
        nv.thr = cur.thr * bs.throughput * edge.b_cos;
        if (kEnd == 2 || (kEnd == 0 && st == kLightWait)) nv.thr = nv.thr * roulette_inv;
        if (!(l1norm(nv.thr) < kFltEpsilon)) {
          nv_ok = true;
          nv.thr /= bs.density;
          cur.finite = min(cur.finite, bs.finite);
          nv.finite = bs.finite;
          nv.w0 = 1.0f / beta_fn(edge.f_geometry * bs.density, beta);
          nv.w1 = (cur.w1 * beta_fn(bs.density_rev, beta) + cur.w0 * float(cur.finite)) * beta_fn(edge.b_geometry, beta) * nv.w0;
        }
      }

      if (kEnd != 1 && st == kLightWait) {
        // ---- _traceLight loop body after intersectMesh (BPT.cpp:139-176)
        if (tri == kInvalid) {
          st = kLightFinal;
        } else {
          if (!nv_ok) {
            st = kLightFinal;  // path.pop_back(); break;  — prv is untouched
          } else {
            if (bs.finite == 0) {
              store_pool_vertex(pool, head, nv);  // path[prv] = path[itr]; path.pop_back();
              cur = nv; cur_dirty = true;
              st = kLightLoop;
            } else {
              // path[prv].finite was updated in place
              pool.lv[size_t(head) * kVertexFloat4s + 1].w = __uint_as_float(cur.finite);
              unsigned idx = kInvalid;
              if (count < max_subpath) idx = atomicAdd(&cnt->lv_count, 1u);
              if (idx >= pool.lv_capacity) {
                atomicAdd(&cnt->capacity_events, 1ull);  // truncated here: the stored prefix stays valid
                cur_dirty = true;
                st = kLightFinal;
              } else {
                store_pool_vertex(pool, idx, nv);
                pool.lv_link[idx] = head;
                if (count < kLpTable) pool.lp_idx[size_t(slot) * kLpTable + count] = idx;
                head = idx; ++count;
                cur = nv; cur_dirty = true;
                st = kLightLoop;
              }
            }
          }
        }
      }

      if (kEnd != 1 && st == kLightLoop) {
        if (roulette < rng.sample()) st = kLightFinal;  // while (!_russian_roulette(generator))
        else sample_kind = 1;
      }
      if (kEnd != 1 && st == kLightFinal) sample_kind = 2;  // BPT.cpp:184-191: one more BSDF sample at the last vertex

      if (kEnd != 2 && st0 == kEyeWait) {
        // ---- BPT.cpp:45-97: consume the closest hit of the extension ray
        const uint32_t tri = __float_as_uint(hv.w);
        if (tri == kInvalid) {
          if (cur.s.is_camera()) {  // sky_gradient(bsdf.omega) *= _roulette_inv  (Technique.cpp:87-89)
            const float3 h = f3(cfg.params.sky_horizon[0], cfg.params.sky_horizon[1], cfg.params.sky_horizon[2]);
            const float3 z = f3(cfg.params.sky_zenith[0], cfg.params.sky_zenith[1], cfg.params.sky_zenith[2]);
            rad = (h * (1 - bs.omega.z) + z * bs.omega.z) * roulette_inv;  // nothing was added to a camera ray's radiance before
            rad_add = true;
          }
          finished = true;
        } else {
          if (!nv_ok) {
            finished = true;
          } else {
            if (nv.s.is_light()) {
              rad = connect_light(sc, nv, beta);
              rad_add = true;
              // surface = _scene->intersect(surface, bsdf.omega): on through the light, eye[prv] stays
              O = offset_origin(nv.s.pos, nv.s.gnormal, bs.omega);
              cur_dirty = true;  // eye[prv].finite
              ray_mesh_only = 0;
              want_trace = true;
            } else {
              cur = nv; cur_dirty = true;  // std::swap(itr, prv)
              if (roulette < rng.sample()) {
                finished = true;
              } else {
                cur.thr = cur.thr * roulette_inv;
                // ---- _connect(context, eye, path) (BPT.cpp:272-291)
                if (!(roulette < rng.sample())) {
                  const LightSample ls = sample_light(sc, rng);
                  if (!ls.directional) {
                    emit_nee = true;
                    const uint32_t light_id = sc.materials[ls.surface.material_index()].light_id;
                    nee_light = light_id & 0x00FFFFFFu;
                    rec[kRecNee] = make_float4(ls.surface.pos.x, ls.surface.pos.y, ls.surface.pos.z, __uint_as_float(light_id));
                    // the emitter's lobe restriction at this point: computed here, where every lane of the warp samples a light, instead of
                    // in the connect stage, where one pair in six needs it and the other lanes of its warp would wait for it
                    rec[kRecLadj] = make_float4(light_lobe_adjust(sc, ls.surface), 0.f, 0.f, 0.f);
                  } else {
                    // _connect_directional (BPT.cpp:245-269): queued for the directional stage
                    const float3 n = ls.surface.normal();
                    const BsdfQuery eb = bsdf_query_outofline(sc, cur.s, -n, cur.omega);
                    const float cosine = fabsf(dot(n, cur.s.normal()));
                    const float3 result = ls.radiance / ls.light_density * roulette_inv * cur.thr * eb.throughput * cosine;
                    const float KC = cur.w1 * beta_fn(eb.density, beta) + cur.w0 * float(cur.finite);
                    const float3 org = offset_origin(cur.s.pos, cur.s.gnormal, -n);
                    const unsigned dq = atomicAdd(&cnt->dir_count[parity], 1u);
                    if (dq < pool.size) {
                      float4* e = pool.dir_q + size_t(dq) * 4;
                      e[0] = make_float4(org.x, org.y, org.z, __uint_as_float(slot));
                      e[1] = make_float4(cur.s.pos.x, cur.s.pos.y, cur.s.pos.z, cosine);
                      e[2] = make_float4(result.x, result.y, result.z, KC);
                      e[3] = make_float4(__uint_as_float(sc.materials[ls.surface.material_index()].light_id), 0.f, 0.f, 0.f);
                    } else {
                      atomicAdd(&cnt->capacity_events, 1ull);
                    }
                  }
                }
                emit_n = (count > 0 ? count - 1u : 0u) + (emit_nee ? 1u : 0u);  // light vertices 1 .. n-1 (+ the fresh sample)
                emit_skip_first = true;
                sample_kind = 3;
              }
            }
          }
        }
      }

      // ---- the BSDF sample of the iteration, for whichever branch asked for it.  A lane takes at most one per iteration, and taking
      // it HERE lets the lanes of a warp run Scene::sampleBSDF together: at the light end of the queue nearly every warp holds both
      // paths that go on (kind 1) and paths at their last vertex (kind 2), and two call sites ran the whole function twice per warp.
      if (sample_kind) {
        const BsdfSample smp = bsdf_sample(sc, rng, cur.s, cur.omega);
        if (kEnd != 1 && sample_kind == 2) {
          if (smp.finite == 0) {  // a delta vertex cannot be connected
            head = pool.lv_link[head];
            --count;
          }
          st = kEyeStart;
        } else {
          bs = smp;
          O = offset_origin(cur.s.pos, cur.s.gnormal, bs.omega);  // Scene::intersectMesh (Scene.cpp:204-208) / Scene::intersect
          ray_mesh_only = sample_kind == 1 ? 1u : 0u;
          want_trace = true;
          if (sample_kind == 1) st = kLightWait;
        }
      }

      if (kEnd != 1 && st == kEyeStart) {
        // ---- BPT.cpp:26-34: the camera vertex
        Rng first; first.key = rng.key; first.ctr = 0;  // the two jitter draws open the stream (Technique.cpp:322)
        const float3 dir = shoot_primary(cfg, first, slot % cfg.pixels);
        cur.s = cfg.camera;
        cur.ladj = -1.0f;
        cur.omega = -dir;
        cur.thr = f3(1.0f) * roulette_inv;
        cur.finite = 1; cur.w0 = 0.0f; cur.w1 = 0.0f;
        cur_dirty = true;
        emit_n = count; emit_flags = kPairCamera;  // _connect_eye: every light vertex, no next-event sample
        bs = camera_sample(cur.s, cur.omega);  // CameraBSDF::sample: no draws
        O = offset_origin(cur.s.pos, cur.s.gnormal, bs.omega);
        ray_mesh_only = 0;
        want_trace = true;
        st = kEyeWait;
      }

      // the path's radiance lives in its record only: the any-hit stage adds the visible connections with float atomics, this stage
      // adds its own terms the same way (the same IEEE additions a load, add and store would make; nothing else touches the word now)
      if (rad_add) {
        atomicAdd(&rec[kRecRad], make_float4(rad.x, rad.y, rad.z, 0.0f));
      }
      if (finished) {
        st = kDone;
      } else {
        if (cur_dirty) store_slot_vertex(pool, slot, cur);
        rec[kRecCtl] = make_float4(__uint_as_float(rng.key), __uint_as_float(rng.ctr), __uint_as_float(head), __uint_as_float(count));
        if (want_trace) {
          rec[kRecOmega] = make_float4(bs.omega.x, bs.omega.y, bs.omega.z, bs.density);
          rec[kRecThr] = make_float4(bs.throughput.x, bs.throughput.y, bs.throughput.z, bs.density_rev);
          rec[kRecRayO] = make_float4(O.x, O.y, O.z, __uint_as_float((bs.finite & 1u) | (ray_mesh_only << 1)));
        }
      }
    }

    // ---- converged: queue appends
    {
      const unsigned mask = __ballot_sync(0xFFFFFFFFu, finished);
      if (mask && (threadIdx.x & 31u) == 0) atomicAdd(&cnt->done, unsigned(__popc(mask)));
    }
    // The three reservations of the iteration — the eye end and the light end of the live queue, the pair queue — are made by lane 0 with
    // three atomics issued back to back and broadcast afterwards: one round trip to the L2 per warp and iteration instead of three in a
    // row (the shuffles that wait for them were 15 % of the stage's warp-state samples, profiles/r02_bpt_logic_full_summary.md).
    const bool to_eye_end = want_trace && st == kEyeWait;
    const unsigned lane_id = threadIdx.x & 31u;
    const unsigned m_eye = __ballot_sync(0xFFFFFFFFu, to_eye_end), m_light = __ballot_sync(0xFFFFFFFFu, want_trace && !to_eye_end);
    unsigned pair_incl = emit_n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(0xFFFFFFFFu, pair_incl, o);
      if (lane_id >= unsigned(o)) pair_incl += t;
    }
    const unsigned pair_total = __shfl_sync(0xFFFFFFFFu, pair_incl, 31);
    unsigned base_eye = 0, base_light = 0, base_pair = 0;
    if (lane_id == 0) {
      if (m_eye) base_eye = atomicAdd(&cnt->trace_count[parity], unsigned(__popc(m_eye)));
      if (m_light) base_light = atomicAdd(&cnt->light_count[parity], unsigned(__popc(m_light)));
      if (pair_total) base_pair = atomicAdd(&cnt->pair_count[parity], pair_total);
    }
    base_eye = __shfl_sync(0xFFFFFFFFu, base_eye, 0);
    base_light = __shfl_sync(0xFFFFFFFFu, base_light, 0);
    base_pair = __shfl_sync(0xFFFFFFFFu, base_pair, 0);
    const unsigned tq = base_eye + __popc(m_eye & ((1u << lane_id) - 1u));
    const unsigned lq = base_light + __popc(m_light & ((1u << lane_id) - 1u));
    if (want_trace) {
      const uint32_t at = to_eye_end ? tq : pool.size - 1u - lq;
      next_q[at] = slot;
      if (sort == 1) {
        const uint32_t key = ray_sort_key(sc, O, bs.omega);
        pool.key_q[at] = key;
        atomicAdd(&pool.hist[key + (to_eye_end ? 0u : kSortBins)], 1u);
      }
    }
    if (sort == 2) {  // every lane of the warp is here: one atomic per octant present in the warp
      const uint32_t oct = (bs.omega.x >= 0.0f ? 1u : 0u) | (bs.omega.y >= 0.0f ? 2u : 0u) | (bs.omega.z >= 0.0f ? 4u : 0u);
      const unsigned active = __ballot_sync(0xFFFFFFFFu, want_trace);
      if (want_trace) {
        const unsigned peers = __match_any_sync(active, oct);
        const unsigned lane = threadIdx.x & 31u;
        const int leader = __ffs(int(peers)) - 1;
        unsigned base = 0;
        if (lane == unsigned(leader)) base = atomicAdd(&cnt->oct_count[parity][oct], unsigned(__popc(peers)));
        base = __shfl_sync(peers, base, leader);
        pool.oct_q[size_t(oct) * pool.size + base + __popc(peers & ((1u << lane) - 1u))] = slot;
      }
    }
    unsigned pq = base_pair + pair_incl - emit_n;
    if (emit_n) {
      if (pq + emit_n > pool.q_capacity) {
        atomicAdd(&cnt->capacity_events, 1ull);
      } else {
        if (emit_nee) pool.pair_q[pq++] = make_uint2(slot, kNeeVertex | nee_light);
        // vertices kLpTable .. count-1 hang off the head of the chain; the rest come from the position table, four loads in flight
        uint32_t v = head;
        for (uint32_t k = count; k > kLpTable; --k) {
          pool.pair_q[pq++] = make_uint2(slot | emit_flags, v);
          v = pool.lv_link[v];
        }
        const uint32_t kt = count < kLpTable ? count : kLpTable;
        const uint4* table = reinterpret_cast<const uint4*>(pool.lp_idx + size_t(slot) * kLpTable);
        for (uint32_t k = 0; k < kt; k += 4u) {
          const uint4 a4 = table[k >> 2];
          const uint32_t a[4] = {k == 0u ? lp_first : a4.x, a4.y, a4.z, a4.w};
#pragma unroll
          for (uint32_t j = 0; j < 4u; ++j)
            if (k + j < kt && !(emit_skip_first && k + j == 0u)) pool.pair_q[pq++] = make_uint2(slot | emit_flags, a[j]);
        }
      }
    }
  }
}

// one (eye vertex, light vertex) pair: BPTBase::_connect / _connect_eye without the visibility test; fills the shadow-queue entry
struct PairOut {
  bool want;
  float3 a, d, c;
  uint32_t target, flags;
};
__device__ __forceinline__ void connect_pair(const RenderConfig& cfg, const BptPool& pool, uint32_t slot, bool camera, const Vertex& eye,
                                             const Vertex& light, float beta, PairOut& o) {
  const DeviceScene& sc = cfg.scene;
  o.want = false; o.a = f3(0.f); o.d = f3(0.f); o.c = f3(0.f); o.target = 0; o.flags = 0;
  // ONE evaluation of the connection for camera pairs and eye-vertex pairs alike (two call sites were two copies of both BSDF switches:
  // the stage spent 10 % of its warp-state samples waiting for instructions)
  bool inside = true;
  float cc = 1.0f;
  if (camera) {
    // Technique::_accumulate (Technique.cpp:276-305): which pixel the light vertex projects to
    const float3 omega = normalize(light.s.pos - eye.s.pos);
    mat3 w2v;
    const float* m = cfg.params.camera.world_to_view;
    w2v.c0 = f3(m[0], m[1], m[2]); w2v.c1 = f3(m[3], m[4], m[5]); w2v.c2 = f3(m[6], m[7], m[8]);
    const float3 vd = mul(w2v, omega);
    const float2 pp = pixel_position(vd, cfg.width_f, cfg.height_f, cfg.res_y_inv, cfg.params.camera.focal_length_y);
    inside = (0 <= pp.x) && (pp.x < cfg.width_f) && (0 <= pp.y) && (pp.y < cfg.height_f);
    if (inside) {
      const uint32_t pixel = uint32_t(int(pp.y)) * cfg.params.width + uint32_t(int(pp.x));
      o.target = (slot / cfg.pixels) * cfg.pixels + pixel;  // same pass, other pixel
      o.flags = kShadowSplat;
      cc = camera_coefficient(light.omega, light.s.gnormal, light.s.normal(), omega, eye.s.normal());
    }
  } else {
    o.target = slot;
  }
  if (inside) {
    o.c = connect_vertices(sc, light, eye, beta);
    if (camera) o.c = o.c * cfg.focal_factor_y * cc;
  }
  if (inside) {
    const float l1 = l1norm(o.c);
    if (!isfinite(l1)) {
      // the reference multiplies by occluded() first: 0 * inf = NaN, so a non-finite connection poisons the
      // sample whether or not it is visible, and _commit_images drops it (Technique.cpp:224-230)
      if (o.flags & kShadowSplat) {
        atomicAdd(pool.splat + size_t(o.target) * 3, double(NAN));
      } else {
        atomicAdd(&pool.rec[size_t(o.target) * kRecFloat4s + kRecRad].x, NAN);
      }
    } else if (l1 != 0.0f) {
      // Scene::occluded(eye.surface, light.surface) (Scene.cpp:151-171)
      const float3 direction = normalize(light.s.pos - eye.s.pos);
      o.a = eye.s.pos + (dot(eye.s.gnormal, direction) > 0.0f ? 1.0f : -1.0f) * eye.s.gnormal * 0.0001f;
      const float3 b = light.s.pos + (dot(light.s.gnormal, direction) < 0.0f ? 1.0f : -1.0f) * light.s.gnormal * 0.0001f;
      o.d = b - o.a;
      o.want = true;
    }
  }
}
__device__ __forceinline__ void connect_emit(const BptPool& pool, BptCounters* cnt, int parity, const PairOut& o) {
  const unsigned sq = warp_append(&cnt->shadow_count[parity], o.want);
  if (o.want) {
    if (sq < pool.q_capacity) {
      float4* e = pool.shadow_q + size_t(sq) * 3;
      e[0] = make_float4(o.a.x, o.a.y, o.a.z, __uint_as_float(o.target));
      e[1] = make_float4(o.d.x, o.d.y, o.d.z, __uint_as_float(o.flags));
      e[2] = make_float4(o.c.x, o.c.y, o.c.z, 0.f);
    } else {
      atomicAdd(&cnt->capacity_events, 1ull);
    }
  }
}

// kStaged: the two vertices of the NEXT pair of a thread are copied to shared memory while it evaluates the current one (above);
// the launch is then one wave of resident CTAs (148 x B2PT_LOGIC_MIN_BLOCKS) and needs kConnectStageBytes of dynamic shared memory
constexpr int kConnectStageWords = 2 * kVertexFloat4s;
constexpr size_t kConnectStageBytes = size_t(kConnectStageWords) * 256 * sizeof(float4);
template <bool kStaged>
__global__ void __launch_bounds__(kStaged ? 256 : B2PT_CONNECT_THREADS, kStaged ? 2 : B2PT_CONNECT_MIN_BLOCKS) bpt_connect_kernel(RenderConfig cfg, BptPool pool, BptCounters* cnt, int parity) {
  extern __shared__ float4 s_stage[];
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = min(cnt->pair_count[parity], pool.q_capacity);
  const float beta = cfg.params.beta;
  const float roulette_inv = 1.0f / cfg.params.roulette;
  const uint32_t stride = gridDim.x * blockDim.x;
  const uint32_t rounds = (n + stride - 1) / stride;
  float4* mine = s_stage + threadIdx.x;
  // copy the words of pair `desc` into this thread's column: eye vertex 0..6, then the light vertex — or, for the next-event sample,
  // the two words of the path record it is rebuilt from
  auto stage = [&](uint2 desc) {
    const float4* rec = pool.rec + size_t(desc.x & ~kPairCamera) * kRecFloat4s;
#pragma unroll
    for (int k = 0; k < kVertexFloat4s; ++k) cp_async16(mine + k * 256, rec + k);
    if (desc.y >= kNeeVertex) {
      cp_async16(mine + kVertexFloat4s * 256, rec + kRecNee);
      cp_async16(mine + (kVertexFloat4s + 1) * 256, rec + kRecLadj);
    } else {
      const float4* lv = pool.lv + size_t(desc.y) * kVertexFloat4s;
#pragma unroll
      for (int k = 0; k < kVertexFloat4s; ++k) cp_async16(mine + (kVertexFloat4s + k) * 256, lv + k);
    }
  };
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint2 desc = make_uint2(0u, 0u), desc_next = make_uint2(0u, 0u);
  if (kStaged) {
    if (i < n) { desc = pool.pair_q[i]; stage(desc); }
    cp_async_commit();
    if (i + stride < n && i + stride >= i) desc_next = pool.pair_q[i + stride];
  }
  for (uint32_t r = 0; r < rounds; ++r, i += stride) {
    PairOut o;
    o.want = false; o.a = f3(0.f); o.d = f3(0.f); o.c = f3(0.f); o.target = 0; o.flags = 0;
    if (kStaged) {
      cp_async_wait_all();
      const bool on = i < n;
      const uint2 dcur = desc;
      float4 q[kConnectStageWords];
      if (on) {
#pragma unroll
        for (int k = 0; k < kConnectStageWords; ++k) q[k] = mine[k * 256];
      }
      // the column is free again: ask for the next pair's words, and for the descriptor after that
      desc = desc_next;
      if (i + stride < n) stage(desc);
      cp_async_commit();
      desc_next = make_uint2(0u, 0u);
      if (r + 2 < rounds) { const uint32_t i2 = i + 2 * stride; if (i2 < n) desc_next = pool.pair_q[i2]; }
      if (on) {
        const uint32_t slot = dcur.x & ~kPairCamera;
        const Vertex eye = unpack_vertex(q);
        Vertex light;
        if (dcur.y >= kNeeVertex) {
          const float4 nr = q[kVertexFloat4s];
          light = sample_to_vertex(light_sample_at(sc, dcur.y & 0x00FFFFFFu, f3(nr.x, nr.y, nr.z)), roulette_inv, beta, q[kVertexFloat4s + 1].x);
        } else {
          light = unpack_vertex(q + kVertexFloat4s);
        }
        connect_pair(cfg, pool, slot, (dcur.x & kPairCamera) != 0u, eye, light, beta, o);
      }
    } else if (i < n) {
      const uint2 d = pool.pair_q[i];
      const uint32_t slot = d.x & ~kPairCamera;
      const Vertex eye = load_slot_vertex(pool, slot);
      Vertex light;
      if (d.y >= kNeeVertex) {
        const float4 nr = pool.rec[size_t(slot) * kRecFloat4s + kRecNee];
        light = sample_to_vertex(light_sample_at(sc, d.y & 0x00FFFFFFu, f3(nr.x, nr.y, nr.z)), roulette_inv, beta,
                                 pool.rec[size_t(slot) * kRecFloat4s + kRecLadj].x);
      } else {
        light = load_pool_vertex(pool, d.y);
      }
      connect_pair(cfg, pool, slot, (d.x & kPairCamera) != 0u, eye, light, beta, o);
    }
    connect_emit(pool, cnt, parity, o);
  }
}

// ray sources / hit sinks of the two traversal stages (trace.cuh: trace_stream8)
struct BptTraceIo {
  const BptPool* pool;
  const uint32_t* q;
  uint32_t n_eye;  // entries [0, n_eye) from the front of the queue, the rest from its back
  const unsigned* oct_count;  // non-null: q is the array of eight octant queues and ray i is found through the prefix of the counts
  bool by_position;  // the hit goes to hit_q[queue position] (the logic stage reads it with the queue entry); else into the path record
  __device__ __forceinline__ uint32_t load(uint32_t i, float3& O, float3& D, float& tnear, float& tfar, uint32_t& mask) const {
    uint32_t slot, pos = 0u;
    if (oct_count) {
      uint32_t o = 0, j = i;
#pragma unroll
      for (int k = 0; k < 7; ++k) { const uint32_t c = oct_count[k]; if (o == uint32_t(k) && j >= c) { j -= c; ++o; } }
      slot = q[size_t(o) * pool->size + j];
    } else {
      pos = i < n_eye ? i : pool->size - 1u - (i - n_eye);
      slot = q[pos];
    }
    const float4* rec = pool->rec + size_t(slot) * kRecFloat4s;
    const float4 o = rec[kRecRayO], d = rec[kRecOmega];
    O = f3(o.x, o.y, o.z); D = f3(d.x, d.y, d.z); tnear = 0.0f; tfar = INFINITY;
    mask = (__float_as_uint(o.w) & 2u) ? kMaskMesh : kMaskAll;  // Scene::intersectMesh (light subpaths) / Scene::intersect
    return by_position ? pos : slot;
  }
  __device__ __forceinline__ void store(uint32_t tag, const Hit& h) const {
    const float4 v = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri));
    if (by_position) pool->hit_q[tag] = v;
    else pool->rec[size_t(tag) * kRecFloat4s + kRecHit] = v;
  }
};
struct BptShadowIo {
  const BptPool* pool;
  __device__ __forceinline__ uint32_t load(uint32_t i, float3& O, float3& D, float& tnear, float& tfar, uint32_t& mask) const {
    const float4* e = pool->shadow_q + size_t(i) * 3;
    const float4 o = e[0], d = e[1];
    O = f3(o.x, o.y, o.z); D = f3(d.x, d.y, d.z); tnear = 0.0f; tfar = 1.0f;  // Scene::occluded: t in (0, 1), meshes only
    mask = kMaskMesh;
    return i;
  }
  __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
    if (h.tri != kInvalid) return;  // blocked
    const float4* e = pool->shadow_q + size_t(i) * 3;
    const uint32_t target = __float_as_uint(e[0].w), flags = __float_as_uint(e[1].w);
    const float4 c = e[2];
    if (flags & kShadowSplat) {
      double* s = pool->splat + size_t(target) * 3;
      atomicAdd(s + 0, double(c.x)); atomicAdd(s + 1, double(c.y)); atomicAdd(s + 2, double(c.z));
    } else {
      // ONE vector reduction (red.global.add.v4.f32, sm_90+) instead of three scalar ones: the same three IEEE additions, a third of the
      // atomic operations at the L2 (w of the radiance word is 0 and stays 0)
      atomicAdd(&pool->rec[size_t(target) * kRecFloat4s + kRecRad], make_float4(c.x, c.y, c.z, 0.0f));
    }
  }
};

template <int kStream>
__global__ void __launch_bounds__(256, B2PT_TRACE_MIN_BLOCKS) bpt_trace_kernel(RenderConfig cfg, BptPool pool, BptCounters* cnt, int parity, int sorted) {
  const uint32_t n_eye = cnt->trace_count[parity];
  const uint32_t n = n_eye + cnt->light_count[parity];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cnt->basic_rays, (unsigned long long)n);
    cnt->trace_count[parity ^ 1] = 0;  // consumed by this iteration's logic stage; the next one appends to it
    cnt->light_count[parity ^ 1] = 0;
  }
  BptTraceIo io{&pool, sorted == 2 ? pool.oct_q : sorted == 1 ? pool.sort_q : pool.trace_q[parity], n_eye, sorted == 2 ? cnt->oct_count[parity] : nullptr, sorted == 0};
  if (kStream == 3) { trace_queue8<false>(cfg.scene, n, &cnt->ticket[0], io); return; }
  if (kStream == 2) { trace_dyn8<false>(cfg.scene, n, &cnt->ticket[0], io); return; }
  if (kStream == 1) { trace_stream8<false>(cfg.scene, n, &cnt->ticket[0], io); return; }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float3 O, D; float tnear, tfar; uint32_t mask;
    const uint32_t tag = io.load(i, O, D, tnear, tfar, mask);
    io.store(tag, trace_any_bvh<false>(cfg.scene, O, D, tnear, tfar, mask));
  }
}

template <int kStream>
__global__ void __launch_bounds__(256, B2PT_TRACE_MIN_BLOCKS) bpt_shadow_kernel(RenderConfig cfg, BptPool pool, BptCounters* cnt, int parity) {
  const uint32_t n = min(cnt->shadow_count[parity], pool.q_capacity);
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->shadow_rays, (unsigned long long)n);
  BptShadowIo io{&pool};
  if (kStream == 3) { trace_queue8<true>(cfg.scene, n, &cnt->ticket[1], io); return; }
  if (kStream == 2) { trace_dyn8<true>(cfg.scene, n, &cnt->ticket[1], io); return; }
  if (kStream == 1) { trace_stream8<true>(cfg.scene, n, &cnt->ticket[1], io); return; }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float3 O, D; float tnear, tfar; uint32_t mask;
    const uint32_t tag = io.load(i, O, D, tnear, tfar, mask);
    io.store(tag, trace_any_bvh<true>(cfg.scene, O, D, tnear, tfar, mask));
  }
}

// Technique::_commit_images (Technique.cpp:194-244) for the passes of one batch: per pixel and pass,
// view[p] += (light[p] + eye[p], 1) unless non-finite; the helpers are zeroed for the next batch.
__global__ void __launch_bounds__(256) bpt_commit_kernel(RenderConfig cfg, BptPool pool, BptCounters* cnt, uint32_t batch_pass0,
                                                         uint32_t batch_passes) {
  for (uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < cfg.pixels; pixel += gridDim.x * blockDim.x) {
    double sum[4] = {0.0, 0.0, 0.0, 0.0};
    for (uint32_t k = 0; k < batch_passes; ++k) {
      const uint32_t slot = k * cfg.pixels + pixel;
      const float4 e = pool.rec[size_t(slot) * kRecFloat4s + kRecRad];
      double* s = pool.splat + size_t(slot) * 3;
      const double r = s[0] + double(e.x), g = s[1] + double(e.y), b = s[2] + double(e.z);
      s[0] = 0.0; s[1] = 0.0; s[2] = 0.0;
      if (cfg.path_radiance) {
        float* o = cfg.path_radiance + (size_t(batch_pass0 + k) * cfg.pixels + pixel) * 3;
        o[0] = e.x; o[1] = e.y; o[2] = e.z;
      }
      const bool ok = isfinite(fabs(r) + fabs(g) + fabs(b));
      if (!ok) atomicAdd(&cnt->numeric_errors, 1ull);
      if (cfg.accum_pass_stride) {  // per-pass films (host pipeline): this pass owns its film
        if (ok) {
          double* a = cfg.accum + size_t(batch_pass0 + k) * cfg.accum_pass_stride + size_t(pixel) * 4;
          a[0] += r; a[1] += g; a[2] += b; a[3] += 1.0;
        }
      } else if (ok) {
        sum[0] += r; sum[1] += g; sum[2] += b; sum[3] += 1.0;
      }
    }
    if (cfg.accum_pass_stride) continue;
    // two batches (lanes) may commit at the same time: red.global.add.f64
    double* a = cfg.accum + size_t(pixel) * 4;
    atomicAdd(a + 0, sum[0]); atomicAdd(a + 1, sum[1]); atomicAdd(a + 2, sum[2]); atomicAdd(a + 3, sum[3]);
  }
}

// Two batches are kept in flight on two streams ("lanes"): the late iterations of a batch have few live paths,
// so its kernels are small and end in long single-ray tails; the other lane's kernels fill the machine meanwhile.
constexpr int kMaxLanes = 8;
constexpr int kDefaultLanes = 2;
constexpr int kSortRaysDefault = 0;
// ray ordering of the closest-hit queue: 0 = pool order, 1 = counting sort by (octant, cell) (ray_sort_key), 2 = one queue per octant
inline int bpt_sort_mode() {
  const char* e = getenv("B2PT_SORT_RAYS");
  return e ? (e[0] == '1' ? 1 : e[0] == '2' ? 2 : 0) : kSortRaysDefault;
}
inline int bpt_lanes() {  // B2PT_BPT_LANES overrides the number of concurrent batches for A/B runs
  const char* e = getenv("B2PT_BPT_LANES");
  const int v = e ? atoi(e) : kDefaultLanes;
  return v < 1 ? 1 : v > kMaxLanes ? kMaxLanes : v;
}

struct BptLane {
  BptPool pool{};
  BptCounters* counters = nullptr;
  BptCounters* host_counters = nullptr;  // pinned
  cudaStream_t stream = nullptr;         // lane 0 runs on the caller's stream, lane 1 on `own_stream`
  cudaStream_t own_stream = nullptr;
  cudaEvent_t done = nullptr;
};

struct BptWorkspace {
  BptLane lane[kMaxLanes];
  uint32_t size = 0, per_path = 0;
  char* scan_tmp = nullptr; size_t scan_bytes = 0;  // cub scratch of the ray sort, one slice per lane
  cudaEvent_t fork = nullptr;
  std::vector<void*> blocks;
};

void bpt_workspace_free(void* p) {
  BptWorkspace* w = (BptWorkspace*)p;
  for (void* q : w->blocks) cudaFree(q);
  for (auto& l : w->lane) {
    if (l.host_counters) cudaFreeHost(l.host_counters);
    if (l.own_stream) cudaStreamDestroy(l.own_stream);
    if (l.done) cudaEventDestroy(l.done);
  }
  if (w->fork) cudaEventDestroy(w->fork);
  delete w;
}

template <class T>
int ws_alloc(BptWorkspace* w, T** p, size_t count) {
  void* q = nullptr;
  B2_CUDA(cudaMalloc(&q, count * sizeof(T)));
  w->blocks.push_back(q);
  *p = (T*)q;
  return B2PT_OK;
}

int bpt_workspace(b2pt_scene_s* scene, uint32_t size, uint32_t per_path, int lanes, BptWorkspace** out) {
  BptWorkspace* w = (BptWorkspace*)scene->workspace;
  // a workspace that is large enough is reused (b2pt_render_host ramps its batch sizes; start_batch sets the size a batch really uses)
  if (w && scene->workspace_free == bpt_workspace_free && w->size >= size && w->per_path == per_path && w->lane[lanes - 1].counters) { *out = w; return B2PT_OK; }
  if (scene->workspace && scene->workspace_free) scene->workspace_free(scene->workspace);
  scene->workspace = nullptr;
  w = new BptWorkspace();
  scene->workspace = w; scene->workspace_free = bpt_workspace_free;
  w->size = size; w->per_path = per_path;
  B2_CUDA(cudaEventCreateWithFlags(&w->fork, cudaEventDisableTiming));
  for (int k = 0; k < lanes; ++k) {
    BptLane& l = w->lane[k];
    BptPool& p = l.pool;
    p.size = size;
    p.lv_capacity = uint32_t(std::min<uint64_t>(std::max<uint64_t>(uint64_t(size) * per_path, 4096u), 0x7FFFFF00u));
    p.q_capacity = p.lv_capacity + size;
    int rc;
    if ((rc = ws_alloc(w, &p.rec, size_t(size) * kRecFloat4s))) return rc;
    if ((rc = ws_alloc(w, &p.splat, size_t(size) * 3))) return rc;
    if ((rc = ws_alloc(w, &p.lv, size_t(p.lv_capacity) * kVertexFloat4s))) return rc;
    if ((rc = ws_alloc(w, &p.lv_link, p.lv_capacity))) return rc;
    if ((rc = ws_alloc(w, &p.lp_idx, size_t(size) * kLpTable))) return rc;
    if ((rc = ws_alloc(w, &p.hit_q, size))) return rc;
    if ((rc = ws_alloc(w, &p.trace_q[0], size))) return rc;
    if ((rc = ws_alloc(w, &p.trace_q[1], size))) return rc;
    if (bpt_sort_mode() == 2) { if ((rc = ws_alloc(w, &p.oct_q, size_t(size) * 8))) return rc; }
    if ((rc = ws_alloc(w, &p.sort_q, size))) return rc;
    if ((rc = ws_alloc(w, &p.key_q, size))) return rc;
    if ((rc = ws_alloc(w, &p.hist, 2 * kSortBins))) return rc;
    if ((rc = ws_alloc(w, &p.offs, 2 * kSortBins))) return rc;
    B2_CUDA(cudaMemset(p.hist, 0, size_t(2 * kSortBins) * 4));
    if (!w->scan_tmp) {
      B2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, w->scan_bytes, p.hist, p.offs, int(2 * kSortBins)));
      w->scan_bytes = (w->scan_bytes + 255) & ~size_t(255);
      char* t = nullptr;
      if ((rc = ws_alloc(w, &t, w->scan_bytes * kMaxLanes))) return rc;
      w->scan_tmp = t;
    }
    if ((rc = ws_alloc(w, &p.pair_q, p.q_capacity))) return rc;
    if ((rc = ws_alloc(w, &p.shadow_q, size_t(p.q_capacity) * 3))) return rc;
    if ((rc = ws_alloc(w, &p.dir_q, size_t(size) * 4))) return rc;
    if ((rc = ws_alloc(w, &l.counters, 1))) return rc;
    B2_CUDA(cudaMallocHost((void**)&l.host_counters, sizeof(BptCounters)));
    B2_CUDA(cudaMemset(p.splat, 0, size_t(size) * 3 * sizeof(double)));
    B2_CUDA(cudaEventCreateWithFlags(&l.done, cudaEventDisableTiming));
    if (k > 0) { B2_CUDA(cudaStreamCreateWithFlags(&l.own_stream, cudaStreamNonBlocking)); scene_l2_window(scene, l.own_stream); }
  }
  *out = w;
  return B2PT_OK;
}

}  // namespace

int render_bpt(b2pt_scene_s* scene, const RenderConfig& base, cudaStream_t stream, B2ptStats* stats) {
  const int block = 256;
  const uint32_t pixels = base.pixels;
  const uint32_t num_passes = base.params.num_samples;
  // a batch is a whole number of passes, at most 2^23 paths (B2PT_BPT_BATCH_LOG2 overrides the exponent; one pass if the window is larger):
  // measured on the LivingRoom stand-in, 16 passes of 1024 x 1024 run at 46 / 53 / 60 Msamples/s with 2^21 / 2^22 / 2^23 paths per batch
  // (the late iterations of a batch have few live paths; the more paths a batch starts with, the longer its launches stay large).
  // With more than one pass to render the passes are spread over bpt_lanes() concurrent batches.
  uint32_t batch_paths = 1u << 23;
  { const char* e = getenv("B2PT_BPT_BATCH_LOG2"); if (e) { int v = atoi(e); if (v >= 10 && v <= 26) batch_paths = 1u << v; } }
  // the light-vertex pool grows with 1 / (1 - roulette) (render_common.cuh: lv_per_path): keep it at what 2^23 paths x 12 vertices take
  const uint32_t per_path = lv_per_path(base.params.roulette);
  while (batch_paths > pixels && uint64_t(batch_paths) * per_path > (uint64_t(12) << 23)) batch_paths >>= 1;
  uint32_t passes_per_batch = batch_paths / pixels;
  if (passes_per_batch < 1) passes_per_batch = 1;
  B2_CHECK(base.scene.num_lights < (1u << 24), B2PT_ERR_INVALID, "render_bpt: more than 2^24 lights (the pair descriptor carries the light id in 24 bits)");
  const int kLanes = bpt_lanes();
  if (passes_per_batch > (num_passes + kLanes - 1) / kLanes) passes_per_batch = (num_passes + kLanes - 1) / kLanes;
  const uint32_t num_batches = (num_passes + passes_per_batch - 1) / passes_per_batch;
  const int lanes = int(num_batches) < kLanes ? int(num_batches) : kLanes;
  const uint32_t size = (passes_per_batch * pixels + block - 1) / block * block;
  BptWorkspace* w = nullptr;
  if (int rc = bpt_workspace(scene, size, per_path, lanes, &w)) return rc;
  RenderConfig cfg = base;
  cfg.pool_size = size;
  const int grid = grid_for(size, block);
  int grid_t = persistent_grid(size, block);  // traversal stages: persistent warps fetch rays from a ticket counter
  if (B2PT_TRACE_MIN_BLOCKS != 4) grid_t = int(std::min((size_t(size) + block - 1) / block, size_t(kNumSMs) * B2PT_TRACE_MIN_BLOCKS));  // A/B builds: one wave at another register cap
  // Co-residency experiment (A/B): B2PT_TRACE_CTAS=n caps the persistent traversal grids at n CTAs per SM (4 fill the register file),
  // B2PT_SHADE_BLOCK=128 launches the shading stages with 128-thread CTAs, so that a shading CTA of the OTHER lane fits beside the
  // traversal CTAs and the latency-bound shading overlaps the issue-bound traversal on the same SM.
  int shade_block = block, shade_grid = grid;
  { const char* e = getenv("B2PT_TRACE_CTAS"); if (e) { const int v = atoi(e); if (v >= 1 && v <= 8) grid_t = std::min(grid_t, kNumSMs * v); } }
  { const char* e = getenv("B2PT_SHADE_BLOCK"); if (e) { const int v = atoi(e); if (v == 64 || v == 128) {
      shade_block = v;
      const size_t want = (size_t(size) + v - 1) / v, cap = size_t(kNumSMs) * 8 * (block / v);  // the same number of threads in flight
      shade_grid = int(std::min(want < 1 ? size_t(1) : want, cap));
  } } }
  // the connect stage may be compiled for smaller CTAs and a lower register cap than the logic stage (-DB2PT_CONNECT_THREADS / _MIN_BLOCKS)
  const int connect_block = std::min(shade_block, int(B2PT_CONNECT_THREADS));
  const int connect_grid = int(std::min((size_t(size) + connect_block - 1) / connect_block, size_t(kNumSMs) * 8 * (block / connect_block)));
  // Connect stage with its gathers staged through shared memory (cp.async, see above; measured negative): B2PT_STAGED=2
  bool staged_connect = false;
  { const char* e = getenv("B2PT_STAGED"); if (e && shade_block == block) staged_connect = e[0] == '2'; }
  int staged_grid = std::min(grid, kNumSMs * B2PT_LOGIC_MIN_BLOCKS);
  { const char* e = getenv("B2PT_STAGED_GRID"); if (e) { const int v = atoi(e); if (v >= 1 && v <= 8) staged_grid = std::min(grid, kNumSMs * v); } }
  if (staged_connect) {
    static bool configured = false;  // per process; the attribute is per function
    if (!configured) {
      B2_CUDA(cudaFuncSetAttribute(bpt_connect_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kConnectStageBytes)));
      configured = true;
    }
  }
  // B2PT_LOGIC_SPLIT=1: the logic stage as two kernels, one per end of the live queue (light subpaths, then eye subpaths)
  bool logic_split = false;
  { const char* e = getenv("B2PT_LOGIC_SPLIT"); if (e) logic_split = e[0] == '1'; }
  // traversal form per stage: 0 = one ray per thread, 1 = persistent warps with dynamic fetch, 2 = the same plus postponed triangle
  // groups (trace.cuh: trace_dyn8), 3 = warp-synchronous rounds with a shared triangle queue (trace_queue8, the default).
  // B2PT_TRACE_STREAM / B2PT_SHADOW_STREAM select them for A/B runs; measured on the LivingRoom stand-in with 2^23-path batches:
  // 0 / 0 76.3, 2 / 2 79.8, 3 / 2 87.4, 2 / 3 83.2, 3 / 3 92.8 Msamples/s (closest-hit / any-hit stage).
  const char* tv = getenv("B2PT_TRACE_STREAM");
  const char* sv = getenv("B2PT_SHADOW_STREAM");
  const bool queue_ok = cfg.scene.num_tris < (1u << 27);  // trace_queue8 packs (lane, triangle) into 32 bits
  const int stream_trace = tv ? (tv[0] == '1' ? 1 : tv[0] == '0' ? 0 : tv[0] == '2' || !queue_ok ? 2 : 3) : (queue_ok ? 3 : 2);
  const int stream_shadow = sv ? (sv[0] == '1' ? 1 : sv[0] == '0' ? 0 : sv[0] == '3' && queue_ok ? 3 : 2) : stream_trace;
  // ray sorting of the closest-hit queue (see ray_sort_key): B2PT_SORT_RAYS=0/1
  int sort_rays = bpt_sort_mode();
  if (sort_rays == 2 && !w->lane[0].pool.oct_q) sort_rays = 0;  // the workspace was built before the mode was switched on
  unsigned long long basic = 0, shadow = 0, numeric = 0, capacity = 0;
  const int kBatchIters = 8;
  const bool log_iterations = getenv("B2PT_LOG_ITER") != nullptr;

  // per-lane progress
  struct Progress { bool active = false; uint32_t pass0 = 0, passes = 0; int it = 0; BptPool pool; };
  Progress prog[kMaxLanes];
  w->lane[0].stream = stream;
  std::vector<std::unique_ptr<StageTimer>> timers;
  timers.emplace_back(new StageTimer(stream));
  if (lanes > 1) B2_CUDA(cudaEventRecord(w->fork, stream));
  for (int k = 1; k < lanes; ++k) {
    w->lane[k].stream = w->lane[k].own_stream;
    timers.emplace_back(new StageTimer(w->lane[k].stream));
    // the extra lanes start after whatever the caller queued before this call
    B2_CUDA(cudaStreamWaitEvent(w->lane[k].stream, w->fork, 0));
  }
  uint32_t next_pass = 0;
  auto start_batch = [&](int k) -> int {
    Progress& g = prog[k];
    BptLane& l = w->lane[k];
    g.active = true; g.pass0 = next_pass; g.passes = min(passes_per_batch, num_passes - next_pass); g.it = 0;
    next_pass += g.passes;
    g.pool = l.pool;
    g.pool.valid = g.passes * pixels;  // the last batch may be shorter; SoA strides shrink with it, which is fine
    g.pool.size = (g.pool.valid + block - 1) / block * block;
    B2_CUDA(cudaMemsetAsync(l.counters, 0, sizeof(BptCounters), l.stream));
    return B2PT_OK;
  };
  for (int k = 0; k < lanes && next_pass < num_passes; ++k)
    if (int rc = start_batch(k)) return rc;
  for (;;) {
    bool any = false;
    // enqueue a burst of stage iterations on every active lane, then look at the counters
    for (int k = 0; k < lanes; ++k) {
      Progress& g = prog[k];
      if (!g.active) continue;
      any = true;
      BptLane& l = w->lane[k];
      StageTimer& timer = *timers[k];
      for (int j = 0; j < kBatchIters; ++j, ++g.it) {
        const int parity = g.it & 1;
        timer.begin(kStageShade);
        if (g.it == 0) {
          if (sort_rays) bpt_logic_kernel<0, true, true><<<shade_grid, shade_block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, g.pass0, sort_rays);
          else bpt_logic_kernel<0, true><<<shade_grid, shade_block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, g.pass0, 0);
        } else if (sort_rays) {
          bpt_logic_kernel<0, false, true><<<shade_grid, shade_block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, g.pass0, sort_rays);
        } else if (logic_split) {
          bpt_logic_kernel<2, false><<<shade_grid, shade_block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, g.pass0, 0);
          bpt_logic_kernel<1, false><<<shade_grid, shade_block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, g.pass0, 0);
        } else bpt_logic_kernel<0, false><<<shade_grid, shade_block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, g.pass0, 0);
        timer.end();
        timer.begin(kStageShade);
        if (staged_connect) bpt_connect_kernel<true><<<staged_grid, block, kConnectStageBytes, l.stream>>>(cfg, g.pool, l.counters, parity);
        else bpt_connect_kernel<false><<<connect_grid, connect_block, 0, l.stream>>>(cfg, g.pool, l.counters, parity);
        timer.end();
        if (cfg.scene.has_directional) {
          timer.begin(kStageTrace);
          directional_kernel<<<grid, block, 0, l.stream>>>(cfg.scene, g.pool.dir_q, &l.counters->dir_count[parity], g.pool.size,
                                                          DirectionalParams{cfg.params.beta, 1.0f, 0.0f, 0.0f}, g.pool.rec + kRecRad, uint32_t(kRecFloat4s), &l.counters->basic_rays);
          timer.end();
        }
        if (sort_rays == 1) {
          timer.begin(kStageShade);
          size_t bytes = w->scan_bytes;
          B2_CUDA(cub::DeviceScan::ExclusiveSum(w->scan_tmp + w->scan_bytes * k, bytes, g.pool.hist, g.pool.offs, int(2 * kSortBins), l.stream));
          bpt_sort_scatter_kernel<<<grid, block, 0, l.stream>>>(g.pool, l.counters, parity);
          B2_CUDA(cudaMemsetAsync(g.pool.hist, 0, size_t(2 * kSortBins) * 4, l.stream));
          timer.end();
        }
        timer.begin(kStageTrace);
        if (stream_trace == 3) bpt_trace_kernel<3><<<grid_t, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, sort_rays);
        else if (stream_trace == 2) bpt_trace_kernel<2><<<grid_t, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, sort_rays);
        else if (stream_trace == 1) bpt_trace_kernel<1><<<grid_t, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, sort_rays);
        else bpt_trace_kernel<0><<<grid, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity, sort_rays);
        timer.end();
        timer.begin(kStageShadow);
        if (stream_shadow == 3) bpt_shadow_kernel<3><<<grid_t, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity);
        else if (stream_shadow == 2) bpt_shadow_kernel<2><<<grid_t, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity);
        else if (stream_shadow == 1) bpt_shadow_kernel<1><<<grid_t, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity);
        else bpt_shadow_kernel<0><<<grid, block, 0, l.stream>>>(cfg, g.pool, l.counters, parity);
        timer.end();
        if (log_iterations) {  // B2PT_LOG_ITER=1: queue lengths of every iteration (synchronises: for matching an ncu launch with its ray count)
          B2_CUDA(cudaMemcpyAsync(l.host_counters, l.counters, sizeof(BptCounters), cudaMemcpyDeviceToHost, l.stream));
          B2_CUDA(cudaStreamSynchronize(l.stream));
          fprintf(stderr, "b2pt_iter lane %d pass0 %u it %d closest %u shadow %u pairs %u\n", k, g.pass0, g.it,
                  l.host_counters->trace_count[parity] + l.host_counters->light_count[parity],
                  min(l.host_counters->shadow_count[parity], g.pool.q_capacity), l.host_counters->pair_count[parity]);
        }
      }
      B2_CUDA(cudaMemcpyAsync(l.host_counters, l.counters, sizeof(BptCounters), cudaMemcpyDeviceToHost, l.stream));
    }
    if (!any) break;
    for (int k = 0; k < lanes; ++k) {
      Progress& g = prog[k];
      if (!g.active) continue;
      BptLane& l = w->lane[k];
      B2_CUDA(cudaStreamSynchronize(l.stream));
      timers[k]->resolve();
      if (g.it > (1 << 20)) { set_error("render_bpt: the batch did not drain"); return B2PT_ERR_CUDA; }
      if (l.host_counters->done < g.pool.size) continue;
      timers[k]->begin(kStageShade);
      bpt_commit_kernel<<<grid_for(pixels, block), block, 0, l.stream>>>(cfg, g.pool, l.counters, g.pass0, g.passes);
      timers[k]->end();
      B2_CUDA(cudaMemcpyAsync(l.host_counters, l.counters, sizeof(BptCounters), cudaMemcpyDeviceToHost, l.stream));
      B2_CUDA(cudaStreamSynchronize(l.stream));
      timers[k]->resolve();
      basic += l.host_counters->basic_rays; shadow += l.host_counters->shadow_rays;
      numeric += l.host_counters->numeric_errors; capacity += l.host_counters->capacity_events;
      g.active = false;
      if (next_pass < num_passes)
        if (int rc = start_batch(k)) return rc;
    }
  }
  for (int k = 1; k < lanes; ++k) {  // the caller's stream continues after the extra lanes
    B2_CUDA(cudaEventRecord(w->lane[k].done, w->lane[k].stream));
    B2_CUDA(cudaStreamWaitEvent(stream, w->lane[k].done, 0));
  }
  B2_CUDA(cudaGetLastError());
  if (stats) {
    stats->num_basic_rays = basic;
    stats->num_shadow_rays = shadow;
    stats->numeric_errors = numeric;
    stats->capacity_events = capacity;
    timers[0]->fill(stats);
    for (int k = 1; k < lanes; ++k) {
      B2ptStats s1{};
      timers[k]->fill(&s1);
      stats->trace_ms += s1.trace_ms; stats->shadow_ms += s1.shadow_ms; stats->shade_ms += s1.shade_ms;
      stats->trace_launches += s1.trace_launches; stats->shadow_launches += s1.shadow_launches;
      stats->kernel_launches += s1.kernel_launches;
    }
  }
  if (capacity) return capacity_error(capacity, "render_bpt");
  return B2PT_OK;
}

}  // namespace b2pt
