// vcm.cu — wavefront VCM / UPG: UPGBase<Beta>::_preprocess / _scatter / _traceLight / _traceEye / _connect* /
// _gather* / _merge* and the VC / VM MIS sub-weights (UPG.cpp:40-831).
//
// One pass (= one Technique::render call of the reference) is three device phases:
//
//   photons  (vcm_photon_kernel + trace)   num_photons light paths, one slot each, walk the reference's _traceLight
//            state machine; every accepted vertex is bump-allocated in ONE pool in HBM (112 B records) and linked
//            to its predecessor, so `_light_paths[index - 1]` of the reference is `link[index]` here and
//            `_light_offsets[p] .. [p + 1]` is the chain hanging off head[p].
//   grid     (grid.cuh)                    the non-light vertices are radix-sorted by cell into the photon index.
//   eye      (vcm_eye_kernel -> vcm_gather_kernel -> vcm_connect_kernel -> trace -> shadow)
//            per pixel the reference's _traceEye state machine.  Vertex CONNECTION work (own light path of the
//            pixel + one next-event sample) is expanded into (eye vertex, light vertex) pairs exactly as in bpt.cu;
//            vertex MERGING is a range query of the photon index per new eye vertex, evaluated in the gather stage.
//
// VCM (biased, shrinking radius r_i = r_0 (i + 1)^((alpha - 1) / 2), UPG.cpp:180-183) draws nothing inside a merge, so
// given the same per-(pixel, sample) and per-(photon, sample) streams the eye-path radiance is the reference's.
// UPG (unbiased) estimates the merge density by rejection ray casting (BSDF::gathering_density, BSDF.cpp:42-66,
// 264-289, 393-434): a data-dependent loop of bounded rays (Intersector::intersectFast, Intersector.cpp:17-46) that
// consumes random numbers in an order the reference leaves to std::sort; here every merge owns a sub-stream keyed by
// (path stream, eye vertex, photon), so UPG agrees with the reference in distribution, not per path.
//
// Both merge directions of the reference are built: from the light side (Options::from_light = true, the default:
// the photon's predecessor is connected to the tentative eye vertex) and from the camera side (--from-camera: the
// eye vertex BEFORE the tentative one is connected to the photon itself; UPG.cpp:687-719).
#include "grid.cuh"
#include <string>
#include <thread>

#include "../../include/b2pt_debug.h"
#include "render_common.cuh"

namespace b2pt {

namespace {

enum : uint32_t { kNew = 0, kLightWait = 1, kEyeStart = 2, kEyeWait = 3, kDone = 4 };
constexpr uint32_t kPairCamera = 0x80000000u;
constexpr uint32_t kNeeVertex = 0xFFFFFFFFu;
constexpr uint32_t kShadowSplat = 1u;
constexpr int kVertexFloat4s = 7;
constexpr uint32_t kFlagFinite = 1u, kFlagDirectional = 2u, kFlagDead = 0x80000000u;
constexpr float kDensityTestLimit = 16777216.f;  // DENSITY_ESTIMATION_TEST_LIMIT, BSDF.cpp:9

struct VcmCounters {
  unsigned long long basic_rays, shadow_rays, numeric_errors, capacity_events, density_rays;
  unsigned int done;
  unsigned int lv_count;
  unsigned int trace_count[2], pair_count[2], shadow_count[2], gather_count[2], dir_count[2];
  unsigned int merge_count;  // (eye vertex, photon) pairs of the current iteration; cleared by the eye stage
  unsigned int long_count, long_ticket;  // UPG: density loops handed to whole warps (upg_density_long_kernel)
  unsigned int wide_count;               // UPG: density loops handed to the whole grid (upg_density_wide_kernel)
  unsigned int job_count, job_ticket;  // UPG: merges whose density is still to be estimated, and the fetch counter of the stage that does it
  unsigned int ray_ticket;             // ray fetch counter of the persistent closest-hit stages, zeroed by the logic stage before them
  unsigned int shadow_ticket;          // the same of the persistent any-hit stage, zeroed by the eye stage
};

// the per-pass constants of UPGBase (UPG.hpp:119-137)
struct VcmParams {
  float beta, roulette, roulette_inv;
  float radius, circle, clamp_const;
  float num_scattered, num_scattered_inv;
  uint32_t unbiased, enable_vc, enable_vm, from_light;
  uint32_t num_photons, max_subpath;
  uint32_t brute_vertices;  // diagnostic (B2PT_VCM_BRUTE=1): gather scans every stored vertex instead of the grid
};

// UPGBase::LightVertex / EyeVertex (UPG.hpp:19-43): w0 = a | c, w1 = A | C, w2 = B | D
struct Vertex {
  Surface s;
  float3 omega, thr;
  float w0, w1, w2, bgeo;
  uint32_t flags;  // kFlagFinite | kFlagDirectional | length << 2

  __device__ __forceinline__ uint32_t finite() const { return flags & kFlagFinite; }
  __device__ __forceinline__ float length() const { return float((flags & ~kFlagDead) >> 2); }
  __device__ __forceinline__ void set(uint32_t finite, uint32_t directional, uint32_t length) {
    flags = (finite & 1u) | (directional ? kFlagDirectional : 0u) | (length << 2);
  }
  __device__ __forceinline__ void set_finite(uint32_t f) { flags = (flags & ~kFlagFinite) | (f & 1u); }
};

struct VcmPool {
  uint32_t photons;      // photon slots (rounded up to the block size)
  uint32_t pixels_pad;   // eye slots (rounded up)
  uint32_t lv_capacity, q_capacity;
  // photon phase, per photon slot
  uint32_t* p_state; uint2* p_rng; float4* p_cv; float4* p_bs_omega; float4* p_bs_thr; float4* p_ray_o; float4* p_hit;
  uint32_t* lp_head; uint32_t* lp_count;
  // eye phase, per pixel slot
  uint32_t* state; uint2* rng; float4* cv; float4* pv; float4* bs_omega; float4* bs_thr; float4* ray_o; float4* hit; float4* rad; float4* nee;
  double* splat;
  // vertex pool
  float4* lv; uint32_t* lv_link; float4* grid_points;
  // queues
  float4* job_q;  // UPG density jobs, 3 float4 each: (unoccluded * weight, bits slot) (target, bits vertex) (bits key, -, -, -); m_capacity entries
  uint2* merge_q; uint32_t m_capacity;  // (eye slot, photon index) pairs found by the gather stage, evaluated one per thread
  uint2* wide_q;      // UPG: (job, first untried trial) of the density loops the whole grid finishes; kWideCap entries
  uint2* wide_state;  // per entry: (first trial that landed, chunk ticket); (0xFFFFFFFF, 0) between uses
  uint32_t* trace_q[2];  // live slots (extension ray in flight), written by iteration i into [i & 1]: the next logic stage walks it
  uint2* pair_q; float4* shadow_q; uint32_t* gather_q;
  float4* dir_q;  // connections to sun lights, 4 float4 per entry (render_common.cuh: directional_kernel)
};

__device__ __forceinline__ void pack_vertex(const Vertex& v, float4* q) {
  q[0] = make_float4(v.s.pos.x, v.s.pos.y, v.s.pos.z, __uint_as_float(v.s.material_id));
  q[1] = make_float4(v.s.gnormal.x, v.s.gnormal.y, v.s.gnormal.z, __uint_as_float(v.flags));
  q[2] = make_float4(v.s.tan.c0.x, v.s.tan.c0.y, v.s.tan.c0.z, v.w0);
  q[3] = make_float4(v.s.tan.c1.x, v.s.tan.c1.y, v.s.tan.c1.z, v.w1);
  q[4] = make_float4(v.s.tan.c2.x, v.s.tan.c2.y, v.s.tan.c2.z, v.w2);
  q[5] = make_float4(v.omega.x, v.omega.y, v.omega.z, v.bgeo);
  q[6] = make_float4(v.thr.x, v.thr.y, v.thr.z, 0.f);
}
__device__ __forceinline__ Vertex unpack_vertex(const float4* q) {
  Vertex v;
  v.s.pos = f3(q[0].x, q[0].y, q[0].z); v.s.material_id = __float_as_uint(q[0].w);
  v.s.gnormal = f3(q[1].x, q[1].y, q[1].z); v.flags = __float_as_uint(q[1].w);
  v.s.tan.c0 = f3(q[2].x, q[2].y, q[2].z); v.w0 = q[2].w;
  v.s.tan.c1 = f3(q[3].x, q[3].y, q[3].z); v.w1 = q[3].w;
  v.s.tan.c2 = f3(q[4].x, q[4].y, q[4].z); v.w2 = q[4].w;
  v.omega = f3(q[5].x, q[5].y, q[5].z); v.bgeo = q[5].w;
  v.thr = f3(q[6].x, q[6].y, q[6].z);
  return v;
}
// slot vertices are SoA (stride = number of slots), pool vertices AoS
// A slot's vertex is ONE 128-byte-aligned record (seven words + one spare): the stages reach slots through queues, and the same words as
// seven SoA streams were seven half-used sectors per vertex (the PT and BPT pools made the same move).  `stride` (the slot count) is unused.
constexpr int kSlotVertexFloat4s = 8;
__device__ __forceinline__ void store_slot_vertex(float4* base, uint32_t /*stride*/, uint32_t slot, const Vertex& v) {
  float4 q[kVertexFloat4s];
  pack_vertex(v, q);
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) base[size_t(slot) * kSlotVertexFloat4s + k] = q[k];
}
__device__ __forceinline__ Vertex load_slot_vertex(const float4* base, uint32_t /*stride*/, uint32_t slot) {
  float4 q[kVertexFloat4s];
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) q[k] = base[size_t(slot) * kSlotVertexFloat4s + k];
  return unpack_vertex(q);
}
__device__ __forceinline__ void store_pool_vertex(const VcmPool& p, uint32_t idx, const Vertex& v) {
  float4 q[kVertexFloat4s];
  pack_vertex(v, q);
  float4* o = p.lv + size_t(idx) * kVertexFloat4s;
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) o[k] = q[k];
}
__device__ __forceinline__ Vertex load_pool_vertex(const VcmPool& p, uint32_t idx) {
  float4 q[kVertexFloat4s];
  const float4* o = p.lv + size_t(idx) * kVertexFloat4s;
#pragma unroll
  for (int k = 0; k < kVertexFloat4s; ++k) q[k] = __ldg(o + k);
  return unpack_vertex(q);
}

__device__ __forceinline__ float clamp_vm(const VcmParams& P, float x) { return fminf(P.clamp_const, x); }  // UPGBase::_clamp

// UPGBase::_sample_to_vertex (UPG.cpp:188-201)
__device__ __forceinline__ Vertex sample_to_vertex(const LightSample& sample, const VcmParams& P) {
  Vertex v;
  v.s = sample.surface;
  v.omega = v.s.normal();
  v.thr = sample.radiance / sample.combined_density() * P.roulette_inv;
  v.w0 = sample.directional ? 0.0f : 1.0f / sample.combined_density();
  v.w1 = 0.0f; v.w2 = 0.0f; v.bgeo = 0.0f;
  v.set(1u, sample.directional ? 1u : 0u, 0u);
  return v;
}

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

// UPGBase::Connection (UPG.hpp:45-53)
struct Connection {
  const Vertex& light;
  const Vertex& eye;
  BsdfQuery lb, eb;
  Edge edge;
  __device__ __forceinline__ Connection(const DeviceScene& sc, const Vertex& l, const Vertex& e, float3 omega)
      : light(l), eye(e), lb(bsdf_query(sc, l.s, l.omega, omega)), eb(bsdf_query(sc, e.s, -omega, e.omega)),
        edge(l.s.pos, l.s.normal(), e.s.pos, e.s.normal(), omega) {}
  // _connect(connection) without the visibility factor (UPG.cpp:438-447)
  __device__ __forceinline__ float3 unoccluded() const { return light.thr * lb.throughput * eye.thr * eb.throughput * edge.b_cos * edge.f_geometry; }
};

// _vc_subweight_inv (UPG.cpp:294-305)
__device__ __forceinline__ float vc_subweight_inv(const Connection& c, const VcmParams& P) {
  const float b = P.beta;
  const float Ap = (c.light.w1 * beta_fn(c.lb.density_rev, b) + float(c.light.finite()) * beta_fn(c.light.w0, b)) * beta_fn(c.edge.b_geometry * c.eb.density_rev, b);
  const float Cp = (c.eye.w1 * beta_fn(c.eb.density, b) + float(c.eye.finite()) * beta_fn(c.eye.w0, b)) * beta_fn(c.edge.f_geometry * c.lb.density, b);
  return Ap + Cp + 1.0f;
}

// the Bp + Dp part shared by _vm_subweight_inv and _vm_biased_subweight_inv (UPG.cpp:308-377)
__device__ __forceinline__ float vm_bd(const Connection& c, const VcmParams& P) {
  const float b = P.beta;
  float light_vm, eye_vm;
  if (P.from_light) {
    light_vm = clamp_vm(P, beta_fn(P.circle / c.light.w0, b));
    eye_vm = clamp_vm(P, beta_fn(P.circle * c.eye.bgeo * c.eb.density, b)) * (c.eye.length() <= 1.0f ? 0.0f : 1.0f);
  } else {
    light_vm = clamp_vm(P, beta_fn(P.circle * c.light.bgeo * c.lb.density_rev, b)) * (c.light.length() <= 1.0f ? 0.0f : 1.0f);
    eye_vm = clamp_vm(P, beta_fn(P.circle / c.eye.w0, b));
  }
  const float Bp = (c.light.w2 * beta_fn(c.lb.density_rev, b) + float(c.lb.finite) * light_vm * beta_fn(c.light.w0, b) * (c.light.length() <= 1.0f ? 0.0f : 1.0f)) *
                   beta_fn(c.edge.b_geometry * c.eb.density_rev, b);
  const float Dp = (c.eye.w2 * beta_fn(c.eb.density, b) + float(c.eb.finite) * eye_vm * beta_fn(c.eye.w0, b) * (c.eye.length() <= 1.0f ? 0.0f : 1.0f)) *
                   beta_fn(c.edge.f_geometry * c.lb.density, b);
  return Bp + Dp;
}
__device__ __forceinline__ float vm_subweight_inv(const Connection& c, const VcmParams& P) {
  const float area_density = P.from_light ? c.edge.f_geometry * c.lb.density : c.edge.b_geometry * c.eb.density_rev;
  const float connect_vm = clamp_vm(P, beta_fn(P.circle * area_density, P.beta)) * (c.eye.length() * c.light.length() == 0.0f ? 0.0f : 1.0f);
  return beta_fn(P.num_scattered, P.beta) * (vm_bd(c, P) + connect_vm);
}
__device__ __forceinline__ float vm_biased_subweight_inv(const Connection& c, const VcmParams& P, float connect_vm) {
  return beta_fn(P.num_scattered, P.beta) * (vm_bd(c, P) + connect_vm);
}
// _vc_weight / _vc_biased_weight (UPG.cpp:379-403)
__device__ __forceinline__ float vc_weight(const Connection& c, const VcmParams& P) {
  if (c.eye.length() + c.light.length() < 2.0f) return 1.0f / vc_subweight_inv(c, P);
  return 1.0f / (float(P.enable_vc) * vc_subweight_inv(c, P) + float(P.enable_vm) * vm_subweight_inv(c, P));
}
__device__ __forceinline__ float vc_biased_weight(const Connection& c, const VcmParams& P, float vm_current) {
  if (c.eye.length() + c.light.length() < 2.0f) return 1.0f / vc_subweight_inv(c, P);
  return 1.0f / (float(P.enable_vc) * vc_subweight_inv(c, P) + float(P.enable_vm) * vm_biased_subweight_inv(c, P, vm_current));
}

// UPGBase::_connect(light, eye) without the visibility factor (UPG.cpp:527-551)
__device__ __forceinline__ float3 connect_vertices(const DeviceScene& sc, const Vertex& light, const Vertex& eye, const VcmParams& P) {
  const float3 omega = normalize(eye.s.pos - light.s.pos);
  const Connection c(sc, light, eye, omega);
  const float3 throughput = c.unoccluded();
  if (!(l1norm(throughput) > kFltEpsilon)) return f3(0.0f);
  const float vm_current = clamp_vm(P, beta_fn(P.circle * c.edge.f_geometry * c.lb.density, P.beta)) * (eye.length() == 0.0f ? 0.0f : 1.0f);
  const float weight = P.unbiased ? vc_weight(c, P) : vc_biased_weight(c, P, vm_current);
  return throughput * weight;
}

// Technique::_camera_coefficient (Technique.cpp:246-274)
__device__ __forceinline__ float camera_coefficient(float3 light_omega, float3 light_gnormal, float3 light_normal, float3 eye_omega, float3 eye_normal) {
  const float normal_c = fabsf(dot(eye_omega, light_gnormal) * dot(light_omega, light_normal) / (dot(eye_omega, light_normal) * dot(light_omega, light_gnormal)));
  const float c = fabsf(dot(eye_normal, eye_omega));
  return normal_c * (1.0f / cr_cubef(c));
}

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

// ---------------------------------------------------------------- photon phase: UPGBase::_traceLight (UPG.cpp:208-292)
// One slot per photon path; stream (seed, kDomainPhoton, photon, sample).  States: kNew -> kLightWait (an
// intersectMesh ray is in flight) -> kDone.
// Like every logic stage here it walks the LIVE paths only (the trace queue of the previous iteration; `first` walks all slots).
__global__ void __launch_bounds__(256) vcm_photon_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt, int parity, uint32_t sample, int first) {
  const DeviceScene& sc = cfg.scene;
  if (blockIdx.x == 0 && threadIdx.x == 0) cnt->ray_ticket = 0;
  const float beta = P.beta;
  const uint32_t n_live = first ? pool.photons : cnt->trace_count[parity ^ 1];
  const uint32_t n_round = (n_live + 31u) & ~31u;
  const uint32_t* live_q = pool.trace_q[parity ^ 1];
  uint32_t* next_q = pool.trace_q[parity];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
    const bool lane_on = i < n_live;
    const uint32_t slot = lane_on ? (first ? i : live_q[i]) : 0u;
    uint32_t st = lane_on ? pool.p_state[slot] : uint32_t(kDone);
    bool want_trace = false, finished = false, terminate = false;
    uint32_t head = kInvalid, count = 0;
    Rng rng; rng.key = 0; rng.ctr = 0;
    Vertex cur;
    BsdfSample bs;
    bs.omega = f3(0.f); bs.throughput = f3(0.f); bs.density = 0.f; bs.density_rev = 0.f; bs.finite = 1;
    float3 O = f3(0.f);
    if (st != kDone) {
      if (st == kNew && slot >= P.num_photons) {
        finished = true;  // padding slot
      } else if (st == kNew) {
        rng.key = stream_key(cfg.params.seed, kDomainPhoton, slot, sample);
        pool.lp_head[slot] = kInvalid; pool.lp_count[slot] = 0;
        if (P.roulette < rng.sample()) {  // UPG.cpp:210-212
          finished = true;
        } else {
          cur = sample_to_vertex(sample_light(sc, rng), P);
          const unsigned idx = atomicAdd(&cnt->lv_count, 1u);
          if (idx >= pool.lv_capacity) {
            atomicAdd(&cnt->capacity_events, 1ull);
            finished = true;
          } else {
            store_pool_vertex(pool, idx, cur);
            pool.lv_link[idx] = kInvalid;
            head = idx; count = 1;
            bs = bsdf_sample(sc, rng, cur.s, cur.omega);
            st = kLightWait;
            // while (!_russian_roulette(generator)) -> intersectMesh
            if (P.roulette < rng.sample()) terminate = true; else want_trace = true;
          }
        }
      } else {  // kLightWait: consume the closest mesh hit (UPG.cpp:227-285)
        const uint2 rs = pool.p_rng[slot];
        rng.key = rs.x; rng.ctr = rs.y;
        head = pool.lp_head[slot]; count = pool.lp_count[slot];
        cur = load_slot_vertex(pool.p_cv, pool.photons, slot);
        const float4 bo = pool.p_bs_omega[slot], bt = pool.p_bs_thr[slot], ro = pool.p_ray_o[slot], hv = pool.p_hit[slot];
        bs.omega = f3(bo.x, bo.y, bo.z); bs.density = bo.w;
        bs.throughput = f3(bt.x, bt.y, bt.z); bs.density_rev = bt.w;
        bs.finite = __float_as_uint(ro.w) & 1u;
        O = f3(ro.x, ro.y, ro.z);
        const uint32_t tri = __float_as_uint(hv.w);
        if (tri == kInvalid) {
          terminate = true;
        } else {
          Vertex nv;
          nv.s = query_surface(sc, tri, hv.y, hv.z, hv.x, O, bs.omega);
          nv.omega = -bs.omega;
          const BsdfSample nbs = bsdf_sample(sc, rng, nv.s, nv.omega);  // new_bsdf is drawn before the throughput test
          const Edge edge(cur.s.pos, cur.s.normal(), nv.s.pos, nv.s.normal(), nv.omega);
          nv.thr = cur.thr * bs.throughput * edge.b_cos * P.roulette_inv;
          if (l1norm(nv.thr) < kFltEpsilon) {
            terminate = true;
          } else {
            nv.thr /= bs.density;
            nv.bgeo = edge.b_geometry;
            const uint32_t len = uint32_t(cur.length()) + 1u;
            cur.set_finite(min(cur.finite(), bs.finite));
            nv.set(bs.finite, 0u, len);
            nv.w0 = 1.0f / (edge.f_geometry * bs.density);
            nv.w1 = (cur.w1 * beta_fn(bs.density_rev, beta) + float(cur.finite()) * beta_fn(cur.w0, beta)) * beta_fn(edge.b_geometry * nv.w0, beta);
            const float vertex_merging = P.from_light ? clamp_vm(P, beta_fn(P.circle / cur.w0, beta))
                                                      : clamp_vm(P, beta_fn(P.circle * cur.bgeo * bs.density_rev, beta)) * (cur.length() <= 1.0f ? 0.0f : 1.0f);
            nv.w2 = (cur.w2 * beta_fn(bs.density_rev, beta) + float(bs.finite) * vertex_merging * (cur.length() <= 1.0f ? 0.0f : 1.0f) * beta_fn(cur.w0, beta)) *
                    beta_fn(edge.b_geometry * nv.w0, beta);
            pool.lv[size_t(head) * kVertexFloat4s + 1].w = __uint_as_float(cur.flags);  // prv->finite, updated in place
            unsigned idx = kInvalid;
            if (count < P.max_subpath) idx = atomicAdd(&cnt->lv_count, 1u);
            if (idx >= pool.lv_capacity) {
              atomicAdd(&cnt->capacity_events, 1ull);  // truncated: the stored prefix stays valid
              terminate = true;
            } else {
              store_pool_vertex(pool, idx, nv);
              pool.lv_link[idx] = head;
              head = idx; ++count;
              cur = nv; bs = nbs;
              if (P.roulette < rng.sample()) terminate = true; else want_trace = true;
            }
          }
        }
      }
      if (terminate) {
        // UPG.cpp:287-289: a path may not end in a vertex whose pending BSDF sample is a delta lobe
        if (bs.finite == 0) {
          pool.lv[size_t(head) * kVertexFloat4s + 1].w = __uint_as_float(cur.flags | kFlagDead);
          head = pool.lv_link[head];
          --count;
        }
        finished = true;
      }
      if (finished) {
        st = kDone;
        if (head != kInvalid || count) { pool.lp_head[slot] = head; pool.lp_count[slot] = count; }
      } else {
        O = offset_origin(cur.s.pos, cur.s.gnormal, bs.omega);  // Scene::intersectMesh (Scene.cpp:204-208)
        store_slot_vertex(pool.p_cv, pool.photons, slot, cur);
        pool.p_rng[slot] = make_uint2(rng.key, rng.ctr);
        pool.lp_head[slot] = head; pool.lp_count[slot] = count;
        pool.p_bs_omega[slot] = make_float4(bs.omega.x, bs.omega.y, bs.omega.z, bs.density);
        pool.p_bs_thr[slot] = make_float4(bs.throughput.x, bs.throughput.y, bs.throughput.z, bs.density_rev);
        pool.p_ray_o[slot] = make_float4(O.x, O.y, O.z, __uint_as_float(bs.finite & 1u));
      }
      pool.p_state[slot] = st;
    }
    {
      const unsigned mask = __ballot_sync(0xFFFFFFFFu, finished);
      if (mask && (threadIdx.x & 31u) == 0) atomicAdd(&cnt->done, unsigned(__popc(mask)));
    }
    const unsigned tq = warp_append(&cnt->trace_count[parity], want_trace);
    if (want_trace) next_q[tq] = slot;
  }
}

// closest-hit queries of a logic stage as trace_dyn8 reads and writes them (trace.cuh): rays from (origin, direction) streams by slot
struct VcmTraceIo {
  const uint32_t* q;
  const float4* ray_o; const float4* ray_d;
  float4* hit;
  uint32_t mask;
  __device__ __forceinline__ uint32_t load(uint32_t i, float3& O, float3& D, float& tnear, float& tfar, uint32_t& m) const {
    const uint32_t slot = q[i];
    const float4 o = ray_o[slot], d = ray_d[slot];
    O = f3(o.x, o.y, o.z); D = f3(d.x, d.y, d.z); tnear = 0.0f; tfar = INFINITY; m = mask;
    return slot;
  }
  __device__ __forceinline__ void store(uint32_t slot, const Hit& h) const { hit[slot] = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri)); }
};
// kPhoton: light subpaths (Scene::intersectMesh), else eye subpaths (Scene::intersect); persistent warps, dynamic fetch;
// kQueue: trace_queue8 (the default), else trace_dyn8
template <bool kPhoton, bool kQueue>
__global__ void __launch_bounds__(256, 4) vcm_trace_dyn_kernel(RenderConfig cfg, VcmPool pool, VcmCounters* cnt, int parity) {
  const uint32_t n = cnt->trace_count[parity];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cnt->basic_rays, (unsigned long long)n);
    cnt->trace_count[parity ^ 1] = 0;  // consumed by this iteration's logic stage
  }
  VcmTraceIo io{pool.trace_q[parity], kPhoton ? pool.p_ray_o : pool.ray_o, kPhoton ? pool.p_bs_omega : pool.bs_omega,
                kPhoton ? pool.p_hit : pool.hit, kPhoton ? kMaskMesh : kMaskAll};
  if (kQueue) trace_queue8<false>(cfg.scene, n, &cnt->ray_ticket, io);
  else trace_dyn8<false>(cfg.scene, n, &cnt->ray_ticket, io);
}

__global__ void __launch_bounds__(256) vcm_photon_trace_kernel(RenderConfig cfg, VcmPool pool, VcmCounters* cnt, int parity) {
  const uint32_t n = cnt->trace_count[parity];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cnt->basic_rays, (unsigned long long)n);
    cnt->trace_count[parity ^ 1] = 0;  // consumed by this iteration's logic stage
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t slot = pool.trace_q[parity][i];
    const float4 o = pool.p_ray_o[slot], d = pool.p_bs_omega[slot];
    const Hit h = trace_ray<false, true>(cfg.scene, f3(o.x, o.y, o.z), f3(d.x, d.y, d.z), 0.0f, INFINITY);
    pool.p_hit[slot] = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri));
  }
}

// the points of the photon index: every stored vertex that is not on a light (HashGrid3D.hpp:252-255) and was not
// dropped at the end of its path
__global__ void vcm_grid_points_kernel(VcmPool pool, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 q0 = pool.lv[size_t(i) * kVertexFloat4s + 0], q1 = pool.lv[size_t(i) * kVertexFloat4s + 1];
  const bool is_light = (__float_as_uint(q0.w) & 3u) == B2PT_ENTITY_LIGHT;
  const bool dead = (__float_as_uint(q1.w) & kFlagDead) != 0u;
  pool.grid_points[i] = make_float4(q0.x, q0.y, q0.z, __uint_as_float(is_light || dead ? kInvalidPoint : i));
}

// ---------------------------------------------------------------- eye phase: UPGBase::_traceEye (UPG.cpp:40-174)
// UPGBase::_connect_light (UPG.cpp:449-476)
__device__ __forceinline__ float3 connect_light(const DeviceScene& sc, const Vertex& eye, float Dp, const VcmParams& P) {
  const BsdfQuery cb = bsdf_query(sc, eye.s, f3(0.0f), eye.omega);
  if (l1norm(cb.throughput) < kFltMin) return f3(0.0f);
  float3 lr; float ld;
  query_lsdf(sc, eye.s, eye.omega, lr, ld);
  const float Cp = (eye.w1 * beta_fn(cb.density, P.beta) + beta_fn(eye.w0, P.beta) * float(eye.finite())) * beta_fn(ld, P.beta);
  const float weight_inv = Cp + 1.0f + beta_fn(P.num_scattered, P.beta) * Dp;
  return lr * eye.thr / weight_inv;
}

__global__ void __launch_bounds__(256) vcm_eye_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt, int parity, uint32_t sample, int first) {
  const DeviceScene& sc = cfg.scene;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    cnt->pair_count[parity ^ 1] = 0; cnt->merge_count = 0; cnt->job_count = 0; cnt->job_ticket = 0; cnt->long_count = 0; cnt->long_ticket = 0; cnt->wide_count = 0;
    cnt->shadow_count[parity ^ 1] = 0; cnt->gather_count[parity ^ 1] = 0; cnt->dir_count[parity ^ 1] = 0;
    cnt->ray_ticket = 0; cnt->shadow_ticket = 0;
  }
  const float beta = P.beta;
  const uint32_t n_live = first ? pool.pixels_pad : cnt->trace_count[parity ^ 1];
  const uint32_t n_round = (n_live + 31u) & ~31u;
  const uint32_t* live_q = pool.trace_q[parity ^ 1];
  uint32_t* next_q = pool.trace_q[parity];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
    const bool lane_on = i < n_live;
    const uint32_t slot = lane_on ? (first ? i : live_q[i]) : 0u;
    uint32_t st = lane_on ? pool.state[slot] : uint32_t(kDone);
    bool want_trace = false, want_gather = false, finished = false, store_cur = false;
    uint32_t emit_n = 0, emit_flags = 0;
    bool emit_nee = false, emit_skip_first = false;
    uint32_t head = kInvalid, count = 0;
    float3 O = f3(0.f), rad = f3(0.f);
    Rng rng; rng.key = 0; rng.ctr = 0;
    Vertex cur;
    BsdfSample bs;
    bs.omega = f3(0.f); bs.throughput = f3(0.f); bs.density = 0.f; bs.density_rev = 0.f; bs.finite = 1;
    if (st != kDone) {
      if (st == kNew && slot >= cfg.pixels) {
        finished = true;  // padding slot
      } else if (st == kNew) {
        rng.key = stream_key(cfg.params.seed, kDomainPixel, slot, sample);
        const float3 dir = shoot_primary(cfg, rng, slot);  // Technique::_for_each_ray (Technique.cpp:321-331)
        if (P.roulette < rng.sample()) {  // UPG.cpp:45-47
          finished = true;
        } else {
          cur.s = cfg.camera;
          cur.omega = -dir;
          cur.thr = f3(1.0f) * P.roulette_inv;
          cur.w0 = 0.0f; cur.w1 = 0.0f; cur.w2 = 0.0f; cur.bgeo = 0.0f;
          cur.set(1u, 0u, 0u);
          store_cur = true;
          head = pool.lp_head[slot]; count = pool.lp_count[slot];
          if (P.enable_vc) { emit_n = count; emit_flags = kPairCamera; }  // _connect(context, camera vertex), UPG.cpp:559-578
          bs = bsdf_sample(sc, rng, cur.s, cur.omega);  // CameraBSDF::sample: no draws
          O = offset_origin(cur.s.pos, cur.s.gnormal, bs.omega);
          want_trace = true;
          st = kEyeWait;
        }
      } else {  // kEyeWait: UPG.cpp:71-143
        const uint2 rs = pool.rng[slot];
        rng.key = rs.x; rng.ctr = rs.y;
        head = pool.lp_head[slot]; count = pool.lp_count[slot];
        const float4 rv = pool.rad[slot];
        rad = f3(rv.x, rv.y, rv.z);
        cur = load_slot_vertex(pool.cv, pool.pixels_pad, slot);
        const float4 bo = pool.bs_omega[slot], bt = pool.bs_thr[slot], ro = pool.ray_o[slot], hv = pool.hit[slot];
        bs.omega = f3(bo.x, bo.y, bo.z); bs.density = bo.w;
        bs.throughput = f3(bt.x, bt.y, bt.z); bs.density_rev = bt.w;
        bs.finite = __float_as_uint(ro.w) & 1u;
        O = f3(ro.x, ro.y, ro.z);
        const uint32_t tri = __float_as_uint(hv.w);
        if (tri == kInvalid) {
          if (cur.s.is_camera()) {  // return sky_gradient(bsdf.omega) *= _roulette_inv (UPG.cpp:76-78)
            const float3 h = f3(cfg.params.sky_horizon[0], cfg.params.sky_horizon[1], cfg.params.sky_horizon[2]);
            const float3 z = f3(cfg.params.sky_zenith[0], cfg.params.sky_zenith[1], cfg.params.sky_zenith[2]);
            rad = (h * (1 - bs.omega.z) + z * bs.omega.z) * P.roulette_inv;
          }
          finished = true;
        } else {
          Vertex nv;
          nv.s = query_surface(sc, tri, hv.y, hv.z, hv.x, O, bs.omega);
          nv.omega = -bs.omega;
          const BsdfSample nbs = bsdf_sample(sc, rng, nv.s, nv.omega);  // new_bsdf: drawn on every hit, lights included
          const Edge edge(cur.s.pos, cur.s.normal(), nv.s.pos, nv.s.normal(), nv.omega);
          nv.thr = cur.thr * bs.throughput * edge.b_cos;
          if (l1norm(nv.thr) < kFltEpsilon) {
            finished = true;
          } else {
            nv.thr /= bs.density;
            nv.bgeo = edge.b_geometry;
            const uint32_t len = uint32_t(cur.length()) + 1u;
            cur.set_finite(min(cur.finite(), bs.finite));
            nv.set(bs.finite, 0u, len);
            nv.w0 = 1.0f / (edge.f_geometry * bs.density);
            nv.w1 = (cur.w1 * beta_fn(bs.density_rev, beta) + float(cur.finite()) * beta_fn(cur.w0, beta)) * beta_fn(edge.b_geometry * nv.w0, beta);
            const float trim = cur.length() <= 1.0f ? 0.0f : 1.0f;
            const float vertex_merging = P.from_light ? clamp_vm(P, beta_fn(P.circle * cur.bgeo * bs.density_rev, beta)) * trim
                                                      : clamp_vm(P, beta_fn(P.circle / cur.w0, beta));
            const float dterm = cur.w2 * beta_fn(bs.density_rev, beta) + float(cur.finite()) * vertex_merging * trim * beta_fn(cur.w0, beta);
            nv.w2 = dterm * beta_fn(edge.b_geometry * nv.w0, beta);
            if (nv.s.is_light()) {
              if (P.enable_vc) {  // UPG.cpp:121-139
                float3 lr; float ld;
                query_lsdf(sc, nv.s, nv.omega, lr, ld);
                const BsdfQuery cb = bsdf_query(sc, nv.s, f3(0.0f), nv.omega);
                float Dp = dterm * beta_fn(edge.b_geometry * cb.density, beta);
                Dp *= (nv.length() <= 2.0f ? 0.0f : 1.0f) * float(P.enable_vm);
                Dp *= beta_fn(ld * nv.w0, beta);
                rad += connect_light(sc, nv, Dp, P);
              }
              // on through the light: prv and bsdf stay, the next ray starts at the light's surface
              O = offset_origin(nv.s.pos, nv.s.gnormal, bs.omega);
              store_cur = true;  // prv->finite
              want_trace = true;
            } else {
              nv.thr = nv.thr * P.roulette_inv;
              // _gather / _gather_biased(context, *prv, *itr), UPG.cpp:148-156; from the camera side the merge is evaluated at
              // *prv (with prv->finite already updated) and a camera prv never merges (UPG.cpp:690,716)
              want_gather = P.enable_vm != 0u && (P.from_light || !cur.s.is_camera());
              if (!P.from_light && want_gather) store_slot_vertex(pool.pv, pool.pixels_pad, slot, cur);
              cur = nv; bs = nbs;
              store_cur = true;
              if (P.roulette < rng.sample()) {  // UPG.cpp:160-162
                finished = true;
              } else {
                if (P.enable_vc) {  // _connect(context, eye) on a surface vertex (UPG.cpp:579-595)
                  if (!(P.roulette < rng.sample())) {
                    const LightSample ls = sample_light(sc, rng);
                    if (!ls.directional) {
                      emit_nee = true;
                      const uint32_t light_id = sc.materials[ls.surface.material_index()].light_id;
                      pool.nee[slot] = make_float4(ls.surface.pos.x, ls.surface.pos.y, ls.surface.pos.z, __uint_as_float(light_id));
                    } else {
                      // _connect_directional (UPG.cpp:480-525): queued for the directional stage
                      const float3 n = ls.surface.normal();
                      const BsdfQuery cb = bsdf_query(sc, cur.s, -n, cur.omega);
                      const float trim1 = cur.length() <= 1.0f ? 0.0f : 1.0f;
                      const float eye_vm = P.from_light ? clamp_vm(P, beta_fn(P.circle * cur.bgeo * cb.density, beta)) * trim1
                                                        : clamp_vm(P, beta_fn(P.circle / cur.w0, beta));
                      const float cosine = fabsf(dot(n, cur.s.normal()));
                      const float KD = cur.w2 * beta_fn(cb.density, beta) + float(cb.finite) * eye_vm * beta_fn(cur.w0, beta) * trim1;
                      const float KC = cur.w1 * beta_fn(cb.density, beta) + beta_fn(cur.w0, beta) * float(cur.finite());
                      const float3 result = ls.radiance / ls.light_density * P.roulette_inv * cur.thr * cb.throughput * cosine;
                      const float3 org = offset_origin(cur.s.pos, cur.s.gnormal, -n);
                      const unsigned dq = atomicAdd(&cnt->dir_count[parity], 1u);
                      if (dq < pool.pixels_pad) {
                        float4* e = pool.dir_q + size_t(dq) * 4;
                        e[0] = make_float4(org.x, org.y, org.z, __uint_as_float(slot));
                        e[1] = make_float4(cur.s.pos.x, cur.s.pos.y, cur.s.pos.z, cosine);
                        e[2] = make_float4(result.x, result.y, result.z, KC);
                        e[3] = make_float4(__uint_as_float(sc.materials[ls.surface.material_index()].light_id), KD, 0.f, 0.f);
                      } else {
                        atomicAdd(&cnt->capacity_events, 1ull);
                      }
                    }
                  }
                  emit_n = (count > 0 ? count - 1u : 0u) + (emit_nee ? 1u : 0u);
                  emit_skip_first = true;
                }
                O = offset_origin(cur.s.pos, cur.s.gnormal, bs.omega);
                want_trace = true;
              }
            }
          }
        }
      }
      if (store_cur) store_slot_vertex(pool.cv, pool.pixels_pad, slot, cur);
      if (finished) {
        pool.rad[slot] = make_float4(rad.x, rad.y, rad.z, 0.f);
        st = kDone;
      } else {
        pool.rng[slot] = make_uint2(rng.key, rng.ctr);
        if (st == kEyeWait) pool.rad[slot] = make_float4(rad.x, rad.y, rad.z, 0.f);
        if (want_trace) {
          pool.bs_omega[slot] = make_float4(bs.omega.x, bs.omega.y, bs.omega.z, bs.density);
          pool.bs_thr[slot] = make_float4(bs.throughput.x, bs.throughput.y, bs.throughput.z, bs.density_rev);
          pool.ray_o[slot] = make_float4(O.x, O.y, O.z, __uint_as_float(bs.finite & 1u));
        }
      }
      pool.state[slot] = st;
    }
    {
      const unsigned mask = __ballot_sync(0xFFFFFFFFu, finished);
      if (mask && (threadIdx.x & 31u) == 0) atomicAdd(&cnt->done, unsigned(__popc(mask)));
    }
    const unsigned tq = warp_append(&cnt->trace_count[parity], want_trace);
    if (want_trace) next_q[tq] = slot;
    const unsigned gq = warp_append(&cnt->gather_count[parity], want_gather);
    if (want_gather) pool.gather_q[gq] = slot;
    unsigned pq = warp_reserve(&cnt->pair_count[parity], emit_n);
    if (emit_n) {
      if (pq + emit_n > pool.q_capacity) {
        atomicAdd(&cnt->capacity_events, 1ull);
      } else {
        if (emit_nee) pool.pair_q[pq++] = make_uint2(slot, kNeeVertex);
        for (uint32_t v = head; v != kInvalid;) {
          const uint32_t prev = pool.lv_link[v];
          if (!(emit_skip_first && prev == kInvalid)) pool.pair_q[pq++] = make_uint2(slot | emit_flags, v);
          v = prev;
        }
      }
    }
  }
}

// ---------------------------------------------------------------- connection stage (one thread per pair)
__global__ void __launch_bounds__(256) vcm_connect_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt, int parity) {
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = min(cnt->pair_count[parity], pool.q_capacity);
  const uint32_t rounds = (n + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
  for (uint32_t r = 0; r < rounds; ++r) {
    const uint32_t i = (r * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
    bool want = false;
    float3 a = f3(0.f), d = f3(0.f), c = f3(0.f);
    uint32_t target = 0, flags = 0;
    if (i < n) {
      const uint2 desc = pool.pair_q[i];
      const uint32_t slot = desc.x & ~kPairCamera;
      const bool camera = (desc.x & kPairCamera) != 0u;
      const Vertex eye = load_slot_vertex(pool.cv, pool.pixels_pad, slot);
      Vertex light;
      if (desc.y == kNeeVertex) {
        const float4 nr = pool.nee[slot];
        light = sample_to_vertex(light_sample_at(sc, __float_as_uint(nr.w), f3(nr.x, nr.y, nr.z)), P);
      } else {
        light = load_pool_vertex(pool, desc.y);
      }
      bool inside = true;
      if (camera) {  // Technique::_accumulate (Technique.cpp:276-305)
        const float3 omega = normalize(light.s.pos - eye.s.pos);
        mat3 w2v;
        const float* m = cfg.params.camera.world_to_view;
        w2v.c0 = f3(m[0], m[1], m[2]); w2v.c1 = f3(m[3], m[4], m[5]); w2v.c2 = f3(m[6], m[7], m[8]);
        const float2 pp = pixel_position(mul(w2v, omega), cfg.width_f, cfg.height_f, cfg.res_y_inv, cfg.params.camera.focal_length_y);
        inside = (0 <= pp.x) && (pp.x < cfg.width_f) && (0 <= pp.y) && (pp.y < cfg.height_f);
        if (inside) {
          target = uint32_t(int(pp.y)) * cfg.params.width + uint32_t(int(pp.x));
          flags = kShadowSplat;
          const float cc = camera_coefficient(light.omega, light.s.gnormal, light.s.normal(), omega, eye.s.normal());
          c = connect_vertices(sc, light, eye, P) * cfg.focal_factor_y * cc;
        }
      } else {
        target = slot;
        c = connect_vertices(sc, light, eye, P);
      }
      if (inside) {
        const float l1 = l1norm(c);
        if (!isfinite(l1)) {
          // a non-finite connection poisons the sample whether or not it is visible (0 * inf); _commit_images drops it
          if (flags & kShadowSplat) atomicAdd(pool.splat + size_t(target) * 3, double(NAN));
          else atomicAdd(&pool.rad[target].x, NAN);
        } else if (l1 != 0.0f) {
          // Scene::occluded(eye.surface, light.surface) (Scene.cpp:151-171)
          const float3 direction = normalize(light.s.pos - eye.s.pos);
          a = eye.s.pos + (dot(eye.s.gnormal, direction) > 0.0f ? 1.0f : -1.0f) * eye.s.gnormal * 0.0001f;
          const float3 b = light.s.pos + (dot(light.s.gnormal, direction) < 0.0f ? 1.0f : -1.0f) * light.s.gnormal * 0.0001f;
          d = b - a;
          want = true;
        }
      }
    }
    const unsigned sq = warp_append(&cnt->shadow_count[parity], want);
    if (want) {
      if (sq < pool.q_capacity) {
        float4* e = pool.shadow_q + size_t(sq) * 3;
        e[0] = make_float4(a.x, a.y, a.z, __uint_as_float(target));
        e[1] = make_float4(d.x, d.y, d.z, __uint_as_float(flags));
        e[2] = make_float4(c.x, c.y, c.z, 0.f);
      } else {
        atomicAdd(&cnt->capacity_events, 1ull);
      }
    }
  }
}

__global__ void __launch_bounds__(256) vcm_trace_kernel(RenderConfig cfg, VcmPool pool, VcmCounters* cnt, int parity) {
  const uint32_t n = cnt->trace_count[parity];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cnt->basic_rays, (unsigned long long)n);
    cnt->trace_count[parity ^ 1] = 0;  // consumed by this iteration's logic stage
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t slot = pool.trace_q[parity][i];
    const float4 o = pool.ray_o[slot], d = pool.bs_omega[slot];
    const Hit h = trace_ray<false, false>(cfg.scene, f3(o.x, o.y, o.z), f3(d.x, d.y, d.z), 0.0f, INFINITY);  // Scene::intersect
    pool.hit[slot] = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri));
  }
}

__global__ void __launch_bounds__(256) vcm_shadow_kernel(RenderConfig cfg, VcmPool pool, VcmCounters* cnt, int parity) {
  const uint32_t n = min(cnt->shadow_count[parity], pool.q_capacity);
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->shadow_rays, (unsigned long long)n);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4* e = pool.shadow_q + size_t(i) * 3;
    const float4 o = e[0], d = e[1], c = e[2];
    const uint32_t target = __float_as_uint(o.w), flags = __float_as_uint(d.w);
    const Hit h = trace_ray<true, true>(cfg.scene, f3(o.x, o.y, o.z), f3(d.x, d.y, d.z), 0.0f, 1.0f);
    if (h.tri == kInvalid) {
      if (flags & kShadowSplat) {
        double* s = pool.splat + size_t(target) * 3;
        atomicAdd(s + 0, double(c.x)); atomicAdd(s + 1, double(c.y)); atomicAdd(s + 2, double(c.z));
      } else {
        float* r = &pool.rad[target].x;
        atomicAdd(r + 0, c.x); atomicAdd(r + 1, c.y); atomicAdd(r + 2, c.z);
      }
    }
  }
}

// the any-hit stage through trace_queue8: entry i of the shadow queue is ray i, an unoccluded one adds its contribution
struct VcmShadowIo {
  const float4* shadow_q;
  double* splat; float4* rad;
  __device__ __forceinline__ uint32_t load(uint32_t i, float3& O, float3& D, float& tnear, float& tfar, uint32_t& m) const {
    const float4* e = shadow_q + size_t(i) * 3;
    const float4 o = e[0], d = e[1];
    O = f3(o.x, o.y, o.z); D = f3(d.x, d.y, d.z); tnear = 0.0f; tfar = 1.0f; m = kMaskMesh;
    return i;
  }
  __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
    if (h.tri != kInvalid) return;
    const float4* e = shadow_q + size_t(i) * 3;
    const float4 c = e[2];
    const uint32_t target = __float_as_uint(e[0].w), flags = __float_as_uint(e[1].w);
    if (flags & kShadowSplat) {
      double* s = splat + size_t(target) * 3;
      atomicAdd(s + 0, double(c.x)); atomicAdd(s + 1, double(c.y)); atomicAdd(s + 2, double(c.z));
    } else {
      float* r = &rad[target].x;
      atomicAdd(r + 0, c.x); atomicAdd(r + 1, c.y); atomicAdd(r + 2, c.z);
    }
  }
};
__global__ void __launch_bounds__(256, 4) vcm_shadow_queue_kernel(RenderConfig cfg, VcmPool pool, VcmCounters* cnt, int parity) {
  const uint32_t n = min(cnt->shadow_count[parity], pool.q_capacity);
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cnt->shadow_rays, (unsigned long long)n);
  VcmShadowIo io{pool.shadow_q, pool.splat, pool.rad};
  trace_queue8<true>(cfg.scene, n, &cnt->shadow_ticket, io);
}

// ---------------------------------------------------------------- merging stage
// Intersector::intersectFast (Intersector.cpp:17-46): bounded mesh-only closest hit, then "hit point inside the sphere"
__device__ __forceinline__ bool intersect_fast(const DeviceScene& sc, const Surface& s, float3 direction, float target_length, float3 world_center,
                                               float radius) {
  const float3 org = offset_origin(s.pos, s.gnormal, direction);
  const Hit h = trace_ray<false, true>(sc, org, direction, 0.0f, target_length + radius);
  const float3 isect = org + direction * h.t;
  return h.tri != kInvalid && distance2(world_center, isect) < radius * radius;
}

// BSDF::gathering_density of the surface's material (DiffuseBSDF BSDF.cpp:264-289, PhongBSDF :393-434): the number of
// bounded-lobe samples until one lands inside the merge sphere, over the lobe restriction factor.  Kept as a resumable
// loop (begin / step) so that the deferred form below can run one rejection sample per lane per round.
struct DensityLoop {
  Surface s;
  float3 center, omega, target;
  float target_length, N, diffuse_probability, combined_adjust, power;
  // The restricted lobes of the loop: sample_lambert(generator, sphere, omega) and sample_phong(generator, omega, power, sphere)
  // (Sample.inl:58-82, 151-205) recompute these ranges in every call from arguments that do not change inside the loop — three asin,
  // an atan2, two cos (and two pow for the Phong lobe), all correctly rounded, per TRIAL.  The same functions on the same arguments
  // return the same bits, so they are evaluated once per loop here.  (angular_bound reads center.y only squared, so the sign flip of
  // Sample.inl:61 does not enter.)
  LambertRange lr;
  PhongRange pr;
  uint32_t type;
  uint32_t key;  // trial t draws from its own stream (key, t): the outcome of a trial does not depend on who runs it, or when
};
__device__ __forceinline__ void density_begin(DensityLoop& L, const DeviceScene& sc, const Rng& rng, const Surface& s, float3 target, float radius,
                                              float3 omega_world) {
  const B2ptMaterial& m = sc.materials[s.material_index()];
  L.s = s; L.target = target; L.key = rng.key;
  L.center = s.to_surface(target - s.pos);
  L.omega = s.to_surface(omega_world);
  L.target_length = length(L.center);
  L.N = 1.0f;
  L.type = m.type; L.power = m.power;
  L.diffuse_probability = 1.0f; L.combined_adjust = 1.0f;
  if (m.type == B2PT_BSDF_DIFFUSE || m.type == B2PT_BSDF_PHONG) L.lr = lambert_range(L.center, radius);
  if (m.type == B2PT_BSDF_PHONG) {
    L.pr = phong_range(L.omega, m.power, L.center, radius);
    const float diffuse_adjust = L.lr.theta_range * L.lr.phi_range;      // lambert_adjust (Sample.inl:119-131)
    const float specular_adjust = L.pr.theta_range * L.pr.phi_range;     // phong_adjust (Sample.inl:197-205)
    L.combined_adjust = diffuse_adjust * m.diffuse_probability + specular_adjust * (1.0f - m.diffuse_probability);
    L.diffuse_probability =
        diffuse_adjust * m.diffuse_probability / (diffuse_adjust * m.diffuse_probability + specular_adjust * (1.0f - m.diffuse_probability));
  }
}
// delta lobes never reach the estimate (their connection throughput is zero); anything else the reference answers with infinity
__device__ __forceinline__ bool density_has_lobe(const DensityLoop& L) { return L.type == B2PT_BSDF_DIFFUSE || L.type == B2PT_BSDF_PHONG; }
// trial number t (1, 2, ...): one bounded-lobe sample and one bounded ray; true if it lands in the merge sphere (density = t / adjust)
__device__ __forceinline__ bool density_trial(const DensityLoop& L, const DeviceScene& sc, float radius, float t, float& density) {
  Rng r; r.key = mix32(L.key ^ mix32(uint32_t(t) * 0x9E3779B9u + 0x85EBCA6Bu)); r.ctr = 0;
  float3 d;
  float adjust = 1.0f;
  // the two sampling functions with their ranges taken from the loop; ONE polar conversion after the branch (the draws, the operations and
  // their order are those of sample_lambert_bounded / sample_phong_bounded in shading.cuh)
  const bool lambert = L.type == B2PT_BSDF_DIFFUSE || r.sample() < L.diffuse_probability;
  float y, phi;
  if (lambert) {
    y = sqrtf(r.sample() * L.lr.theta_range + L.lr.theta_inf) * sign_glm(L.omega.y);
    phi = kTwoPi * (r.sample() * L.lr.phi_range + L.lr.phi_inf);
  } else {
    y = cr_powf(r.sample() * L.pr.theta_range + L.pr.theta_inf, 1.0f / (L.power + 1.0f));
    phi = kTwoPi * (r.sample() * L.pr.phi_range + L.pr.phi_inf);
  }
  const float rr = sqrtf(1 - y * y);
  d = cr_polar(rr, y, phi);
  if (!lambert) d = mul(L.pr.reflection, d);
  adjust = L.type == B2PT_BSDF_DIFFUSE ? L.lr.theta_range * L.lr.phi_range : L.combined_adjust;
  density = t / adjust;
  return intersect_fast(sc, L.s, L.s.to_world(d), L.target_length, L.target, radius);
}
// one rejection sample; true when the loop is over (density holds the result)
__device__ __forceinline__ bool density_step(DensityLoop& L, const DeviceScene& sc, float radius, float& density, unsigned long long& rays) {
  if (!density_has_lobe(L) || !(L.N < kDensityTestLimit)) { density = INFINITY; return true; }  // DENSITY_ESTIMATION_TEST_LIMIT, BSDF.cpp:9
  ++rays;
  if (density_trial(L, sc, radius, L.N, density)) return true;
  L.N += 1.0f;
  return false;
}
__device__ float gathering_density(const DeviceScene& sc, Rng& rng, const Surface& s, float3 target, float radius, float3 omega_world,
                                   unsigned long long& rays) {
  DensityLoop L;
  density_begin(L, sc, rng, s, target, radius, omega_world);
  float density = INFINITY;
  while (!density_step(L, sc, radius, density, rays)) {}
  return density;
}

// _merge_biased (UPG.cpp:783-817) with light = _light_paths[index - 1], tentative = _light_paths[index]
__device__ __forceinline__ float3 merge_biased(const DeviceScene& sc, const Vertex& light, const Vertex& tentative, const Vertex& eye, const VcmParams& P) {
  const float3 omega = normalize(eye.s.pos - light.s.pos);
  const Connection c(sc, light, eye, omega);
  const float3 throughput = tentative.thr * eye.thr * c.eb.throughput * P.roulette;
  if (l1norm(throughput) < kFltEpsilon) return f3(0.0f);
  const float vm_current = beta_fn(P.circle / tentative.w0, P.beta);
  // _vm_biased_weight (UPG.cpp:405-410)
  const float weight = beta_fn(P.num_scattered, P.beta) * vm_current * vc_biased_weight(c, P, vm_current * (eye.length() == 0.0f ? 0.0f : 1.0f));
  const float density = 1.0f / P.circle;
  return throughput * density * weight;
}

// _merge_light (UPG.cpp:727-753), unbiased mode
// `deferred` != nullptr: return unoccluded * weight and leave the density to upg_density_kernel (*deferred = true when there is one to estimate)
__device__ __forceinline__ float3 merge_light(const DeviceScene& sc, Rng& rng, const Vertex& light, const Vertex& eye, const VcmParams& P,
                                              unsigned long long& shadow_rays, unsigned long long& density_rays, bool* deferred = nullptr) {
  const float3 omega = normalize(eye.s.pos - light.s.pos);
  const Connection c(sc, light, eye, omega);
  const float3 unoccluded = c.unoccluded();
  if (!(l1norm(unoccluded) >= kFltEpsilon)) return f3(0.0f);
  // Scene::occluded(eye.surface, light.surface)
  const float3 direction = normalize(light.s.pos - eye.s.pos);
  const float3 a = eye.s.pos + (dot(eye.s.gnormal, direction) > 0.0f ? 1.0f : -1.0f) * eye.s.gnormal * 0.0001f;
  const float3 b = light.s.pos + (dot(light.s.gnormal, direction) < 0.0f ? 1.0f : -1.0f) * light.s.gnormal * 0.0001f;
  ++shadow_rays;
  if (trace_ray<true, true>(sc, a, b - a, 0.0f, 1.0f).tri != kInvalid) return f3(0.0f);
  // _weight_vm_light (UPG.cpp:420-426)
  const float weight = beta_fn(P.num_scattered * clamp_vm(P, P.circle * c.edge.f_geometry * c.lb.density), P.beta) * vc_weight(c, P);
  if (deferred) { *deferred = true; return unoccluded * weight; }
  const float density = gathering_density(sc, rng, light.s, eye.s.pos, P.radius, light.omega, density_rays);
  return unoccluded * density * weight;
}

// _merge_eye (UPG.cpp:755-781), unbiased mode, merging from the camera side: the density is estimated from the eye vertex
__device__ __forceinline__ float3 merge_eye(const DeviceScene& sc, Rng& rng, const Vertex& light, const Vertex& eye, const VcmParams& P,
                                            unsigned long long& shadow_rays, unsigned long long& density_rays, bool* deferred = nullptr) {
  const float3 omega = normalize(eye.s.pos - light.s.pos);
  const Connection c(sc, light, eye, omega);
  const float3 unoccluded = c.unoccluded();
  if (!(l1norm(unoccluded) >= kFltEpsilon)) return f3(0.0f);
  const float3 direction = normalize(light.s.pos - eye.s.pos);
  const float3 a = eye.s.pos + (dot(eye.s.gnormal, direction) > 0.0f ? 1.0f : -1.0f) * eye.s.gnormal * 0.0001f;
  const float3 b = light.s.pos + (dot(light.s.gnormal, direction) < 0.0f ? 1.0f : -1.0f) * light.s.gnormal * 0.0001f;
  ++shadow_rays;
  if (trace_ray<true, true>(sc, a, b - a, 0.0f, 1.0f).tri != kInvalid) return f3(0.0f);
  // _weight_vm_eye (UPG.cpp:412-418)
  const float weight = beta_fn(P.num_scattered * clamp_vm(P, P.circle * c.edge.b_geometry * c.eb.density_rev), P.beta) * vc_weight(c, P);
  if (deferred) { *deferred = true; return unoccluded * weight; }
  const float density = gathering_density(sc, rng, eye.s, light.s.pos, P.radius, eye.omega, density_rays);
  return unoccluded * density * weight;
}

// what upg_density_kernel needs to finish an unbiased merge: whose surface samples (a pool vertex, or the path's previous eye
// vertex), towards which point, with which random stream; vertex == kJobNone: nothing was deferred
constexpr uint32_t kJobNone = 0xFFFFFFFFu, kJobEyeVertex = 0xFFFFFFFEu;
struct DensityJob { uint32_t vertex = kJobNone; float3 target; uint32_t key = 0; };

// One merge of _gather / _gather_biased (UPG.cpp:684-693, 708-719): photon `index` against the eye path in `slot`.
// eye = the tentative vertex (throughput already * 1 / roulette), prev_eye = the vertex before it (camera-side merging only).
__device__ __forceinline__ float3 merge_one(const DeviceScene& sc, const VcmParams& P, const VcmPool& pool, uint32_t index, const Vertex& eye,
                                            const Vertex& prev_eye, uint32_t path_key, unsigned long long& shadow_rays, unsigned long long& density_rays,
                                            DensityJob* job = nullptr) {
  if (!P.from_light) {  // UPG.cpp:690-692, 716-718: the photon itself against the eye vertex before the tentative one
    const Vertex light = load_pool_vertex(pool, index);
    if (P.unbiased) {
      Rng sub; sub.key = mix32(path_key ^ mix32(index * 0x9E3779B9u + uint32_t(eye.length()))); sub.ctr = 0;
      bool deferred = false;
      const float3 v = merge_eye(sc, sub, light, prev_eye, P, shadow_rays, density_rays, job ? &deferred : nullptr) * P.num_scattered_inv;
      if (deferred) { job->vertex = kJobEyeVertex; job->target = light.s.pos; job->key = sub.key; }
      return v;
    }
    return merge_biased(sc, light, light, prev_eye, P) * P.num_scattered_inv;
  }
  const uint32_t prev = pool.lv_link[index];
  if (prev == kInvalid) return f3(0.0f);
  const Vertex light = load_pool_vertex(pool, prev);
  if (light.s.is_light()) return f3(0.0f);  // UPG.cpp:689,713: the predecessor must be a surface vertex
  if (P.unbiased) {
    Rng sub; sub.key = mix32(path_key ^ mix32(index * 0x9E3779B9u + uint32_t(eye.length()))); sub.ctr = 0;
    bool deferred = false;
    const float3 v = merge_light(sc, sub, light, eye, P, shadow_rays, density_rays, job ? &deferred : nullptr) * P.num_scattered_inv;
    if (deferred) { job->vertex = prev; job->target = eye.s.pos; job->key = sub.key; }
    return v;
  }
  const Vertex tentative = load_pool_vertex(pool, index);
  return merge_biased(sc, light, tentative, eye, P) * P.num_scattered_inv;
}

// _gather / _gather_biased (UPG.cpp:675-724) in two stages.
//
// vcm_gather_kernel: one thread per new eye vertex runs the range query (rQuery, HashGrid3D.hpp:181-204) TWICE — once to
// count the photons inside the radius, once to write them as (eye slot, photon index) pairs into a range of the merge queue
// reserved with one atomic per warp — and evaluates nothing.  vcm_merge_kernel: one thread per pair evaluates the merge.
// A query finds anything between 0 and hundreds of photons, and a merge is two BSDF queries plus the MIS weights (UPG also a
// shadow ray and the density-estimation loop): evaluated inside the query loop, a warp runs as long as its longest query with
// whatever lanes still have photons left; as pairs, every lane always has one.  Queries that do not fit the queue any more
// (m_capacity) evaluate their merges in place, so the result never depends on the capacity.
constexpr uint32_t kFoundMax = 48;  // photons per query kept from the counting pass (queries that find more walk the cells a second time)
__global__ void __launch_bounds__(256) vcm_gather_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, DeviceGrid grid, VcmCounters* cnt, int parity) {
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = cnt->gather_count[parity];
  const uint32_t n_round = (n + 31u) & ~31u;
  const float radius_sq = P.radius * P.radius;
  unsigned long long shadow_rays = 0, density_rays = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
    const bool on = i < n;
    const uint32_t slot = on ? pool.gather_q[i] : 0u;
    const float4 p0 = pool.cv[size_t(slot) * kSlotVertexFloat4s];  // float4 0 of the tentative vertex: its position
    const float3 q = f3(p0.x, p0.y, p0.z);
    uint32_t count = 0;
    uint32_t found[kFoundMax];
    if (on) {
      if (P.brute_vertices) {  // diagnostic: every stored vertex instead of the grid
        for (uint32_t k = 0; k < P.brute_vertices; ++k) {
          const float4 g = pool.grid_points[k];
          if (__float_as_uint(g.w) != kInvalidPoint && distance2(q, f3(g.x, g.y, g.z)) < radius_sq) ++count;
        }
      } else {
        // the first kFoundMax photons are remembered, so that the usual query walks its 27 cells once, not twice
        grid_query(grid, q, radius_sq, [&](uint32_t index, float3) { if (count < kFoundMax) found[count] = index; ++count; });
      }
    }
    unsigned base = warp_reserve(&cnt->merge_count, count);
    if (count == 0) continue;
    if (base + count <= pool.m_capacity && !P.brute_vertices) {
      if (count <= kFoundMax) {
        for (uint32_t k = 0; k < count; ++k) pool.merge_q[base + k] = make_uint2(slot, found[k]);
      } else {
        grid_query(grid, q, radius_sq, [&](uint32_t index, float3) { pool.merge_q[base++] = make_uint2(slot, index); });
      }
    } else {
      // no room (or the diagnostic mode): blank whatever part of the reservation lies inside the queue, merge in place
      for (unsigned k = base; k < base + count && k < pool.m_capacity; ++k) pool.merge_q[k] = make_uint2(kInvalid, 0u);
      const Vertex eye = load_slot_vertex(pool.cv, pool.pixels_pad, slot);
      Vertex prev_eye = eye;
      if (!P.from_light) prev_eye = load_slot_vertex(pool.pv, pool.pixels_pad, slot);
      const uint32_t path_key = pool.rng[slot].x;
      float3 sum = f3(0.0f);
      auto merge = [&](uint32_t index, float3) { sum += merge_one(sc, P, pool, index, eye, prev_eye, path_key, shadow_rays, density_rays); };
      if (P.brute_vertices) {
        for (uint32_t k = 0; k < P.brute_vertices; ++k) {
          const float4 g = pool.grid_points[k];
          if (__float_as_uint(g.w) != kInvalidPoint && distance2(q, f3(g.x, g.y, g.z)) < radius_sq) merge(k, f3(g.x, g.y, g.z));
        }
      } else {
        grid_query(grid, q, radius_sq, merge);
      }
      if (sum.x != 0.0f || sum.y != 0.0f || sum.z != 0.0f) {
        float* r = &pool.rad[slot].x;
        atomicAdd(r + 0, sum.x); atomicAdd(r + 1, sum.y); atomicAdd(r + 2, sum.z);
      }
    }
  }
  if (shadow_rays) atomicAdd(&cnt->shadow_rays, shadow_rays);
  if (density_rays) { atomicAdd(&cnt->density_rays, density_rays); atomicAdd(&cnt->basic_rays, density_rays); }
}

__global__ void __launch_bounds__(256) vcm_merge_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt) {
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = min(cnt->merge_count, pool.m_capacity);
  unsigned long long shadow_rays = 0, density_rays = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint2 e = pool.merge_q[i];
    if (e.x == kInvalid) continue;
    const uint32_t slot = e.x;
    const Vertex eye = load_slot_vertex(pool.cv, pool.pixels_pad, slot);
    Vertex prev_eye = eye;
    if (!P.from_light) prev_eye = load_slot_vertex(pool.pv, pool.pixels_pad, slot);
    DensityJob job;
    const float3 v = merge_one(sc, P, pool, e.y, eye, prev_eye, pool.rng[slot].x, shadow_rays, density_rays, P.unbiased ? &job : nullptr);
    if (job.vertex != kJobNone) {
      // UPG: the density estimate is a rejection loop of unpredictable length — its own stage (upg_density_kernel)
      const unsigned j = warp_append(&cnt->job_count, true);
      float4* o = pool.job_q + size_t(j) * 3;  // job_count <= merge_count <= m_capacity
      o[0] = make_float4(v.x, v.y, v.z, __uint_as_float(slot));
      o[1] = make_float4(job.target.x, job.target.y, job.target.z, __uint_as_float(job.vertex));
      o[2] = make_float4(__uint_as_float(job.key), 0.f, 0.f, 0.f);
    } else if (v.x != 0.0f || v.y != 0.0f || v.z != 0.0f) {
      float* r = &pool.rad[slot].x;
      atomicAdd(r + 0, v.x); atomicAdd(r + 1, v.y); atomicAdd(r + 2, v.z);
    }
  }
  if (shadow_rays) atomicAdd(&cnt->shadow_rays, shadow_rays);
  if (density_rays) { atomicAdd(&cnt->density_rays, density_rays); atomicAdd(&cnt->basic_rays, density_rays); }
}

constexpr float kLaneTrials = 16.0f;  // trials a lane runs on its own before the loop goes to a whole warp
constexpr float kWarpTrials = 16384.0f;  // trials a warp runs before the loop goes to the whole grid
constexpr uint32_t kWideCap = 4096;      // loops the grid stage takes per eye iteration (the rest stay with their warp)
constexpr int kWideChunk = 128;          // trials per ticket in the grid stage

// The unbiased merge density (BSDF::gathering_density, BSDF.cpp:42-66): sample the bounded lobe, cast a bounded ray, count the
// tries until one lands in the merge sphere — a geometric number of iterations, a handful on average, 2^24 at most.  Persistent
// warps: every round runs ONE trial for all lanes that hold a job, a warp refills its idle lanes from the job queue once a quarter of
// them are idle, and a loop that has not ended after kLaneTrials trials moves to upg_density_long_kernel (measured on the LivingRoom
// stand-in: 60 density rays per pixel-sample in total, but single loops of tens of thousands of trials, which used to set the duration
// of the whole launch).
__global__ void __launch_bounds__(256, 2) upg_density_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt) {
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = min(cnt->job_count, pool.m_capacity);
  const unsigned lane = threadIdx.x & 31u, lane_lt = (1u << lane) - 1u;
  unsigned long long density_rays = 0;
  DensityLoop L;
  float3 coef = f3(0.0f);
  uint32_t slot = 0, job = 0;
  bool has = false, exhausted = false;
  for (;;) {
    __syncwarp();
    const unsigned idle = __ballot_sync(0xFFFFFFFFu, !has);
    if (!exhausted && (__popc(idle) >= 8 || idle == 0xFFFFFFFFu)) {
      const unsigned want = __popc(idle);
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(&cnt->job_ticket, want);
      base = __shfl_sync(0xFFFFFFFFu, base, 0);
      exhausted = base + want >= n;
      const uint32_t i = base + __popc(idle & lane_lt);
      if (!has && i < n) {
        job = i;
        const float4* o = pool.job_q + size_t(i) * 3;
        const float4 a = o[0], b = o[1], c = o[2];
        coef = f3(a.x, a.y, a.z); slot = __float_as_uint(a.w);
        const uint32_t vertex = __float_as_uint(b.w);
        const Vertex v = vertex == kJobEyeVertex ? load_slot_vertex(pool.pv, pool.pixels_pad, slot) : load_pool_vertex(pool, vertex);
        Rng sub; sub.key = __float_as_uint(c.x); sub.ctr = 0;
        density_begin(L, sc, sub, v.s, f3(b.x, b.y, b.z), P.radius, v.omega);
        has = true;
      }
    }
    if (__ballot_sync(0xFFFFFFFFu, has) == 0u) break;
    if (has) {
      float density;
      if (density_step(L, sc, P.radius, density, density_rays)) {
        const float3 v = coef * density;
        float* r = &pool.rad[slot].x;
        atomicAdd(r + 0, v.x); atomicAdd(r + 1, v.y); atomicAdd(r + 2, v.z);
        has = false;
      } else if (L.N > kLaneTrials) {
        // a long loop (the sphere is mostly hidden from this surface): a whole warp takes it over, 32 trials per round
        const unsigned j = atomicAdd(&cnt->long_count, 1u);
        pool.merge_q[j] = make_uint2(job, uint32_t(L.N));  // the merge queue is free again: vcm_merge_kernel has consumed it
        has = false;
      }
    }
  }
  if (density_rays) { atomicAdd(&cnt->density_rays, density_rays); atomicAdd(&cnt->basic_rays, density_rays); }
}

// Long density loops, one WARP per loop: the trials are independent (trial t has its own stream), so 32 of them run per round and
// the first one that lands is the answer — the same estimate in 1/32 of the rounds.  Only the trials up to that one count as rays.
__global__ void __launch_bounds__(256, 2) upg_density_long_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt) {
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = min(cnt->long_count, pool.m_capacity);
  const unsigned lane = threadIdx.x & 31u;
  unsigned long long density_rays = 0;
  for (;;) {
    unsigned j = 0;
    if (lane == 0) j = atomicAdd(&cnt->long_ticket, 1u);
    j = __shfl_sync(0xFFFFFFFFu, j, 0);
    if (j >= n) break;
    const uint2 e = pool.merge_q[j];
    const float4* o = pool.job_q + size_t(e.x) * 3;
    const float4 a = o[0], b = o[1], c = o[2];
    const uint32_t slot = __float_as_uint(a.w), vertex = __float_as_uint(b.w);
    const Vertex v = vertex == kJobEyeVertex ? load_slot_vertex(pool.pv, pool.pixels_pad, slot) : load_pool_vertex(pool, vertex);
    Rng sub; sub.key = __float_as_uint(c.x); sub.ctr = 0;
    DensityLoop L;
    density_begin(L, sc, sub, v.s, f3(b.x, b.y, b.z), P.radius, v.omega);
    float N = float(e.y), density = INFINITY;
    const float N_wide = N + kWarpTrials;
    bool handed_over = false;
    for (;;) {
      if (N >= N_wide) {
        // still nothing after kWarpTrials trials: most of these run to the test limit (the sphere is hidden from the surface) — 2^24 trials
        // are two seconds for one warp and a millisecond for the grid
        unsigned h = 0;
        if (lane == 0) h = atomicAdd(&cnt->wide_count, 1u);
        h = __shfl_sync(0xFFFFFFFFu, h, 0);
        if (h < kWideCap) { if (lane == 0) pool.wide_q[h] = make_uint2(e.x, uint32_t(N)); handed_over = true; break; }
      }
      const float t = N + float(lane);
      float d = INFINITY;
      const bool ok = t < kDensityTestLimit && density_trial(L, sc, P.radius, t, d);
      const unsigned hit = __ballot_sync(0xFFFFFFFFu, ok);
      if (hit) {
        const int first = __ffs(int(hit)) - 1;
        density = __shfl_sync(0xFFFFFFFFu, d, first);
        if (lane == 0) density_rays += unsigned(first) + 1u;
        break;
      }
      if (lane == 0) density_rays += 32u;
      N += 32.0f;
      if (!(N < kDensityTestLimit)) break;  // DENSITY_ESTIMATION_TEST_LIMIT: the reference answers infinity
    }
    if (lane == 0 && !handed_over) {
      const float3 r3 = f3(a.x, a.y, a.z) * density;
      float* r = &pool.rad[slot].x;
      atomicAdd(r + 0, r3.x); atomicAdd(r + 1, r3.y); atomicAdd(r + 2, r3.z);
    }
  }
  if (density_rays) { atomicAdd(&cnt->density_rays, density_rays); atomicAdd(&cnt->basic_rays, density_rays); }
}

// The longest density loops, one GRID per loop: warps draw chunks of kWideChunk trials from the loop's ticket until a chunk starts past
// the first trial known to have landed (atomicMin) or past the test limit.  The answer is the smallest landing trial, as in the serial loop.
__global__ void __launch_bounds__(256, 2) upg_density_wide_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt) {
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = min(cnt->wide_count, kWideCap);
  const unsigned lane = threadIdx.x & 31u;
  for (uint32_t h = 0; h < n; ++h) {
    const uint2 e = pool.wide_q[h];
    const float4* o = pool.job_q + size_t(e.x) * 3;
    const float4 a = o[0], b = o[1], c = o[2];
    const uint32_t slot = __float_as_uint(a.w), vertex = __float_as_uint(b.w);
    const Vertex v = vertex == kJobEyeVertex ? load_slot_vertex(pool.pv, pool.pixels_pad, slot) : load_pool_vertex(pool, vertex);
    Rng sub; sub.key = __float_as_uint(c.x); sub.ctr = 0;
    DensityLoop L;
    density_begin(L, sc, sub, v.s, f3(b.x, b.y, b.z), P.radius, v.omega);
    unsigned int* best = &pool.wide_state[h].x;
    unsigned int* ticket = &pool.wide_state[h].y;
    for (;;) {
      unsigned chunk = 0;
      if (lane == 0) chunk = atomicAdd(ticket, 1u);
      chunk = __shfl_sync(0xFFFFFFFFu, chunk, 0);
      const float base = float(e.y) + float(chunk) * float(kWideChunk);
      if (!(base < kDensityTestLimit) || base > float(*(volatile unsigned int*)best)) break;
      bool found = false;
      for (int k = 0; k < kWideChunk / 32 && !found; ++k) {
        const float t = base + float(k * 32) + float(lane);
        float d;
        const bool ok = t < kDensityTestLimit && density_trial(L, sc, P.radius, t, d);
        const unsigned hit = __ballot_sync(0xFFFFFFFFu, ok);
        if (hit) {
          if (lane == 0) atomicMin(best, uint32_t(base) + uint32_t(k * 32) + uint32_t(__ffs(int(hit)) - 1));
          found = true;
        }
      }
      if (found) break;
    }
  }
}
// ... and its second half, once every warp is done: one thread per loop adds the estimate and re-arms the loop's state
__global__ void __launch_bounds__(256) upg_density_wide_finish_kernel(RenderConfig cfg, VcmParams P, VcmPool pool, VcmCounters* cnt) {
  const DeviceScene& sc = cfg.scene;
  const uint32_t n = min(cnt->wide_count, kWideCap);
  for (uint32_t h = blockIdx.x * blockDim.x + threadIdx.x; h < n; h += gridDim.x * blockDim.x) {
    const uint2 e = pool.wide_q[h];
    const float4* o = pool.job_q + size_t(e.x) * 3;
    const float4 a = o[0], b = o[1], c = o[2];
    const uint32_t slot = __float_as_uint(a.w), vertex = __float_as_uint(b.w);
    const uint32_t best = pool.wide_state[h].x;
    pool.wide_state[h] = make_uint2(0xFFFFFFFFu, 0u);
    float density = INFINITY;  // DENSITY_ESTIMATION_TEST_LIMIT reached: the reference answers infinity
    unsigned long long rays = (unsigned long long)(kDensityTestLimit) - e.y;
    if (best != 0xFFFFFFFFu) {
      const Vertex v = vertex == kJobEyeVertex ? load_slot_vertex(pool.pv, pool.pixels_pad, slot) : load_pool_vertex(pool, vertex);
      Rng sub; sub.key = __float_as_uint(c.x); sub.ctr = 0;
      DensityLoop L;
      density_begin(L, sc, sub, v.s, f3(b.x, b.y, b.z), P.radius, v.omega);
      density_trial(L, sc, P.radius, float(best), density);
      rays = best - e.y + 1u;
    }
    const float3 r3 = f3(a.x, a.y, a.z) * density;
    float* r = &pool.rad[slot].x;
    atomicAdd(r + 0, r3.x); atomicAdd(r + 1, r3.y); atomicAdd(r + 2, r3.z);
    atomicAdd(&cnt->density_rays, rays); atomicAdd(&cnt->basic_rays, rays);
  }
}

// Technique::_commit_images (Technique.cpp:194-244) for one pass
__global__ void __launch_bounds__(256) vcm_commit_kernel(RenderConfig cfg, VcmPool pool, VcmCounters* cnt, uint32_t pass) {
  for (uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x; pixel < cfg.pixels; pixel += gridDim.x * blockDim.x) {
    const float4 e = pool.rad[pixel];
    double* s = pool.splat + size_t(pixel) * 3;
    const double r = s[0] + double(e.x), g = s[1] + double(e.y), b = s[2] + double(e.z);
    s[0] = 0.0; s[1] = 0.0; s[2] = 0.0;
    if (cfg.path_radiance) {
      float* o = cfg.path_radiance + (size_t(pass) * cfg.pixels + pixel) * 3;
      o[0] = e.x; o[1] = e.y; o[2] = e.z;
    }
    double* a = cfg.accum + size_t(pixel) * 4 + size_t(pass) * cfg.accum_pass_stride;
    if (isfinite(fabs(r) + fabs(g) + fabs(b))) {
      // passes in different lanes may commit the same pixel at the same time (render_vcm)
      atomicAdd(a + 0, r); atomicAdd(a + 1, g); atomicAdd(a + 2, b); atomicAdd(a + 3, 1.0);
    } else {
      atomicAdd(&cnt->numeric_errors, 1ull);
    }
  }
}

// ---------------------------------------------------------------- host driver
inline float beta_fn_host(float x, float beta) {  // Beta::beta on the host (Beta.hpp:24-42)
  if (beta == 2.0f) return x * x;
  if (beta == 1.0f) return x;
  if (beta == 0.0f) return x == 0.0f ? 0.0f : 1.0f;
  return powf(x, beta);
}

struct VcmWorkspace {
  VcmPool pool{};
  uint32_t pixels = 0, photons = 0;
  VcmCounters* counters = nullptr;
  VcmCounters* host_counters = nullptr;  // pinned
  GridStorage grid;
  std::vector<void*> blocks;
};

void vcm_workspace_free(void* p) {
  VcmWorkspace* w = (VcmWorkspace*)p;
  for (void* q : w->blocks) cudaFree(q);
  if (w->host_counters) cudaFreeHost(w->host_counters);
  w->grid.release();
  delete w;
}

template <class T>
int ws_alloc(VcmWorkspace* w, T** p, size_t count) {
  void* q = nullptr;
  B2_CUDA(cudaMalloc(&q, count * sizeof(T)));
  w->blocks.push_back(q);
  *p = (T*)q;
  return B2PT_OK;
}

// Passes are independent (each builds its own photon map), so two of them run at a time, each in its own workspace on its own
// stream, driven by its own host thread (render_vcm): one pass alone leaves the GPU to launches of a few hundred thousand rays.
constexpr int kVcmLanes = 8;       // most passes in flight (B2PT_VCM_LANES)
constexpr int kVcmDefaultLanes = 4;  // measured on the LivingRoom stand-in, 1024^2: 1 lane 15.2, 2: 22.8, 3: 25.8, 4: 29.1, 8: 29.6 Msamples/s (VCM)
struct VcmWorkspaceSet {
  VcmWorkspace* lane[kVcmLanes] = {};
  cudaStream_t stream[kVcmLanes] = {};  // lanes 1.. (lane 0 runs on the caller's stream)
  cudaEvent_t fork = nullptr;
  uint32_t pixels = 0, photons = 0, per_path = 0;
};
void vcm_workspace_set_free(void* p) {
  VcmWorkspaceSet* set = (VcmWorkspaceSet*)p;
  for (VcmWorkspace* w : set->lane) if (w) vcm_workspace_free(w);
  for (cudaStream_t st : set->stream) if (st) cudaStreamDestroy(st);
  if (set->fork) cudaEventDestroy(set->fork);
  delete set;
}

int vcm_workspace_fill(VcmWorkspace* w, uint32_t pixels, uint32_t photons, uint32_t per_path);

int vcm_workspace(b2pt_scene_s* scene, uint32_t pixels, uint32_t photons, uint32_t per_path, int* lanes_io, VcmWorkspaceSet** out) {
  int lanes = *lanes_io;
  VcmWorkspaceSet* set = (VcmWorkspaceSet*)scene->workspace;
  if (!(set && scene->workspace_free == vcm_workspace_set_free && set->pixels == pixels && set->photons == photons && set->per_path == per_path)) {
    if (scene->workspace && scene->workspace_free) scene->workspace_free(scene->workspace);
    scene->workspace = nullptr;
    set = new VcmWorkspaceSet();
    scene->workspace = set; scene->workspace_free = vcm_workspace_set_free;
    set->pixels = pixels; set->photons = photons; set->per_path = per_path;
  }
  for (int k = 0; k < lanes; ++k) {
    if (set->lane[k]) continue;
    set->lane[k] = new VcmWorkspace();
    if (int rc = vcm_workspace_fill(set->lane[k], pixels, photons, per_path)) {
      // not enough memory for another pass in flight: run with the lanes that fit
      vcm_workspace_free(set->lane[k]); set->lane[k] = nullptr;
      cudaGetLastError();
      if (k == 0) return rc;
      lanes = k;
      break;
    }
  }
  *lanes_io = lanes;
  for (int k = 1; k < lanes; ++k) if (!set->stream[k]) { B2_CUDA(cudaStreamCreateWithFlags(&set->stream[k], cudaStreamNonBlocking)); scene_l2_window(scene, set->stream[k]); }
  if (lanes > 1 && !set->fork) B2_CUDA(cudaEventCreateWithFlags(&set->fork, cudaEventDisableTiming));
  *out = set;
  return B2PT_OK;
}

int vcm_workspace_fill(VcmWorkspace* w, uint32_t pixels, uint32_t photons, uint32_t per_path) {
  w->pixels = pixels; w->photons = photons;
  const uint32_t block = 256;
  VcmPool& p = w->pool;
  p.photons = (photons + block - 1) / block * block;
  p.pixels_pad = (pixels + block - 1) / block * block;
  const size_t np = p.photons, ne = p.pixels_pad, nl = np > ne ? np : ne;
  p.lv_capacity = uint32_t(std::min<uint64_t>(std::max<uint64_t>(uint64_t(np) * per_path, 4096u), 0x7FFFFF00u));  // render_common.cuh: lv_per_path
  p.q_capacity = p.lv_capacity + uint32_t(ne);
  int rc;
  if ((rc = ws_alloc(w, &p.p_state, np))) return rc;
  if ((rc = ws_alloc(w, &p.p_rng, np))) return rc;
  if ((rc = ws_alloc(w, &p.p_cv, np * kSlotVertexFloat4s))) return rc;
  if ((rc = ws_alloc(w, &p.p_bs_omega, np))) return rc;
  if ((rc = ws_alloc(w, &p.p_bs_thr, np))) return rc;
  if ((rc = ws_alloc(w, &p.p_ray_o, np))) return rc;
  if ((rc = ws_alloc(w, &p.p_hit, np))) return rc;
  if ((rc = ws_alloc(w, &p.lp_head, nl))) return rc;
  if ((rc = ws_alloc(w, &p.lp_count, nl))) return rc;
  if ((rc = ws_alloc(w, &p.state, ne))) return rc;
  if ((rc = ws_alloc(w, &p.rng, ne))) return rc;
  if ((rc = ws_alloc(w, &p.cv, ne * kSlotVertexFloat4s))) return rc;
  if ((rc = ws_alloc(w, &p.pv, ne * kSlotVertexFloat4s))) return rc;
  if ((rc = ws_alloc(w, &p.bs_omega, ne))) return rc;
  if ((rc = ws_alloc(w, &p.bs_thr, ne))) return rc;
  if ((rc = ws_alloc(w, &p.ray_o, ne))) return rc;
  if ((rc = ws_alloc(w, &p.hit, ne))) return rc;
  if ((rc = ws_alloc(w, &p.rad, ne))) return rc;
  if ((rc = ws_alloc(w, &p.nee, ne))) return rc;
  if ((rc = ws_alloc(w, &p.splat, ne * 3))) return rc;
  if ((rc = ws_alloc(w, &p.lv, size_t(p.lv_capacity) * kVertexFloat4s))) return rc;
  if ((rc = ws_alloc(w, &p.lv_link, p.lv_capacity))) return rc;
  if ((rc = ws_alloc(w, &p.grid_points, p.lv_capacity))) return rc;
  if ((rc = ws_alloc(w, &p.trace_q[0], nl))) return rc;
  if ((rc = ws_alloc(w, &p.trace_q[1], nl))) return rc;
  if ((rc = ws_alloc(w, &p.pair_q, p.q_capacity))) return rc;
  if ((rc = ws_alloc(w, &p.shadow_q, size_t(p.q_capacity) * 3))) return rc;
  if ((rc = ws_alloc(w, &p.gather_q, ne))) return rc;
  // merge pairs: 64 per pixel (VCM at the thesis radius finds tens of photons per query); what does not fit is merged in place
  p.m_capacity = uint32_t(std::min<size_t>(ne * 64, size_t(1) << 28));
  if ((rc = ws_alloc(w, &p.merge_q, p.m_capacity))) return rc;
  if ((rc = ws_alloc(w, &p.job_q, size_t(p.m_capacity) * 3))) return rc;
  if ((rc = ws_alloc(w, &p.wide_q, kWideCap))) return rc;
  if ((rc = ws_alloc(w, &p.wide_state, kWideCap))) return rc;
  {
    std::vector<uint2> init(kWideCap, make_uint2(0xFFFFFFFFu, 0u));
    if (cudaMemcpy(p.wide_state, init.data(), sizeof(uint2) * kWideCap, cudaMemcpyHostToDevice) != cudaSuccess) return B2PT_ERR_CUDA;
  }
  if ((rc = ws_alloc(w, &p.dir_q, ne * 4))) return rc;
  if ((rc = ws_alloc(w, &w->counters, 1))) return rc;
  B2_CUDA(cudaMallocHost((void**)&w->host_counters, sizeof(VcmCounters)));
  B2_CUDA(cudaMemset(p.splat, 0, ne * 3 * sizeof(double)));
  B2_CUDA(cudaMemset(p.lp_head, 0xFF, nl * 4));
  B2_CUDA(cudaMemset(p.lp_count, 0, nl * 4));
  return B2PT_OK;
}


// b2pt_debug_density (include/b2pt_debug.h): ONE density loop per thread, run by the same density_begin / density_step the
// stages above run, on explicit jobs, so that the trial count N can be set beside the reference's
// (oracle/_ref/ref_harness density).  The surface of a job is the closest mesh hit of its ray.
__global__ void __launch_bounds__(128) debug_density_kernel(DeviceScene sc, const B2ptDensityJob* __restrict__ jobs, uint32_t n,
                                                            B2ptDensityResult* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const B2ptDensityJob j = jobs[i];
  const float3 O = f3(j.org[0], j.org[1], j.org[2]), D = f3(j.dir[0], j.dir[1], j.dir[2]);
  const Hit h = trace_ray<false, true>(sc, O, D, 0.0f, INFINITY);
  B2ptDensityResult r;
  r.trials = 0u; r.density = INFINITY; r.kind = 0u; r.t = h.t;
  if (h.tri != kInvalid) {
    const Surface s = query_surface(sc, h.tri, h.u, h.v, h.t, O, D);
    Rng rng; rng.key = j.key; rng.ctr = 0;
    DensityLoop L;
    density_begin(L, sc, rng, s, f3(j.target[0], j.target[1], j.target[2]), j.radius, -D);
    if (density_has_lobe(L)) {
      r.kind = L.type == B2PT_BSDF_DIFFUSE ? 1u : 2u;
      unsigned long long rays = 0;
      float density = INFINITY;
      while (!density_step(L, sc, j.radius, density, rays)) {}
      r.density = density;
      r.trials = uint32_t(L.N);  // density_step leaves N at the trial that landed (or at the limit)
    }
  }
  out[i] = r;
}

}  // namespace

struct VcmLaneResult {
  int rc = B2PT_OK;
  unsigned long long basic = 0, shadow = 0, numeric = 0, capacity = 0;
  float ms[kNumStages] = {0.f, 0.f, 0.f};
  uint64_t launches[kNumStages] = {0, 0, 0};
  std::string error;
};

// passes first, first + step, ... of the call, one after the other, in workspace w on `stream`
static int vcm_run_passes(const RenderConfig& base, VcmWorkspace* w, cudaStream_t stream, uint32_t first, uint32_t step, uint32_t photons,
                          VcmLaneResult* out) {
  const B2ptRenderParams& rp = base.params;
  const uint32_t pixels = base.pixels;
  const int block = 256;
  VcmPool& pool = w->pool;
  RenderConfig cfg = base;
  StageTimer timer(stream);
  unsigned long long& basic = out->basic; unsigned long long& shadow = out->shadow;
  unsigned long long& numeric = out->numeric; unsigned long long& capacity = out->capacity;
  const int kBatchIters = 8;
  const uint32_t stride = rp.sample_stride ? rp.sample_stride : 1u;
  const int grid_p = grid_for(pool.photons, block), grid_e = grid_for(pool.pixels_pad, block);
  // traversal stages: trace_queue8 (persistent warps, shared triangle queue; the default), B2PT_TRACE_STREAM=2 closest-hit stages
  // through trace_dyn8, =0 one ray per thread (A/B)
  const char* tv = getenv("B2PT_TRACE_STREAM");
  const bool queue_trace = cfg.scene.nodes8 && cfg.scene.num_tris < (1u << 27) && !(tv && (tv[0] == '0' || tv[0] == '2'));
  const bool dyn_trace = tv && tv[0] == '2' && cfg.scene.nodes8;
  for (uint32_t pass = first; pass < rp.num_samples; pass += step) {
    const uint32_t sample = rp.first_sample + pass * stride;
    VcmParams P;
    P.beta = rp.beta; P.roulette = rp.roulette; P.roulette_inv = 1.0f / rp.roulette;
    P.unbiased = rp.technique == B2PT_TECHNIQUE_UPG; P.enable_vc = rp.enable_vc != 0; P.enable_vm = rp.enable_vm != 0; P.from_light = rp.from_light != 0;
    P.radius = rp.radius;
    // UPGBase::_preprocess (UPG.cpp:180-183), double arithmetic like the reference; num_samples = the pass index
    if (!P.unbiased) P.radius = float(rp.radius * pow(double(sample) + 1.0f, double(rp.alpha * 0.5f - 0.5f)));
    P.circle = kPi * P.radius * P.radius;
    P.clamp_const = P.unbiased ? 1.0f : 3.402823466e+38f;
    P.num_scattered = float(photons); P.num_scattered_inv = 1.0f / float(photons);
    P.num_photons = photons;
    P.max_subpath = rp.max_subpath ? rp.max_subpath : 1024u;
    P.brute_vertices = 0;

    // ---- photons
    B2_CUDA(cudaMemsetAsync(w->counters, 0, sizeof(VcmCounters), stream));
    B2_CUDA(cudaMemsetAsync(pool.p_state, 0, size_t(pool.photons) * 4, stream));
    for (int it = 0;;) {
      for (int k = 0; k < kBatchIters; ++k, ++it) {
        const int parity = it & 1;
        timer.begin(kStageShade); vcm_photon_kernel<<<grid_p, block, 0, stream>>>(cfg, P, pool, w->counters, parity, sample, it == 0 ? 1 : 0); timer.end();
        timer.begin(kStageTrace);
        if (queue_trace) vcm_trace_dyn_kernel<true, true><<<persistent_grid(pool.photons, block), block, 0, stream>>>(cfg, pool, w->counters, parity);
        else if (dyn_trace) vcm_trace_dyn_kernel<true, false><<<persistent_grid(pool.photons, block), block, 0, stream>>>(cfg, pool, w->counters, parity);
        else vcm_photon_trace_kernel<<<grid_p, block, 0, stream>>>(cfg, pool, w->counters, parity);
        timer.end();
      }
      B2_CUDA(cudaMemcpyAsync(w->host_counters, w->counters, sizeof(VcmCounters), cudaMemcpyDeviceToHost, stream));
      B2_CUDA(cudaStreamSynchronize(stream));
      timer.resolve();
      if (w->host_counters->done >= pool.photons) break;
      if (it > (1 << 20)) { set_error("render_vcm: the photon pass did not drain"); return B2PT_ERR_CUDA; }
    }
    const uint32_t num_vertices = min(w->host_counters->lv_count, pool.lv_capacity);
    // ---- photon index
    timer.begin(kStageShade);
    if (num_vertices) vcm_grid_points_kernel<<<(num_vertices + block - 1) / block, block, 0, stream>>>(pool, num_vertices);
    timer.end();
    if (int rc = grid_build_device(w->grid, pool.grid_points, num_vertices, P.radius, stream)) return rc;
    { const char* e = getenv("B2PT_VCM_BRUTE"); if (e && e[0] == '1') P.brute_vertices = num_vertices; }
    // ---- eye paths
    B2_CUDA(cudaMemsetAsync(&w->counters->done, 0, sizeof(unsigned int), stream));
    B2_CUDA(cudaMemsetAsync(w->counters->trace_count, 0, sizeof(unsigned int) * 10, stream));
    B2_CUDA(cudaMemsetAsync(pool.state, 0, size_t(pool.pixels_pad) * 4, stream));
    const int grid_m = grid_for(pool.m_capacity, block);
    for (int it = 0;;) {
      for (int k = 0; k < kBatchIters; ++k, ++it) {
        const int parity = it & 1;
        timer.begin(kStageShade); vcm_eye_kernel<<<grid_e, block, 0, stream>>>(cfg, P, pool, w->counters, parity, sample, it == 0 ? 1 : 0); timer.end();
        timer.begin(kStageShade); vcm_gather_kernel<<<grid_e, block, 0, stream>>>(cfg, P, pool, w->grid.dev, w->counters, parity); timer.end();
        timer.begin(kStageShade); vcm_merge_kernel<<<grid_m, block, 0, stream>>>(cfg, P, pool, w->counters); timer.end();
        if (P.unbiased) { timer.begin(kStageShade); upg_density_kernel<<<persistent_grid(pool.m_capacity, block), block, 0, stream>>>(cfg, P, pool, w->counters); timer.end();
                          timer.begin(kStageShade); upg_density_long_kernel<<<persistent_grid(pool.m_capacity, block), block, 0, stream>>>(cfg, P, pool, w->counters);
                          upg_density_wide_kernel<<<kNumSMs * 2, block, 0, stream>>>(cfg, P, pool, w->counters);
                          upg_density_wide_finish_kernel<<<(kWideCap + block - 1) / block, block, 0, stream>>>(cfg, P, pool, w->counters); timer.end(); }
        timer.begin(kStageShade); vcm_connect_kernel<<<grid_e, block, 0, stream>>>(cfg, P, pool, w->counters, parity); timer.end();
        if (cfg.scene.has_directional) {
          timer.begin(kStageTrace);
          directional_kernel<<<grid_e, block, 0, stream>>>(cfg.scene, pool.dir_q, &w->counters->dir_count[parity], pool.pixels_pad,
                                                          DirectionalParams{P.beta, float(P.enable_vc), float(P.enable_vm), beta_fn_host(P.num_scattered, P.beta)},
                                                          pool.rad, 1u, &w->counters->basic_rays);
          timer.end();
        }
        timer.begin(kStageTrace);
        if (queue_trace) vcm_trace_dyn_kernel<false, true><<<persistent_grid(pool.pixels_pad, block), block, 0, stream>>>(cfg, pool, w->counters, parity);
        else if (dyn_trace) vcm_trace_dyn_kernel<false, false><<<persistent_grid(pool.pixels_pad, block), block, 0, stream>>>(cfg, pool, w->counters, parity);
        else vcm_trace_kernel<<<grid_e, block, 0, stream>>>(cfg, pool, w->counters, parity);
        timer.end();
        timer.begin(kStageShadow);
        if (queue_trace) vcm_shadow_queue_kernel<<<persistent_grid(pool.q_capacity, block), block, 0, stream>>>(cfg, pool, w->counters, parity);
        else vcm_shadow_kernel<<<grid_e, block, 0, stream>>>(cfg, pool, w->counters, parity);
        timer.end();
      }
      B2_CUDA(cudaMemcpyAsync(w->host_counters, w->counters, sizeof(VcmCounters), cudaMemcpyDeviceToHost, stream));
      B2_CUDA(cudaStreamSynchronize(stream));
      timer.resolve();
      if (w->host_counters->done >= pool.pixels_pad) break;
      if (it > (1 << 20)) { set_error("render_vcm: the eye pass did not drain"); return B2PT_ERR_CUDA; }
    }
    timer.begin(kStageShade);
    vcm_commit_kernel<<<grid_for(pixels, block), block, 0, stream>>>(cfg, pool, w->counters, pass);
    timer.end();
    B2_CUDA(cudaMemcpyAsync(w->host_counters, w->counters, sizeof(VcmCounters), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    timer.resolve();
    basic += w->host_counters->basic_rays; shadow += w->host_counters->shadow_rays;
    numeric += w->host_counters->numeric_errors; capacity += w->host_counters->capacity_events;
  }
  B2_CUDA(cudaGetLastError());
  for (int k = 0; k < kNumStages; ++k) { out->ms[k] = timer.ms[k]; out->launches[k] = timer.launches[k]; }
  return B2PT_OK;
}

int render_vcm(b2pt_scene_s* scene, const RenderConfig& base, cudaStream_t stream, B2ptStats* stats) {
  const B2ptRenderParams& rp = base.params;
  B2_CHECK(rp.radius > 0.0f, B2PT_ERR_INVALID, "VCM/UPG: radius must be positive");
  const uint32_t pixels = base.pixels;
  const uint32_t photons = rp.num_photons ? rp.num_photons : pixels;  // Options.cpp:914-916
  // UPGBase::_connect reads _light_offsets[pixel_index + 1] (UPG.cpp:556-560): one light path per pixel is assumed
  B2_CHECK(photons >= pixels, B2PT_ERR_INVALID, "VCM/UPG: num_photons must be at least width * height (UPG.cpp:556-560)");
  B2_CHECK(photons <= (1u << 24), B2PT_ERR_CAPACITY, "VCM/UPG: more than 2^24 light paths per pass");
  int lanes = kVcmDefaultLanes;
  { const char* e = getenv("B2PT_VCM_LANES"); if (e) lanes = std::max(1, std::min(kVcmLanes, atoi(e))); }  // A/B runs
  lanes = int(std::min<uint32_t>(uint32_t(lanes), std::max(1u, rp.num_samples)));
  VcmWorkspaceSet* set = nullptr;
  if (int rc = vcm_workspace(scene, pixels, photons, lv_per_path(rp.roulette), &lanes, &set)) return rc;
  VcmLaneResult res[kVcmLanes];
  if (lanes == 1) {
    res[0].rc = vcm_run_passes(base, set->lane[0], stream, 0, 1, photons, &res[0]);
  } else {
    // lane 1 starts after whatever the caller has queued on `stream` (the film may have just been cleared there)
    int device = 0;
    B2_CUDA(cudaGetDevice(&device));
    B2_CUDA(cudaEventRecord(set->fork, stream));
    std::vector<std::thread> others;
    for (int k = 1; k < lanes; ++k) {
      B2_CUDA(cudaStreamWaitEvent(set->stream[k], set->fork, 0));
      others.emplace_back([&, k]() {
        if (cudaSetDevice(device) != cudaSuccess) { res[k].rc = B2PT_ERR_CUDA; res[k].error = "render_vcm: cudaSetDevice failed in a lane thread"; return; }
        res[k].rc = vcm_run_passes(base, set->lane[k], set->stream[k], uint32_t(k), uint32_t(lanes), photons, &res[k]);
        if (res[k].rc != B2PT_OK) res[k].error = b2pt_last_error();
      });
    }
    res[0].rc = vcm_run_passes(base, set->lane[0], stream, 0, uint32_t(lanes), photons, &res[0]);
    if (res[0].rc != B2PT_OK) res[0].error = b2pt_last_error();
    for (auto& t : others) t.join();  // every lane synchronises its stream at the end of each pass: all work is complete here
    for (int k = 0; k < lanes; ++k) if (res[k].rc != B2PT_OK) { set_error(res[k].error); return res[k].rc; }
  }
  if (res[0].rc != B2PT_OK) return res[0].rc;
  if (stats) {
    stats->num_basic_rays = 0; stats->num_shadow_rays = 0; stats->numeric_errors = 0; stats->capacity_events = 0;
    stats->trace_ms = 0; stats->shadow_ms = 0; stats->shade_ms = 0; stats->trace_launches = 0; stats->shadow_launches = 0; stats->kernel_launches = 0;
    for (int k = 0; k < lanes; ++k) {
      stats->num_basic_rays += res[k].basic; stats->num_shadow_rays += res[k].shadow;
      stats->numeric_errors += res[k].numeric; stats->capacity_events += res[k].capacity;
      stats->trace_ms += res[k].ms[kStageTrace]; stats->shadow_ms += res[k].ms[kStageShadow]; stats->shade_ms += res[k].ms[kStageShade];
      stats->trace_launches += res[k].launches[kStageTrace]; stats->shadow_launches += res[k].launches[kStageShadow];
      stats->kernel_launches += res[k].launches[0] + res[k].launches[1] + res[k].launches[2];
    }
  }
  unsigned long long capacity = 0;
  for (int k = 0; k < lanes; ++k) capacity += res[k].capacity;
  if (capacity) return capacity_error(capacity, "render_vcm");
  return B2PT_OK;
}

}  // namespace b2pt

extern "C" int b2pt_debug_density(b2pt_scene_t scene, const B2ptDensityJob* jobs, size_t n, B2ptDensityResult* out) {
  using namespace b2pt;
  B2_CHECK(scene && (n == 0 || (jobs && out)), B2PT_ERR_INVALID, "b2pt_debug_density: null argument");
  B2_CHECK(n < (size_t(1) << 28), B2PT_ERR_CAPACITY, "b2pt_debug_density: too many jobs");
  if (n == 0) return B2PT_OK;
  B2ptDensityJob* d_jobs = nullptr; B2ptDensityResult* d_out = nullptr;
  B2_CUDA(cudaMalloc(&d_jobs, n * sizeof(B2ptDensityJob)));
  cudaError_t e = cudaMalloc(&d_out, n * sizeof(B2ptDensityResult));
  if (e == cudaSuccess) e = cudaMemcpy(d_jobs, jobs, n * sizeof(B2ptDensityJob), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    debug_density_kernel<<<unsigned((n + 127) / 128), 128>>>(scene->dev, d_jobs, uint32_t(n), d_out);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(out, d_out, n * sizeof(B2ptDensityResult), cudaMemcpyDeviceToHost);
  cudaFree(d_jobs); cudaFree(d_out);
  if (e != cudaSuccess) return cuda_fail(e, "b2pt_debug_density", __FILE__, __LINE__);
  return B2PT_OK;
}
