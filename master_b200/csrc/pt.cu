// pt.cu — wavefront path tracer: PathTracing::_traceEye / _connect (PT.cpp:15-120) re-thought as
// three persistent-style stages over a pool of path slots in HBM.
//
// The reference runs one pixel-sample to completion on one thread.  Here the unit of work is one
// STAGE over every path in flight:
//
//   logic   (pt_logic_kernel)   per slot: consume the closest hit found by the previous trace —
//                               light hit: MIS-weighted emission, continue through the light;
//                               surface hit: throughput update, Russian roulette, next-event
//                               estimation set-up (shadow ray + unoccluded contribution into the
//                               shadow queue), BSDF sample (extension ray into the trace queue);
//                               finished paths commit to the film and the slot immediately
//                               REGENERATES a new (pixel, sample) so the pool stays full.
//   trace   (pt_trace_kernel)   closest hit for every slot in the trace queue (mask: meshes + lights)
//   shadow  (pt_shadow_kernel)  any-hit for every entry of the shadow queue (mask: meshes only);
//                               visible -> add the contribution to the path's radiance.
//
// Queues hold slot indices, compacted with one warp-aggregated atomic per warp.  Path state is SoA
// float4 streams so each stage moves 128-bit words.  Every path owns a counter-based RNG stream
// (rng.cuh), so the draw order inside a path is the reference's whatever the stage interleaving.
#include <stdlib.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "render_common.cuh"

namespace b2pt {

enum { kFree = 0, kPrimary = 1, kBounce = 2, kAwaitShadow = 3 };
enum { kQueueTrace = 0, kQueueShadow = 1, kQueueFree = 2, kQueueTicket = 3 };  // [3]: ray ticket of the persistent traversal stage
enum { kShadowCommit = 1u };

// One 96-byte record per slot = three 32-byte sectors, each read and written whole: the stages reach slots through queues, i.e. as
// gathers, and the same words kept as seven SoA streams cost seven half-used sectors per slot and iteration (224 B of DRAM traffic
// for 96 B of state; the BPT pool made the same move in round 1).
//   0 extension ray origin (already offset)        1 extension ray direction                    <- all the closest-hit stage reads
//   2 radiance gathered so far                     3 previous vertex position, w = bsdf.density of the sampled direction
//   4 throughput of the NEXT vertex (already * f * cos / pdf), w = bits(finite | path_size << 1)
//   5 bits(rng key, rng counter, work index low, high)
enum { kPtRayO = 0, kPtRayD = 1, kPtRad = 2, kPtPrev = 3, kPtThr = 4, kPtCtl = 5, kPtRec = 6 };
struct PtPool {
  float4* rec;      // [slot * kPtRec + word]
  float4* hit;      // t, u, v, bits(triangle) of the ray at POSITION i of the trace queue (not by slot)
  uint32_t* trace_q[2];  // slots with an extension ray in flight = the live paths (bit 31: waits for its primary hit); iteration i writes [i & 1], iteration i + 1 walks it
  uint32_t* free_q[2];   // slots the shadow stage committed and freed: the next logic stage refills them
  float4* shadow_q; // 3 float4 per entry: (org, bits slot) (dir, bits flags) (contribution, 0)
};
#define PT_W(pool, slot, word) ((pool).rec[size_t(slot) * kPtRec + (word)])
__device__ __forceinline__ float4 pt_ctl(uint32_t key, uint32_t ctr, unsigned long long work) {
  return make_float4(__uint_as_float(key), __uint_as_float(ctr), __uint_as_float(uint32_t(work)), __uint_as_float(uint32_t(work >> 32)));
}
__device__ __forceinline__ unsigned long long pt_ctl_work(float4 c) {
  return (unsigned long long)__float_as_uint(c.z) | ((unsigned long long)__float_as_uint(c.w) << 32);
}

__device__ __forceinline__ unsigned long long warp_grab_work(unsigned long long* counter, bool want) {
  const unsigned active = __activemask();
  const unsigned mask = __ballot_sync(active, want);
  if (!want) return ~0ull;
  const unsigned lane = threadIdx.x & 31u;
  const int leader = __ffs(mask) - 1;
  unsigned long long base = 0;
  if (lane == unsigned(leader)) base = atomicAdd(counter, (unsigned long long)__popc(mask));
  base = __shfl_sync(mask, base, leader);
  return base + __popc(mask & ((1u << lane) - 1u));
}

// The stage walks the LIVE slots only — the trace queue of the previous iteration — plus the slots the shadow stage freed
// (`first`: every slot, all free).  Walking the pool in slot order costs a full scan per iteration however few paths are
// left, which is what the drain phase of a render call consists of.
#ifndef B2PT_PT_LOGIC_THREADS
#define B2PT_PT_LOGIC_THREADS 256  // CTA size and resident CTAs per SM the logic stage is compiled for (-D... for A/B builds): 256 x 2 = 128 registers
#endif
#ifndef B2PT_PT_LOGIC_MIN_BLOCKS
#define B2PT_PT_LOGIC_MIN_BLOCKS 2
#endif
__global__ void __launch_bounds__(B2PT_PT_LOGIC_THREADS, B2PT_PT_LOGIC_MIN_BLOCKS) pt_logic_kernel(RenderConfig cfg, PtPool pool, int parity, int first) {
  const DeviceScene& sc = cfg.scene;
  RenderCounters* cnt = cfg.counters;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // [parity ^ 1][trace / free] are this launch's inputs; the trace stage clears them afterwards
    cnt->queue_count[parity ^ 1][kQueueShadow] = 0;
    cnt->queue_count[parity][kQueueTicket] = 0;
    cnt->shadow_ticket[parity] = 0;
  }
  const float beta = cfg.params.beta;
  const uint32_t max_path = cfg.params.max_path ? cfg.params.max_path : 0xFFFFFFFFu;
  const uint32_t n_trace = first ? 0u : cnt->queue_count[parity ^ 1][kQueueTrace];
  // freed slots only matter while there is work left to hand out
  const uint32_t n_free = (first || cnt->next_work >= cfg.total_work) ? 0u : cnt->queue_count[parity ^ 1][kQueueFree];
  const uint32_t n_live = first ? cfg.pool_size : n_trace + n_free;
  const uint32_t n_round = (n_live + 31u) & ~31u;  // whole warps: every lane of a warp reaches the ballots below (the padding lanes do nothing)
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
    const bool lane_on = i < n_live;
    // A live slot's queue entry says which kind of vertex it waits for (bit 31: the primary hit), and its closest hit sits at the same
    // POSITION of the hit array (the closest-hit stage stores hits in queue order): queue entry and hit are one coalesced round trip,
    // the slot's streams and the shading record of the hit triangle the next — no state word, no hit gathered by slot in between.
    uint32_t slot, st = kFree;
    float4 hv = make_float4(0.f, 0.f, 0.f, __uint_as_float(kInvalid));
    if (!lane_on) {
      slot = 0;
    } else if (first) {
      slot = i;
    } else if (i < n_trace) {
      const uint32_t e = pool.trace_q[parity ^ 1][i];
      slot = e & 0x7FFFFFFFu;
      st = (e & 0x80000000u) ? uint32_t(kPrimary) : uint32_t(kBounce);
      hv = pool.hit[i];
    } else {
      slot = pool.free_q[parity ^ 1][i - n_trace];
    }
    SurfaceRaw raw;
    const bool has_tri = st != kFree && __float_as_uint(hv.w) != kInvalid;
    if (has_tri) load_surface_early(sc, __float_as_uint(hv.w), raw);
    bool want_trace = false, want_shadow = false;
    float3 sh_o = f3(0.f), sh_d = f3(0.f), sh_c = f3(0.f);
    uint32_t sh_flags = 0;
    float3 O = f3(0.f), D = f3(0.f), rad = f3(0.f);
    Rng rng; rng.key = 0; rng.ctr = 0;
    unsigned long long work = 0;

    if (st == kPrimary || st == kBounce) {
      const float4 ro = PT_W(pool, slot, kPtRayO), rd = PT_W(pool, slot, kPtRayD), rv = PT_W(pool, slot, kPtRad);
      const float4 cw = PT_W(pool, slot, kPtCtl);
      const uint2 rs = make_uint2(__float_as_uint(cw.x), __float_as_uint(cw.y));
      O = f3(ro.x, ro.y, ro.z); D = f3(rd.x, rd.y, rd.z); rad = f3(rv.x, rv.y, rv.z);
      rng.key = rs.x; rng.ctr = rs.y;
      work = pt_ctl_work(cw);
      const uint32_t tri = __float_as_uint(hv.w);
      bool terminated = false, proceed = false;
      float3 thr = f3(1.0f);
      uint32_t path_size = 2;
      Surface s;
      if (tri == kInvalid) {
        terminated = true;  // PT.cpp:49-51 / :28-30: the ray left the scene
      } else {
        load_surface_rest(sc, tri, raw);
        s = query_surface_raw(raw, hv.y, hv.z, hv.x, O, D);
        const float3 omega = -D;
        if (st == kPrimary) {
          if (s.is_light() && max_path > 0) {  // PT.cpp:23-26: lights seen directly, then look through them
            float3 lr; float ld;
            query_lsdf(sc, s, omega, lr, ld);
            rad += cfg.params.lights * lr;
            O = offset_origin(s.pos, s.gnormal, D);
            want_trace = true;
          } else if (max_path < 2) {
            terminated = true;
          } else {
            proceed = true;  // first surface vertex: throughput 1, path_size 2 (PT.cpp:32-38)
          }
        } else {
          const float4 pv = PT_W(pool, slot, kPtPrev), tv = PT_W(pool, slot, kPtThr);
          const float bsdf_density = pv.w;
          const uint32_t packed = __float_as_uint(tv.w);
          const uint32_t finite = packed & 1u;
          path_size = packed >> 1;
          thr = f3(tv.x, tv.y, tv.z);
          // Edge(eye[prv].surface, eye[itr].surface, eye[itr].omega) (PT.cpp:56): only fGeometry is still needed
          const float dist_sq_inv = 1.0f / distance2(f3(pv.x, pv.y, pv.z), s.pos);
          const float f_geometry = dist_sq_inv * fabsf(dot(omega, s.normal()));
          if (s.is_light()) {  // PT.cpp:69-80: emission with the power heuristic, then through the light
            float3 lr; float ld;
            query_lsdf(sc, s, omega, lr, ld);
            float weight_inv = pow_beta(ld, beta) / pow_beta(f_geometry * bsdf_density, beta) + 1.0f;
            if (finite == 0) weight_inv = 1.0f;
            rad += lr * thr / weight_inv;
            O = offset_origin(s.pos, s.gnormal, D);
            want_trace = true;
          } else {  // PT.cpp:84-94: Russian roulette once the path has 3 vertices
            const float roulette = path_size < 3 ? 1.0f : cfg.params.roulette;
            const float uniform = rng.sample();
            if (roulette < uniform) {
              terminated = true;
            } else {
              thr /= roulette;
              ++path_size;
              if (path_size > max_path) terminated = true; else proceed = true;
            }
          }
        }
      }

      if (proceed) {
        const float3 eye_omega = -D;
        // ---- PathTracing::_connect (PT.cpp:100-120): next-event estimation
        const LightSample light = sample_light(sc, rng);
        const float3 omega = normalize(s.pos - light.surface.pos);
        const BsdfQuery lb = bsdf_query_emitter(sc, light.surface, light.surface.normal(), omega);  // a sampled emitter point: LightBSDF or sun_light_bsdf
        if (!(l1norm(lb.throughput) < kFltEpsilon)) {
          const BsdfQuery eb = bsdf_query(sc, s, -omega, eye_omega);
          const Edge edge(light.surface.pos, light.surface.normal(), s.pos, s.normal(), omega);
          const float weight_inv = pow_beta(eb.density_rev * edge.b_geometry, beta) / pow_beta(light.combined_density(), beta) + 1.0f;
          sh_c = light.radiance / light.combined_density() * thr * eb.throughput * edge.b_cos * edge.f_geometry / weight_inv;
          // Scene::occluded(eye.surface, light.surface) (Scene.cpp:151-171)
          const float3 direction = normalize(light.surface.pos - s.pos);
          const float3 a = s.pos + (dot(s.gnormal, direction) > 0.0f ? 1.0f : -1.0f) * s.gnormal * 0.0001f;
          const float3 b = light.surface.pos + (dot(light.surface.gnormal, direction) < 0.0f ? 1.0f : -1.0f) * light.surface.gnormal * 0.0001f;
          sh_o = a; sh_d = b - a;
          want_shadow = !(l1norm(sh_c) == 0.0f);  // a zero contribution needs no visibility test
        }
        // ---- Scene::sampleBSDF + the throughput the next vertex will carry (PT.cpp:43-66)
        const BsdfSample bs = bsdf_sample(sc, rng, s, eye_omega);
        const float b_cos = fabsf(dot(-bs.omega, s.normal()));
        float3 thr_next = thr * bs.throughput * b_cos;
        if (l1norm(thr_next) < kFltEpsilon) {
          terminated = true;  // PT.cpp:61-63 (taken whether or not the extension ray would have hit)
        } else {
          thr_next /= bs.density;
          D = bs.omega;
          O = offset_origin(s.pos, s.gnormal, D);
          PT_W(pool, slot, kPtPrev) = make_float4(s.pos.x, s.pos.y, s.pos.z, bs.density);
          PT_W(pool, slot, kPtThr) = make_float4(thr_next.x, thr_next.y, thr_next.z, __uint_as_float((bs.finite & 1u) | (path_size << 1)));
          PT_W(pool, slot, kPtRayD) = make_float4(D.x, D.y, D.z, 0.f);
          st = kBounce;
          want_trace = true;
        }
      }

      if (terminated) {
        if (want_shadow) {
          sh_flags = kShadowCommit;  // the shadow stage adds the last contribution, then commits
          st = kAwaitShadow;
          PT_W(pool, slot, kPtRad) = make_float4(rad.x, rad.y, rad.z, 0.f);
        } else {
          commit_sample(cfg, work, rad);
          st = kFree;
        }
      } else {
        PT_W(pool, slot, kPtRayO) = make_float4(O.x, O.y, O.z, 0.f);
        PT_W(pool, slot, kPtRad) = make_float4(rad.x, rad.y, rad.z, 0.f);
        PT_W(pool, slot, kPtCtl) = pt_ctl(rng.key, rng.ctr, work);
      }
    }

    // ---- regeneration: a free slot takes the next (sample, pixel) (Technique::_for_each_ray, Technique.cpp:307-352)
    {
      const bool was_free = lane_on && st == kFree;
      const unsigned long long w = warp_grab_work(&cnt->next_work, was_free);
      if (was_free && w < cfg.total_work) {
        const uint32_t pixel = uint32_t(w % cfg.pixels);
        const uint32_t pass = uint32_t(w / cfg.pixels);
        const uint32_t sample = cfg.params.first_sample + pass * (cfg.params.sample_stride ? cfg.params.sample_stride : 1u);
        Rng g; g.key = stream_key(cfg.params.seed, kDomainPixel, pixel, sample); g.ctr = 0;
        const float3 dir = shoot_primary(cfg, g, pixel);
        const float3 org = offset_origin(cfg.camera.pos, cfg.camera.gnormal, dir);
        PT_W(pool, slot, kPtRayO) = make_float4(org.x, org.y, org.z, 0.f);
        PT_W(pool, slot, kPtRayD) = make_float4(dir.x, dir.y, dir.z, 0.f);
        PT_W(pool, slot, kPtRad) = make_float4(0.f, 0.f, 0.f, 0.f);
        PT_W(pool, slot, kPtCtl) = pt_ctl(g.key, g.ctr, w);
        st = kPrimary;
        want_trace = true;
      }
    }
    // both reservations by lane 0, back to back: one round trip to the L2 per warp instead of two in a row (every lane of the warp is
    // here: the loop's trip count is rounded up to whole warps)
    const unsigned lane_id = threadIdx.x & 31u;
    const unsigned m_t = __ballot_sync(0xFFFFFFFFu, want_trace), m_s = __ballot_sync(0xFFFFFFFFu, want_shadow);
    unsigned base_t = 0, base_s = 0;
    if (lane_id == 0) {
      if (m_t) base_t = atomicAdd(&cnt->queue_count[parity][kQueueTrace], unsigned(__popc(m_t)));
      if (m_s) base_s = atomicAdd(&cnt->queue_count[parity][kQueueShadow], unsigned(__popc(m_s)));
    }
    base_t = __shfl_sync(0xFFFFFFFFu, base_t, 0);
    base_s = __shfl_sync(0xFFFFFFFFu, base_s, 0);
    const unsigned tq = base_t + __popc(m_t & ((1u << lane_id) - 1u)), sq = base_s + __popc(m_s & ((1u << lane_id) - 1u));
    if (want_trace) pool.trace_q[parity][tq] = slot | (st == kPrimary ? 0x80000000u : 0u);
    if (want_shadow) {
      float4* e = pool.shadow_q + size_t(sq) * 3;
      e[0] = make_float4(sh_o.x, sh_o.y, sh_o.z, __uint_as_float(slot));
      e[1] = make_float4(sh_d.x, sh_d.y, sh_d.z, __uint_as_float(sh_flags));
      e[2] = make_float4(sh_c.x, sh_c.y, sh_c.z, 0.f);
    }
  }
}

struct PtTraceIo {
  const PtPool* pool;
  const uint32_t* q;
  __device__ __forceinline__ uint32_t load(uint32_t i, float3& O, float3& D, float& tnear, float& tfar, uint32_t& mask) const {
    const uint32_t slot = q[i] & 0x7FFFFFFFu;
    const float4 o = PT_W(*pool, slot, kPtRayO), d = PT_W(*pool, slot, kPtRayD);
    O = f3(o.x, o.y, o.z); D = f3(d.x, d.y, d.z); tnear = 0.0f; tfar = INFINITY; mask = kMaskAll;
    return i;  // the hit goes to position i of the hit array: the logic stage reads it with the queue entry
  }
  __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
    pool->hit[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri));
  }
};

// the same stage with persistent warps, dynamic fetch and postponed triangle groups (trace.cuh: trace_dyn8)
__global__ void __launch_bounds__(256, 4) pt_trace_dyn_kernel(RenderConfig cfg, PtPool pool, int parity) {
  const uint32_t n = cfg.counters->queue_count[parity][kQueueTrace];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cfg.counters->basic_rays, (unsigned long long)n);
    cfg.counters->queue_count[parity ^ 1][kQueueTrace] = 0;
    cfg.counters->queue_count[parity ^ 1][kQueueFree] = 0;
  }
  PtTraceIo io{&pool, pool.trace_q[parity]};
  trace_dyn8<false>(cfg.scene, n, &cfg.counters->queue_count[parity][kQueueTicket], io);
}

// the same stage with a shared triangle queue per warp (trace.cuh: trace_queue8), the default
__global__ void __launch_bounds__(256, 4) pt_trace_queue_kernel(RenderConfig cfg, PtPool pool, int parity) {
  const uint32_t n = cfg.counters->queue_count[parity][kQueueTrace];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cfg.counters->basic_rays, (unsigned long long)n);
    cfg.counters->queue_count[parity ^ 1][kQueueTrace] = 0;
    cfg.counters->queue_count[parity ^ 1][kQueueFree] = 0;
  }
  PtTraceIo io{&pool, pool.trace_q[parity]};
  trace_queue8<false>(cfg.scene, n, &cfg.counters->queue_count[parity][kQueueTicket], io);
}

__global__ void __launch_bounds__(256) pt_trace_kernel(RenderConfig cfg, PtPool pool, int parity) {
  const uint32_t n = cfg.counters->queue_count[parity][kQueueTrace];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    atomicAdd(&cfg.counters->basic_rays, (unsigned long long)n);
    cfg.counters->queue_count[parity ^ 1][kQueueTrace] = 0;  // consumed by this iteration's logic stage
    cfg.counters->queue_count[parity ^ 1][kQueueFree] = 0;
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t slot = pool.trace_q[parity][i] & 0x7FFFFFFFu;
    const float4 o = PT_W(pool, slot, kPtRayO), d = PT_W(pool, slot, kPtRayD);
    // Scene::intersect: tnear 0, tfar inf, mask 0xFFFFFFFF (Scene.cpp:190-197)
    const Hit h = trace_ray<false, false>(cfg.scene, f3(o.x, o.y, o.z), f3(d.x, d.y, d.z), 0.0f, INFINITY);
    pool.hit[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri));
  }
}

__global__ void __launch_bounds__(256) pt_shadow_kernel(RenderConfig cfg, PtPool pool, int parity) {
  const uint32_t n = cfg.counters->queue_count[parity][kQueueShadow];
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cfg.counters->shadow_rays, (unsigned long long)n);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4* e = pool.shadow_q + size_t(i) * 3;
    const float4 o = e[0], d = e[1], c = e[2];
    const uint32_t slot = __float_as_uint(o.w), flags = __float_as_uint(d.w);
    // Scene::occluded: t in (0, 1) along the unnormalised segment, meshes only (Scene.cpp:162-176)
    const Hit h = trace_ray<true, true>(cfg.scene, f3(o.x, o.y, o.z), f3(d.x, d.y, d.z), 0.0f, 1.0f);
    float4 r = PT_W(pool, slot, kPtRad);
    if (h.tri == kInvalid) {
      r.x += c.x; r.y += c.y; r.z += c.z;
      if (!(flags & kShadowCommit)) PT_W(pool, slot, kPtRad) = r;
    }
    if (flags & kShadowCommit) {
      commit_sample(cfg, pt_ctl_work(PT_W(pool, slot, kPtCtl)), f3(r.x, r.y, r.z));
      pool.free_q[parity][warp_append(&cfg.counters->queue_count[parity][kQueueFree], true)] = slot;
    }
  }
}

// The any-hit stage through trace_queue8: entry i of the shadow queue is ray i; what the one-ray-per-thread kernel does after the
// query (add the contribution if visible, commit the sample and free the slot if this was the path's last query) happens in store(),
// whichever lanes of the warp finish in that round (warp_append works on the active lanes).
struct PtShadowIo {
  const RenderConfig* cfg;
  const PtPool* pool;
  int parity;
  __device__ __forceinline__ uint32_t load(uint32_t i, float3& O, float3& D, float& tnear, float& tfar, uint32_t& mask) const {
    const float4* e = pool->shadow_q + size_t(i) * 3;
    const float4 o = e[0], d = e[1];
    O = f3(o.x, o.y, o.z); D = f3(d.x, d.y, d.z); tnear = 0.0f; tfar = 1.0f; mask = kMaskMesh;  // Scene::occluded (Scene.cpp:162-176)
    return i;
  }
  __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
    const float4* e = pool->shadow_q + size_t(i) * 3;
    const uint32_t slot = __float_as_uint(e[0].w), flags = __float_as_uint(e[1].w);
    const float4 c = e[2];
    float4 r = PT_W(*pool, slot, kPtRad);
    if (h.tri == kInvalid) {
      r.x += c.x; r.y += c.y; r.z += c.z;
      if (!(flags & kShadowCommit)) PT_W(*pool, slot, kPtRad) = r;
    }
    const bool commit = (flags & kShadowCommit) != 0u;
    if (commit) {
      commit_sample(*cfg, pt_ctl_work(PT_W(*pool, slot, kPtCtl)), f3(r.x, r.y, r.z));
    }
    const unsigned fq = warp_append(&cfg->counters->queue_count[parity][kQueueFree], commit);
    if (commit) pool->free_q[parity][fq] = slot;
  }
};

__global__ void __launch_bounds__(256, 4) pt_shadow_queue_kernel(RenderConfig cfg, PtPool pool, int parity) {
  const uint32_t n = cfg.counters->queue_count[parity][kQueueShadow];
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&cfg.counters->shadow_rays, (unsigned long long)n);
  PtShadowIo io{&cfg, &pool, parity};
  trace_queue8<true>(cfg.scene, n, &cfg.counters->shadow_ticket[parity], io);
}

// ---------------------------------------------------------------- host driver
struct PtWorkspace {
  PtPool pool{};
  uint32_t pool_size = 0;  // slots in use by the current call (<= capacity)
  uint32_t capacity = 0;   // slots allocated: a call that needs no more reuses the pool (b2pt_render_host ramps its batch sizes)
  RenderCounters* counters = nullptr;
  RenderCounters* host_counters = nullptr;  // pinned
  void* blocks[24] = {nullptr};
  int num_blocks = 0;
};

static void pt_workspace_free(void* p) {
  PtWorkspace* w = (PtWorkspace*)p;
  for (int i = 0; i < w->num_blocks; ++i) cudaFree(w->blocks[i]);
  if (w->host_counters) cudaFreeHost(w->host_counters);
  delete w;
}

template <class T>
static int ws_alloc(PtWorkspace* w, T** p, size_t count) {
  void* q = nullptr;
  B2_CUDA(cudaMalloc(&q, count * sizeof(T)));
  w->blocks[w->num_blocks++] = q;
  *p = (T*)q;
  return B2PT_OK;
}

// Two halves of a call's passes run at the same time, each in its own pool on its own stream with its own host thread (render_pt):
// the latency-bound logic stage of one half overlaps the issue-bound traversal stages of the other.
constexpr int kPtLanes = 4;        // most lanes (B2PT_PT_LANES)
constexpr int kPtDefaultLanes = 2;
struct PtWorkspaceSet {
  PtWorkspace* lane[kPtLanes] = {};
  cudaStream_t stream[kPtLanes] = {};  // lanes 1.. (lane 0 runs on the caller's stream)
  cudaEvent_t fork = nullptr;
};
static void pt_workspace_set_free(void* p) {
  PtWorkspaceSet* set = (PtWorkspaceSet*)p;
  for (PtWorkspace* w : set->lane) if (w) pt_workspace_free(w);
  for (cudaStream_t st : set->stream) if (st) cudaStreamDestroy(st);
  if (set->fork) cudaEventDestroy(set->fork);
  delete set;
}

static int pt_workspace_fill(PtWorkspace* w, uint32_t pool_size);

static int pt_workspace(b2pt_scene_s* scene, int k, uint32_t pool_size, PtWorkspaceSet** out_set, PtWorkspace** out) {
  PtWorkspaceSet* set = (PtWorkspaceSet*)scene->workspace;
  if (!(set && scene->workspace_free == pt_workspace_set_free)) {
    if (scene->workspace && scene->workspace_free) scene->workspace_free(scene->workspace);
    scene->workspace = nullptr;
    set = new PtWorkspaceSet();
    scene->workspace = set; scene->workspace_free = pt_workspace_set_free;
  }
  *out_set = set;
  if (set->lane[k] && set->lane[k]->capacity >= pool_size) { set->lane[k]->pool_size = pool_size; *out = set->lane[k]; return B2PT_OK; }
  if (set->lane[k]) { pt_workspace_free(set->lane[k]); set->lane[k] = nullptr; }
  set->lane[k] = new PtWorkspace();
  *out = set->lane[k];
  return pt_workspace_fill(set->lane[k], pool_size);
}

static int pt_workspace_fill(PtWorkspace* w, uint32_t pool_size) {
  w->pool_size = pool_size; w->capacity = pool_size;
  int rc;
  if ((rc = ws_alloc(w, &w->pool.rec, size_t(pool_size) * kPtRec))) return rc;
  if ((rc = ws_alloc(w, &w->pool.hit, pool_size))) return rc;
  for (int k = 0; k < 2; ++k) {
    if ((rc = ws_alloc(w, &w->pool.trace_q[k], pool_size))) return rc;
    if ((rc = ws_alloc(w, &w->pool.free_q[k], pool_size))) return rc;
  }
  if ((rc = ws_alloc(w, &w->pool.shadow_q, size_t(pool_size) * 3))) return rc;
  if ((rc = ws_alloc(w, &w->counters, 1))) return rc;
  B2_CUDA(cudaMallocHost((void**)&w->host_counters, sizeof(RenderCounters)));
  return B2PT_OK;
}

static uint32_t pt_pool_size(const RenderConfig& base) {
  const int block = 256;
  // pool: enough slots to fill the machine several times over, never more than the work, multiple of the block
  // measured (LivingRoom stand-in, 1024^2): 16 passes run at 110 / 124 / 138 / 161 Msamples/s with 2^21 / 2^22 / 2^23 / 2^24 slots, 64 passes at 154 / 197
  // with 2^24 / 2^26: the more paths in flight the larger the launches, and refilled slots are slower to shade than fresh ones (scattered pixels)
  unsigned long long cap = 1ull << 25;
  { const char* e = getenv("B2PT_PT_POOL_LOG2"); if (e) { int v = atoi(e); if (v >= 10 && v <= 26) cap = 1ull << v; } }  // A/B runs
  unsigned long long want = base.total_work < cap ? base.total_work : cap;
  uint32_t pool_size = uint32_t((want + block - 1) / block * block);
  return pool_size;
}

// one lane: the passes of `base` in pool w on `stream`, synchronous
static int render_pt_lane(const RenderConfig& base, PtWorkspace* w, cudaStream_t stream, B2ptStats* stats) {
  const int block = 256;
  const uint32_t pool_size = w->pool_size;
  RenderConfig cfg = base;
  cfg.pool_size = pool_size;
  cfg.counters = w->counters;
  B2_CUDA(cudaMemsetAsync(w->counters, 0, sizeof(RenderCounters), stream));
  const int grid = grid_for(pool_size, block);
  const int logic_block = B2PT_PT_LOGIC_THREADS;  // the same number of threads in flight whatever the CTA size
  const int logic_grid = int(std::min((size_t(pool_size) + logic_block - 1) / logic_block, size_t(kNumSMs) * 8 * (block / logic_block)));
  // closest-hit stage: persistent warps + dynamic fetch + postponed triangle groups (166.4 -> 170.9 Msamples/s on the LivingRoom
  // stand-in); B2PT_TRACE_STREAM=0 selects one ray per thread for A/B runs
  const char* tv = getenv("B2PT_TRACE_STREAM");
  const bool dyn_trace = !(tv && tv[0] == '0') && cfg.scene.nodes8;
  // 3 (default) = shared triangle queue (trace_queue8) for both stages, 2 = trace_dyn8 for the closest-hit stage, 0 = one ray per thread
  const bool queue_trace = dyn_trace && !(tv && tv[0] == '2') && cfg.scene.num_tris < (1u << 27);
  StageTimer timer(stream);
  const int kBatch = 8;  // stage iterations between two looks at the device counters
  for (int it = 0;; ) {
    for (int k = 0; k < kBatch; ++k, ++it) {
      const int parity = it & 1;
      timer.begin(kStageShade); pt_logic_kernel<<<logic_grid, logic_block, 0, stream>>>(cfg, w->pool, parity, it == 0 ? 1 : 0); timer.end();
      timer.begin(kStageTrace);
      if (queue_trace) pt_trace_queue_kernel<<<persistent_grid(cfg.pool_size, block), block, 0, stream>>>(cfg, w->pool, parity);
      else if (dyn_trace) pt_trace_dyn_kernel<<<persistent_grid(cfg.pool_size, block), block, 0, stream>>>(cfg, w->pool, parity);
      else pt_trace_kernel<<<grid, block, 0, stream>>>(cfg, w->pool, parity);
      timer.end();
      timer.begin(kStageShadow);
      if (queue_trace) pt_shadow_queue_kernel<<<persistent_grid(cfg.pool_size, block), block, 0, stream>>>(cfg, w->pool, parity);
      else pt_shadow_kernel<<<grid, block, 0, stream>>>(cfg, w->pool, parity);
      timer.end();
    }
    B2_CUDA(cudaMemcpyAsync(w->host_counters, w->counters, sizeof(RenderCounters), cudaMemcpyDeviceToHost, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    timer.resolve();
    const RenderCounters& c = *w->host_counters;
    const int last = (it - 1) & 1;
    if (c.next_work >= cfg.total_work && c.queue_count[last][kQueueTrace] == 0 && c.queue_count[last][kQueueShadow] == 0) break;
    if (it > (1 << 22)) { set_error("render_pt: path pool did not drain"); return B2PT_ERR_CUDA; }
  }
  B2_CUDA(cudaGetLastError());
  if (stats) {
    const RenderCounters& c = *w->host_counters;
    stats->num_basic_rays = c.basic_rays;
    stats->num_shadow_rays = c.shadow_rays;
    stats->numeric_errors = c.numeric_errors;
    stats->capacity_events = c.capacity_events;
    timer.fill(stats);
  }
  return B2PT_OK;
}


int render_pt(b2pt_scene_s* scene, const RenderConfig& base, cudaStream_t stream, B2ptStats* stats) {
  const B2ptRenderParams& rp = base.params;
  int lanes = kPtDefaultLanes;
  { const char* e = getenv("B2PT_PT_LANES"); if (e) lanes = std::max(1, std::min(kPtLanes, atoi(e))); }  // A/B runs
  // per-path radiance is written by pass index: only a single lane keeps that layout
  if (base.path_radiance) lanes = 1;
  lanes = int(std::min<uint32_t>(uint32_t(lanes), std::max(1u, rp.num_samples)));
  RenderConfig cfg[kPtLanes];
  PtWorkspaceSet* set = nullptr;
  PtWorkspace* w[kPtLanes] = {};
  const uint32_t stride = rp.sample_stride ? rp.sample_stride : 1u;
  for (int k = 0; k < lanes; ++k) {
    cfg[k] = base;
    if (lanes > 1) {  // lane k takes passes k, k + lanes, ...
      cfg[k].params.first_sample = rp.first_sample + uint32_t(k) * stride;
      cfg[k].params.sample_stride = stride * uint32_t(lanes);
      cfg[k].params.num_samples = (rp.num_samples - uint32_t(k) + uint32_t(lanes) - 1u) / uint32_t(lanes);
      cfg[k].total_work = (unsigned long long)base.pixels * cfg[k].params.num_samples;
      cfg[k].pass_index0 = uint32_t(k); cfg[k].pass_index_step = uint32_t(lanes);
    }
    if (int rc = pt_workspace(scene, k, pt_pool_size(cfg[k]), &set, &w[k])) {
      if (k == 0) return rc;
      cudaGetLastError();  // no memory for a second pool: one lane
      lanes = 1; cfg[0] = base;  // (pass_index0 = 0, pass_index_step = 1)
      if (int rc0 = pt_workspace(scene, 0, pt_pool_size(cfg[0]), &set, &w[0])) return rc0;
      break;
    }
  }
  if (lanes == 1) return render_pt_lane(cfg[0], w[0], stream, stats);
  for (int k = 1; k < lanes; ++k) if (!set->stream[k]) { B2_CUDA(cudaStreamCreateWithFlags(&set->stream[k], cudaStreamNonBlocking)); scene_l2_window(scene, set->stream[k]); }
  if (!set->fork) B2_CUDA(cudaEventCreateWithFlags(&set->fork, cudaEventDisableTiming));
  int device = 0;
  B2_CUDA(cudaGetDevice(&device));
  B2_CUDA(cudaEventRecord(set->fork, stream));  // the other lanes start after whatever the caller queued on `stream` (the film may just have been cleared there)
  B2ptStats st[kPtLanes] = {};
  int rcs[kPtLanes] = {};
  std::string errs[kPtLanes];
  std::vector<std::thread> others;
  for (int k = 1; k < lanes; ++k) {
    B2_CUDA(cudaStreamWaitEvent(set->stream[k], set->fork, 0));
    others.emplace_back([&, k]() {
      if (cudaSetDevice(device) != cudaSuccess) { rcs[k] = B2PT_ERR_CUDA; errs[k] = "render_pt: cudaSetDevice failed in a lane thread"; return; }
      rcs[k] = render_pt_lane(cfg[k], w[k], set->stream[k], &st[k]);
      if (rcs[k] != B2PT_OK) errs[k] = b2pt_last_error();
    });
  }
  rcs[0] = render_pt_lane(cfg[0], w[0], stream, &st[0]);
  for (auto& t : others) t.join();  // every lane ends with a synchronised stream
  if (rcs[0] != B2PT_OK) return rcs[0];
  for (int k = 1; k < lanes; ++k) if (rcs[k] != B2PT_OK) { set_error(errs[k]); return rcs[k]; }
  if (stats) {
    *stats = st[0];
    for (int k = 1; k < lanes; ++k) {
      stats->num_basic_rays += st[k].num_basic_rays; stats->num_shadow_rays += st[k].num_shadow_rays;
      stats->numeric_errors += st[k].numeric_errors; stats->capacity_events += st[k].capacity_events;
      stats->trace_ms += st[k].trace_ms; stats->shadow_ms += st[k].shadow_ms; stats->shade_ms += st[k].shade_ms;
      stats->trace_launches += st[k].trace_launches; stats->shadow_launches += st[k].shadow_launches;
      stats->kernel_launches += st[k].kernel_launches;
    }
  }
  return B2PT_OK;
}

}  // namespace b2pt
