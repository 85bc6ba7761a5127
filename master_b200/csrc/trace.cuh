// trace.cuh — closest-hit and any-hit ray queries against the device BVH.
//
// Replaces rtcIntersect / rtcOccluded as the reference calls them (Scene.cpp:151-226,
// Intersector.cpp:17-46).  One ray per thread; a short traversal stack in registers/local memory;
// every node is fetched with seven 128-bit loads, every triangle with three.
//
// The leaf test follows Embree's MoellerTrumboreIntersector1
// (submodules/embree/kernels/geometry/triangle_intersector_moeller.h:78-111 and the hit
// finalisation :40-50) operation by operation, with Embree's dot/cross association
// (common/math/vec3.h:188,193: dot = a.x*b.x + (a.y*b.y + a.z*b.z)) and unfused multiplies
// (the oracle build caps Embree at the AVX ISA, which has no FMA), so den, U, V, T — and with
// them the accept/reject decision for a given triangle — are bit-identical to the reference's, and so are
// t, u, v: rcp(|den|) goes through the same RCPPS table + one Newton step as Embree's (rcp_embree below).
//
// Box tests are NOT part of the parity contract (the tree is a different tree); they use fused
// multiply-adds and are made conservative: child boxes carry a small absolute pad (scene.cu) and
// the exit distance is stretched by 2 ulp, so a triangle the leaf test would accept is never culled.
#pragma once
#include "scene.cuh"

namespace b2pt {

constexpr int kStackSize = 64;

struct Hit {
  float t, u, v;
  uint32_t tri;  // triangle position in leaf order, kInvalid on a miss
};

__device__ __forceinline__ float xor_sign(float x, uint32_t sgn) { return __uint_as_float(__float_as_uint(x) ^ sgn); }

// Embree dot: madd(a.x, b.x, madd(a.y, b.y, a.z*b.z)) without FMA
__device__ __forceinline__ float dot_embree(float3 a, float3 b) {
  return __fadd_rn(__fmul_rn(a.x, b.x), __fadd_rn(__fmul_rn(a.y, b.y), __fmul_rn(a.z, b.z)));
}
// Embree cross: (msub(a.y,b.z,a.z*b.y), msub(a.z,b.x,a.x*b.z), msub(a.x,b.y,a.y*b.x)) without FMA
__device__ __forceinline__ float3 cross_embree(float3 a, float3 b) {
  return f3(__fsub_rn(__fmul_rn(a.y, b.z), __fmul_rn(a.z, b.y)),
            __fsub_rn(__fmul_rn(a.z, b.x), __fmul_rn(a.x, b.z)),
            __fsub_rn(__fmul_rn(a.x, b.y), __fmul_rn(a.y, b.x)));
}

// Embree's rcp(a) = r * (2 - r * a) with r = _mm_rcp_ps(a) (common/simd/vfloat4_sse2.h:181-188, non-AVX2 branch): RCPPS is
// a 2048-entry table instruction (tools/make_rcp_table.c extracts and verifies the table), so with the table here the hit
// finalisation t = T * rcp(|den|), u = U * rcp, v = V * rcp (triangle_intersector_moeller.h:44-50) is bit-identical.
static __device__ const uint32_t kRcpTable[2048] = {
#include "rcp_table.inc"
};
__device__ __forceinline__ float rcp_embree(float a) {  // a > 0
  const uint32_t b = __float_as_uint(a), e = b >> 23;
  if (e < 2u || e > 252u) return __frcp_rn(a);  // results near the ends of the exponent range: not table territory
  const float r = __uint_as_float(kRcpTable[(b >> 12) & 0x7FFu] - ((e - 127u) << 23));
  return __fmul_rn(r, __fsub_rn(2.0f, __fmul_rn(r, a)));
}

// Returns true and updates (tfar, u, v) when triangle `pos` is hit inside (tnear, tfar).
// ray_mask: RTCRay::mask — 0xFFFFFFFF (Scene::intersect) or 1 << entity_type::mesh (intersectMesh / occluded).
__device__ __forceinline__ bool intersect_triangle(const DeviceScene& sc, uint32_t pos, float3 O, float3 D, float tnear,
                                                   float& tfar, float& u, float& v, uint32_t ray_mask) {
  const float4 q0 = __ldg(&sc.tri_v0[pos]);
  const float4 q1 = __ldg(&sc.tri_e1[pos]);
  const float4 q2 = __ldg(&sc.tri_e2[pos]);
  // geometry mask = 1 << (material_id & 3) (rtcSetMask, Scene.cpp:47); Embree skips the triangle when (mask & ray.mask) == 0
  if ((__float_as_uint(q2.w) & ray_mask) == 0u) return false;
  const float3 v0 = f3(q0.x, q0.y, q0.z), e1 = f3(q1.x, q1.y, q1.z), e2 = f3(q2.x, q2.y, q2.z);
  const float3 Ng = cross_embree(e1, e2);  // Triangle4 stores Ng = cross(e1, e2)
  const float3 C = f3(__fsub_rn(v0.x, O.x), __fsub_rn(v0.y, O.y), __fsub_rn(v0.z, O.z));
  const float3 R = cross_embree(D, C);
  const float den = dot_embree(Ng, D);
  const float absDen = fabsf(den);
  const uint32_t sgnDen = __float_as_uint(den) & 0x80000000u;
  const float U = xor_sign(dot_embree(R, e2), sgnDen);
  const float V = xor_sign(dot_embree(R, e1), sgnDen);
  if (!((den != 0.0f) & (U >= 0.0f) & (V >= 0.0f) & (__fadd_rn(U, V) <= absDen))) return false;
  const float T = xor_sign(dot_embree(Ng, C), sgnDen);
  if (!((T > __fmul_rn(absDen, tnear)) & (T < __fmul_rn(absDen, tfar)))) return false;
  const float rcpAbsDen = rcp_embree(absDen);
  tfar = __fmul_rn(T, rcpAbsDen);
  u = __fmul_rn(U, rcpAbsDen);
  v = __fmul_rn(V, rcpAbsDen);
  return true;
}

__device__ __forceinline__ void cswap(float& ta, uint32_t& ca, float& tb, uint32_t& cb) {
  if (tb < ta) { float t = ta; ta = tb; tb = t; uint32_t c = ca; ca = cb; cb = c; }
}

constexpr uint32_t kMaskAll = 0xFFFFFFFFu;
constexpr uint32_t kMaskMesh = 1u << B2PT_ENTITY_MESH;

// kAnyHit: stop at the first accepted triangle (rtcOccluded).
template <bool kAnyHit>
__device__ __forceinline__ Hit trace_ray_masked(const DeviceScene& sc, float3 O, float3 D, float tnear, float tfar, uint32_t ray_mask) {
  Hit hit; hit.t = tfar; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
  // safe reciprocal direction: a zero component becomes +-huge instead of inf so 0*inf never appears
  const float kHuge = 1e30f;
  const float3 idir = f3(fabsf(D.x) > 1e-30f ? 1.0f / D.x : copysignf(kHuge, D.x),
                         fabsf(D.y) > 1e-30f ? 1.0f / D.y : copysignf(kHuge, D.y),
                         fabsf(D.z) > 1e-30f ? 1.0f / D.z : copysignf(kHuge, D.z));
  const float3 ood = f3(O.x * idir.x, O.y * idir.y, O.z * idir.z);
  uint32_t stack[kStackSize];
  int sp = 0;
  uint32_t cur = 0;  // root is always an inner node
  while (true) {
    if (!child_is_leaf(cur)) {
      const float4* n = sc.nodes + size_t(cur) * kNodeFloat4s;
      const float4 lox = __ldg(n + 0), loy = __ldg(n + 1), loz = __ldg(n + 2);
      const float4 hix = __ldg(n + 3), hiy = __ldg(n + 4), hiz = __ldg(n + 5);
      const float4 lk = __ldg(n + 6);
      float tn[4];
      uint32_t ch[4] = {__float_as_uint(lk.x), __float_as_uint(lk.y), __float_as_uint(lk.z), __float_as_uint(lk.w)};
      const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, ly[4] = {loy.x, loy.y, loy.z, loy.w}, lz[4] = {loz.x, loz.y, loz.z, loz.w};
      const float hx[4] = {hix.x, hix.y, hix.z, hix.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float x0 = __fmaf_rn(lx[k], idir.x, -ood.x), x1 = __fmaf_rn(hx[k], idir.x, -ood.x);
        const float y0 = __fmaf_rn(ly[k], idir.y, -ood.y), y1 = __fmaf_rn(hy[k], idir.y, -ood.y);
        const float z0 = __fmaf_rn(lz[k], idir.z, -ood.z), z1 = __fmaf_rn(hz[k], idir.z, -ood.z);
        const float t0 = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tnear));
        const float t1 = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), hit.t)) * 1.0000003f;
        const bool ok = (t0 <= t1) & (ch[k] != kInvalid);
        tn[k] = ok ? t0 : INFINITY;
      }
      // sort the (up to four) children by entry distance: 5 compare-exchanges
      cswap(tn[0], ch[0], tn[1], ch[1]); cswap(tn[2], ch[2], tn[3], ch[3]);
      cswap(tn[0], ch[0], tn[2], ch[2]); cswap(tn[1], ch[1], tn[3], ch[3]);
      cswap(tn[1], ch[1], tn[2], ch[2]);
      if (tn[0] == INFINITY) {
        if (sp == 0) break;
        cur = stack[--sp];
        continue;
      }
      // far children first onto the stack, nearest is visited next
      if (tn[3] != INFINITY) { if (sp < kStackSize) stack[sp++] = ch[3]; else atomicOr(sc.error_flags, 1u); }
      if (tn[2] != INFINITY) { if (sp < kStackSize) stack[sp++] = ch[2]; else atomicOr(sc.error_flags, 1u); }
      if (tn[1] != INFINITY) { if (sp < kStackSize) stack[sp++] = ch[1]; else atomicOr(sc.error_flags, 1u); }
      cur = ch[0];
    } else {
      const uint32_t first = leaf_first(cur), count = leaf_count(cur);
      for (uint32_t k = 0; k < count; ++k) {
        if (intersect_triangle(sc, first + k, O, D, tnear, hit.t, hit.u, hit.v, ray_mask)) {
          hit.tri = first + k;
          if (kAnyHit) return hit;
        }
      }
      if (sp == 0) break;
      cur = stack[--sp];
    }
  }
  return hit;
}

// kMeshOnly: ray mask = meshes only (compile-time form of the above)
template <bool kAnyHit, bool kMeshOnly>
__device__ __forceinline__ Hit trace_ray(const DeviceScene& sc, float3 O, float3 D, float tnear, float tfar) {
  return trace_ray_masked<kAnyHit>(sc, O, D, tnear, tfar, kMeshOnly ? kMaskMesh : kMaskAll);
}

// ---------------------------------------------------------------- warp-voted persistent traversal
//
// trace_ray_masked above runs one ray per thread to completion; ncu (profiles/r01_bpt_trace_full_summary.md)
// shows what that costs on incoherent rays: 6.9 of 32 lanes active per issued instruction, the triangle test
// running with 2-3 lanes because a lane that reaches a leaf tests its triangles while the other 31 wait at
// the reconvergence point, and finished lanes idling until the slowest ray of the warp ends.
//
// trace_stream keeps the same tree, the same box test and the same (Embree-exact) triangle test, but changes
// who does what when:
//   * ROUNDS.  Every lane holds one pending item `cur`: an inner node, one triangle of a leaf, or nothing.
//     Each round the warp votes (two ballots): if more lanes hold triangles than inner nodes the round is a
//     triangle round, otherwise a node round; only the lanes of the winning class work, each on exactly ONE
//     item, so the instructions of a round run with at least half of the busy lanes.  Lanes of the other class
//     keep their item for a later round ("postponed leaves").
//   * PERSISTENT WARPS, DYNAMIC FETCH.  A warp does not own 32 fixed rays: when kFetchMin or more lanes are
//     idle they take the next rays from a global ticket counter (one atomic per refill), so a long ray no
//     longer holds 31 finished lanes hostage.
// Io supplies the rays and consumes the hits:  uint32_t tag = io.load(i, O, D, tnear, tfar, mask);
// io.store(tag, hit).  Must be entered by complete warps; `ticket` must be zero at kernel start.
constexpr uint32_t kSentinel = 0x7FFFFFFFu;  // bottom of every traversal stack: this lane has no item
constexpr unsigned kFullMask = 0xFFFFFFFFu;
constexpr int kFetchMin = 8;

template <bool kAnyHit, class Io, bool kCount = false>
__device__ __forceinline__ void trace_stream(const DeviceScene& sc, uint32_t n, unsigned* ticket, Io& io) {
  uint32_t n_nodes = 0, n_tris = 0;  // kCount (diagnostic builds only): work items of the current ray
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lane_lt = (1u << lane) - 1u;
  uint32_t stack[kStackSize];
  int sp = 1;
  stack[0] = kSentinel;
  uint32_t cur = kSentinel, tag = kInvalid, ray_mask = 0u;
  bool has_ray = false, exhausted = false;
  float3 O = f3(0.f), D = f3(0.f), idir = f3(0.f), ood = f3(0.f);
  float tnear = 0.f;
  Hit hit; hit.t = 0.f; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
  for (;;) {
    // ---- refill idle lanes
    const unsigned idle = __ballot_sync(kFullMask, cur == kSentinel);
    if (!exhausted && __popc(idle) >= kFetchMin) {
      const unsigned cnt = __popc(idle);
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(ticket, cnt);
      base = __shfl_sync(kFullMask, base, 0);
      exhausted = base + cnt >= n;
      if (cur == kSentinel) {
        const uint32_t i = base + __popc(idle & lane_lt);
        if (i < n) {
          float tfar;
          tag = io.load(i, O, D, tnear, tfar, ray_mask);
          hit.t = tfar; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
          const float kHuge = 1e30f;
          idir = f3(fabsf(D.x) > 1e-30f ? 1.0f / D.x : copysignf(kHuge, D.x),
                    fabsf(D.y) > 1e-30f ? 1.0f / D.y : copysignf(kHuge, D.y),
                    fabsf(D.z) > 1e-30f ? 1.0f / D.z : copysignf(kHuge, D.z));
          ood = f3(O.x * idir.x, O.y * idir.y, O.z * idir.z);
          sp = 1;
          cur = 0;  // the root is always an inner node
          has_ray = true;
          if (kCount) { n_nodes = 0; n_tris = 0; }
        }
      }
    }
    // ---- vote
    const bool is_inner = cur < kSentinel;
    const bool is_leaf = (cur & 0x80000000u) != 0u;
    const unsigned mi = __ballot_sync(kFullMask, is_inner), ml = __ballot_sync(kFullMask, is_leaf);
    if ((mi | ml) == 0u) {
      if (exhausted) break;
      continue;
    }
    if (__popc(ml) > __popc(mi)) {
      // ---- triangle round: one triangle per lane
      if (is_leaf) {
        const uint32_t pos = cur & 0x0FFFFFFFu;
        const bool h = intersect_triangle(sc, pos, O, D, tnear, hit.t, hit.u, hit.v, ray_mask);
        if (h) hit.tri = pos;
        if (kCount) ++n_tris;
        if (kAnyHit && h) cur = kSentinel;
        else if (cur & 0x70000000u) cur += 1u - 0x10000000u;  // next triangle of the same leaf
        else cur = stack[--sp];
      }
    } else if (is_inner) {
      // ---- node round: one 4-wide node per lane
      if (kCount) ++n_nodes;
      const float4* nd = sc.nodes + size_t(cur) * kNodeFloat4s;
      const float4 lox = __ldg(nd + 0), loy = __ldg(nd + 1), loz = __ldg(nd + 2);
      const float4 hix = __ldg(nd + 3), hiy = __ldg(nd + 4), hiz = __ldg(nd + 5);
      const float4 lk = __ldg(nd + 6);
      float tn[4];
      uint32_t ch[4] = {__float_as_uint(lk.x), __float_as_uint(lk.y), __float_as_uint(lk.z), __float_as_uint(lk.w)};
      const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, ly[4] = {loy.x, loy.y, loy.z, loy.w}, lz[4] = {loz.x, loz.y, loz.z, loz.w};
      const float hx[4] = {hix.x, hix.y, hix.z, hix.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float x0 = __fmaf_rn(lx[k], idir.x, -ood.x), x1 = __fmaf_rn(hx[k], idir.x, -ood.x);
        const float y0 = __fmaf_rn(ly[k], idir.y, -ood.y), y1 = __fmaf_rn(hy[k], idir.y, -ood.y);
        const float z0 = __fmaf_rn(lz[k], idir.z, -ood.z), z1 = __fmaf_rn(hz[k], idir.z, -ood.z);
        const float t0 = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), tnear));
        const float t1 = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), hit.t)) * 1.0000003f;
        const bool ok = (t0 <= t1) & (ch[k] != kInvalid);
        tn[k] = ok ? t0 : INFINITY;
      }
      cswap(tn[0], ch[0], tn[1], ch[1]); cswap(tn[2], ch[2], tn[3], ch[3]);
      cswap(tn[0], ch[0], tn[2], ch[2]); cswap(tn[1], ch[1], tn[3], ch[3]);
      cswap(tn[1], ch[1], tn[2], ch[2]);
      if (tn[0] == INFINITY) {
        cur = stack[--sp];
      } else {
        if (tn[3] != INFINITY) { if (sp < kStackSize) stack[sp++] = ch[3]; else atomicOr(sc.error_flags, 1u); }
        if (tn[2] != INFINITY) { if (sp < kStackSize) stack[sp++] = ch[2]; else atomicOr(sc.error_flags, 1u); }
        if (tn[1] != INFINITY) { if (sp < kStackSize) stack[sp++] = ch[1]; else atomicOr(sc.error_flags, 1u); }
        cur = ch[0];
      }
    }
    // ---- a lane whose stack ran empty hands its hit over
    if (has_ray && cur == kSentinel) {
      if (kCount) { hit.u = float(n_nodes); hit.v = float(n_tris); }
      io.store(tag, hit);
      has_ray = false;
    }
  }
}

// blocks of a persistent traversal launch over n rays: one wave of 256-thread CTAs, 4 per SM
inline int persistent_grid(size_t n, int block) {
  size_t want = (n + block - 1) / block;
  size_t cap = size_t(kNumSMs) * 4;
  if (want < 1) want = 1;
  return int(want < cap ? want : cap);
}

// geomID / primID of a triangle in leaf order
__device__ __forceinline__ uint32_t tri_geom(const DeviceScene& sc, uint32_t pos) { return __float_as_uint(__ldg(&sc.tri_v0[pos]).w); }
__device__ __forceinline__ uint32_t tri_prim(const DeviceScene& sc, uint32_t pos) { return __float_as_uint(__ldg(&sc.tri_e1[pos]).w); }

}  // namespace b2pt
