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
#include "bvh8.cuh"
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
__device__ __forceinline__ float rcp_embree(float a, const uint32_t* table = kRcpTable) {  // a > 0; table: kRcpTable or a copy in shared memory
  const uint32_t b = __float_as_uint(a), e = b >> 23;
  if (e < 2u || e > 252u) return __frcp_rn(a);  // results near the ends of the exponent range: not table territory
  const float r = __uint_as_float(table[(b >> 12) & 0x7FFu] - ((e - 127u) << 23));
  return __fmul_rn(r, __fsub_rn(2.0f, __fmul_rn(r, a)));
}

// Returns true and updates (tfar, u, v) when triangle `pos` is hit inside (tnear, tfar).
// ray_mask: RTCRay::mask — 0xFFFFFFFF (Scene::intersect) or 1 << entity_type::mesh (intersectMesh / occluded).
struct TriData { float4 q0, q1, q2; };
__device__ __forceinline__ TriData tri_load(const DeviceScene& sc, uint32_t pos) {
  const float4* tr = sc.tris + size_t(pos) * kTriFloat4s;
  TriData d;
  d.q0 = __ldg(tr + 0); d.q1 = __ldg(tr + 1); d.q2 = __ldg(tr + 2);
  return d;
}
__device__ __forceinline__ bool intersect_triangle_data(const TriData& d, float3 O, float3 D, float tnear,
                                                        float& tfar, float& u, float& v, uint32_t ray_mask, const uint32_t* rcp_table = kRcpTable) {
  const float4 q0 = d.q0, q1 = d.q1, q2 = d.q2;
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
  const float rcpAbsDen = rcp_embree(absDen, rcp_table);
  tfar = __fmul_rn(T, rcpAbsDen);
  u = __fmul_rn(U, rcpAbsDen);
  v = __fmul_rn(V, rcpAbsDen);
  return true;
}
__device__ __forceinline__ bool intersect_triangle(const DeviceScene& sc, uint32_t pos, float3 O, float3 D, float tnear,
                                                   float& tfar, float& u, float& v, uint32_t ray_mask) {
  const TriData d = tri_load(sc, pos);
  return intersect_triangle_data(d, O, D, tnear, tfar, u, v, ray_mask);
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

// ---------------------------------------------------------------- compressed 8-wide BVH (layout: bvh8.cuh)
// One node of the compressed 8-wide tree against one ray: five 128-bit loads, eight quantised slab tests, and the
// hit children split into a node group G (inner children, priority-ordered hit bits | imask) and a triangle group T.
#ifndef B2PT_FFMA2
#define B2PT_FFMA2 1
#endif
// (d0, d1) = (a0, a1) * (b, b) + (c0, c1) as one packed instruction
__device__ __forceinline__ void ffma2(float a0, float a1, float b, float c0, float c1, float& d0, float& d1) {
  asm("{\n .reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2, %3};\n mov.b64 rb, {%4, %4};\n mov.b64 rc, {%5, %6};\n"
      " fma.rn.ftz.f32x2 rd, ra, rb, rc;\n mov.b64 {%0, %1}, rd;\n}"
      : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b), "f"(c0), "f"(c1));
}
struct Node8Data { float4 n0, n1, n2, n3, n4; };
__device__ __forceinline__ Node8Data node8_load(const DeviceScene& sc, uint32_t child) {
  const float4* nd = sc.nodes8 + size_t(child) * kNode8Float4s;
  Node8Data d;
  d.n0 = __ldg(nd + 0); d.n1 = __ldg(nd + 1); d.n2 = __ldg(nd + 2); d.n3 = __ldg(nd + 3); d.n4 = __ldg(nd + 4);
  return d;
}
__device__ __forceinline__ void node8_test(const Node8Data& d, float3 O, float3 idir, uint32_t oct, float tnear, float tmax, uint2& G, uint2& T, uint32_t k47) {
  struct { float t; } hit = {tmax};
  const float4 n0 = d.n0, n1 = d.n1, n2 = d.n2, n3 = d.n3, n4 = d.n4;
  const uint32_t hdr = __float_as_uint(n0.w);
  const uint32_t imask = hdr >> 24;
  // grid step 2^e per axis (biased exponents), folded into the reciprocal direction; origin of the grid relative to the ray
  const float adx = __uint_as_float((hdr & 0xFFu) << 23) * idir.x, ady = __uint_as_float(((hdr >> 8) & 0xFFu) << 23) * idir.y,
              adz = __uint_as_float(((hdr >> 16) & 0xFFu) << 23) * idir.z;
  // Plane k of an axis is at origin + q * 2^e with q a byte.  The byte is turned into the float 2^15 + q by ONE byte permute (mantissa bits
  // 8..15 of 2^15) and the 2^15 is taken out of the constant term instead of out of every plane: t = (2^15 + q) * ad + (o - 2^15 * ad).
  // The constant loses up to |ad| * 2^-9 to rounding, so entry planes move out by |ad| * 2^-8 and exit planes likewise: boxes grow by
  // 1/256 of a grid step, the test stays conservative, and 48 subtractions per node are gone.
  const float ox = (n0.x - O.x) * idir.x, oy = (n0.y - O.y) * idir.y, oz = (n0.z - O.z) * idir.z;
  const float bx = ox - 32768.0f * adx, by = oy - 32768.0f * ady, bz = oz - 32768.0f * adz;
  const float sx = fabsf(adx) * 0.00390625f, sy = fabsf(ady) * 0.00390625f, sz = fabsf(adz) * 0.00390625f;
  const float oxn = bx - sx, oxf = bx + sx, oyn = by - sy, oyf = by + sy, ozn = bz - sz, ozf = bz + sz;
  // entry / exit planes per axis depend on the sign of the direction
  uint32_t nx[2], fx[2], ny[2], fy[2], nz[2], fz[2];
  { const uint32_t l0 = __float_as_uint(n2.x), l1 = __float_as_uint(n2.y), h0 = __float_as_uint(n3.z), h1 = __float_as_uint(n3.w);
    const bool neg = idir.x < 0.0f; nx[0] = neg ? h0 : l0; nx[1] = neg ? h1 : l1; fx[0] = neg ? l0 : h0; fx[1] = neg ? l1 : h1; }
  { const uint32_t l0 = __float_as_uint(n2.z), l1 = __float_as_uint(n2.w), h0 = __float_as_uint(n4.x), h1 = __float_as_uint(n4.y);
    const bool neg = idir.y < 0.0f; ny[0] = neg ? h0 : l0; ny[1] = neg ? h1 : l1; fy[0] = neg ? l0 : h0; fy[1] = neg ? l1 : h1; }
  { const uint32_t l0 = __float_as_uint(n3.x), l1 = __float_as_uint(n3.y), h0 = __float_as_uint(n4.z), h1 = __float_as_uint(n4.w);
    const bool neg = idir.z < 0.0f; nz[0] = neg ? h0 : l0; nz[1] = neg ? h1 : l1; fz[0] = neg ? l0 : h0; fz[1] = neg ? l1 : h1; }
  // meta byte of a child = bits << 5 | position: an inner child contributes ONE bit at 24 + (slot ^ octant) (stored position 24 + slot),
  // a leaf its unary triangle count at its offset from tri_base; both become `hits |= bits << position`, no branch per child
  uint32_t bits4[2], pos4[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const uint32_t m4 = __float_as_uint(h ? n1.w : n1.z);
    const uint32_t inner4 = (m4 & (m4 << 1) & 0x10101010u) >> 4;  // 0x01 in the bytes of inner children (position >= 24: bits 3 and 4 set)
    pos4[h] = (m4 ^ (inner4 * oct)) & 0x1F1F1F1Fu;
    bits4[h] = (m4 >> 5) & 0x07070707u;
  }
  uint32_t hits = 0u;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#if B2PT_FFMA2
      // entry and exit plane of an axis in ONE packed fp32x2 multiply-add (FFMA2, new with sm_100): 24 instead of 48 per node
      float x0, x1, y0, y1, z0, z1;
      ffma2(byte_to_float15(nx[h], k, k47), byte_to_float15(fx[h], k, k47), adx, oxn, oxf, x0, x1);
      ffma2(byte_to_float15(ny[h], k, k47), byte_to_float15(fy[h], k, k47), ady, oyn, oyf, y0, y1);
      ffma2(byte_to_float15(nz[h], k, k47), byte_to_float15(fz[h], k, k47), adz, ozn, ozf, z0, z1);
#else
      const float x0 = __fmaf_rn(byte_to_float15(nx[h], k, k47), adx, oxn), x1 = __fmaf_rn(byte_to_float15(fx[h], k, k47), adx, oxf);
      const float y0 = __fmaf_rn(byte_to_float15(ny[h], k, k47), ady, oyn), y1 = __fmaf_rn(byte_to_float15(fy[h], k, k47), ady, oyf);
      const float z0 = __fmaf_rn(byte_to_float15(nz[h], k, k47), adz, ozn), z1 = __fmaf_rn(byte_to_float15(fz[h], k, k47), adz, ozf);
#endif
      const float t0 = fmaxf(fmaxf(x0, y0), fmaxf(z0, tnear));
      const float t1 = fminf(fminf(x1, y1), fminf(z1, hit.t)) * 1.0000003f;
      if (t0 <= t1) hits |= ((bits4[h] >> (8 * k)) & 0xFFu) << ((pos4[h] >> (8 * k)) & 0xFFu);
    }
  }
  const uint32_t ghits = hits & 0xFF000000u, tmask = hits & 0x00FFFFFFu;
  G = make_uint2(__float_as_uint(n1.x), ghits | imask);
  T = make_uint2(__float_as_uint(n1.y), tmask);
}
__device__ __forceinline__ void node8_step(const DeviceScene& sc, uint32_t child, float3 O, float3 idir, uint32_t oct, float tnear, float tmax,
                                           uint2& G, uint2& T) {
  const Node8Data d = node8_load(sc, child);
  node8_test(d, O, idir, oct, tnear, tmax, G, T, sc.prmt_base);
}

// kCount (b2pt_debug_trace_counts): counts->x += node steps, counts->y += triangle tests of this ray
template <bool kAnyHit, bool kCount = false>
__device__ __forceinline__ Hit trace_ray8(const DeviceScene& sc, float3 O, float3 D, float tnear, float tfar, uint32_t ray_mask,
                                          uint2* counts = nullptr) {
  Hit hit; hit.t = tfar; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
  const float kHuge = 1e24f;  // node8_step adds 2^15 grid steps to the plane offsets: keep that finite
  const float3 idir = f3(fabsf(D.x) > 1e-24f ? 1.0f / D.x : copysignf(kHuge, D.x),
                         fabsf(D.y) > 1e-24f ? 1.0f / D.y : copysignf(kHuge, D.y),
                         fabsf(D.z) > 1e-24f ? 1.0f / D.z : copysignf(kHuge, D.z));
  // slot s was given to the child lying most towards (s&1 ? +x : -x, s&2 ? +y : -y, s&4 ? +z : -z); with `oct` = the
  // octant the ray travels into, slot ^ oct is 7 for the corner the ray enters first and 0 for the one it leaves last
  const uint32_t oct = (idir.x >= 0.0f ? 1u : 0u) | (idir.y >= 0.0f ? 2u : 0u) | (idir.z >= 0.0f ? 4u : 0u);
  uint2 stack[kStack8];
  int sp = 0;
  uint2 G = make_uint2(0u, (1u << (24u + oct)) | 1u);  // node group: the root as the only (inner) child of a virtual parent
  uint2 T = make_uint2(0u, 0u);                        // triangle group: base, bit j = triangle base + j still to test
  for (;;) {
    if (G.y & 0xFF000000u) {
      // ---- next inner child of the group, nearest corner first
      const int pos = 31 - __clz(int(G.y));
      G.y &= ~(1u << pos);
      const uint32_t slot = uint32_t(pos - 24) ^ oct;
      const uint32_t child = G.x + __popc(G.y & 0xFFu & ((1u << slot) - 1u));
      if (G.y & 0xFF000000u) { if (sp < kStack8) stack[sp++] = G; else atomicOr(sc.error_flags, 1u); }
      node8_step(sc, child, O, idir, oct, tnear, hit.t, G, T);
      if (kCount) ++counts->x;
    } else {
      T = G;  // a postponed triangle group (or nothing)
      G = make_uint2(0u, 0u);
    }
    // ---- the triangles hanging off this node
    while (T.y) {
      const uint32_t j = uint32_t(__ffs(int(T.y)) - 1);
      T.y &= T.y - 1u;
      const uint32_t tri = T.x + j;
      if (kCount) ++counts->y;
      if (intersect_triangle(sc, tri, O, D, tnear, hit.t, hit.u, hit.v, ray_mask)) {
        hit.tri = tri;
        if (kAnyHit) return hit;
      }
    }
    if (!(G.y & 0xFF000000u)) {
      if (sp == 0) break;
      G = stack[--sp];
    }
  }
  return hit;
}

// kMeshOnly: ray mask = meshes only (compile-time form of the above)
// ---------------------------------------------------------------- persistent, warp-synchronous traversal of the 8-wide tree
//
// trace_ray8 run by one thread per ray leaves most of the machine idle on incoherent rays (ncu, profiles/): the node
// step runs with 11 of 32 lanes (finished rays wait for the slowest of their warp) and the triangle test with 2 of 32
// (few lanes hold a triangle at any given step).  trace_stream8 keeps the same tree and the same tests but makes the warp
// the unit of scheduling:
//   * a warp is persistent and refills idle lanes from a global ticket counter once kFetchMin8 of them are idle;
//   * every round is warp-synchronous: first every lane with node work advances ONE node (popping its stack if needed),
//     then every lane that holds triangles tests ONE — both code paths run once per round with all the lanes that can
//     use them, instead of once per lane.  Triangles found while earlier ones are still pending are postponed on the
//     stack as a triangle group.
// Io:  uint32_t tag = io.load(i, O, D, tnear, tfar, mask);  io.store(tag, hit).  Must be entered by complete warps;
// *ticket must be zero at kernel start.
constexpr unsigned kFullMask = 0xFFFFFFFFu;
constexpr int kFetchMin8 = 8;

template <bool kAnyHit, class Io>
__device__ __forceinline__ void trace_stream8(const DeviceScene& sc, uint32_t n, unsigned* ticket, Io& io) {
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lane_lt = (1u << lane) - 1u;
  uint2 stack[kStack8];
  int sp = 0;
  uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);
  uint32_t tag = kInvalid, ray_mask = 0u, oct = 0u;
  bool has_ray = false, exhausted = false;
  float3 O = f3(0.f), D = f3(0.f), idir = f3(0.f);
  float tnear = 0.f;
  Hit hit; hit.t = 0.f; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
  for (;;) {
    // ---- refill
    const unsigned idle = __ballot_sync(kFullMask, !has_ray);
    if (!exhausted && __popc(idle) >= kFetchMin8) {
      const unsigned cnt = __popc(idle);
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(ticket, cnt);
      base = __shfl_sync(kFullMask, base, 0);
      exhausted = base + cnt >= n;
      if (!has_ray) {
        const uint32_t i = base + __popc(idle & lane_lt);
        if (i < n) {
          float tfar;
          tag = io.load(i, O, D, tnear, tfar, ray_mask);
          hit.t = tfar; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
          const float kHuge = 1e24f;  // node8_step adds 2^15 grid steps to the plane offsets: keep that finite
          idir = f3(fabsf(D.x) > 1e-24f ? 1.0f / D.x : copysignf(kHuge, D.x),
                    fabsf(D.y) > 1e-24f ? 1.0f / D.y : copysignf(kHuge, D.y),
                    fabsf(D.z) > 1e-24f ? 1.0f / D.z : copysignf(kHuge, D.z));
          oct = (idir.x >= 0.0f ? 1u : 0u) | (idir.y >= 0.0f ? 2u : 0u) | (idir.z >= 0.0f ? 4u : 0u);
          sp = 0;
          G = make_uint2(0u, (1u << (24u + oct)) | 1u);
          T = make_uint2(0u, 0u);
          has_ray = true;
        }
      }
    }
    if (__ballot_sync(kFullMask, has_ray) == 0u) {
      if (exhausted) break;
      continue;
    }
    // ---- node phase: every lane with node work advances one node; triangles found while others are still pending are
    // postponed onto the stack as a triangle group (they are popped like any other entry)
    if (has_ray) {
      if (!(G.y & 0xFF000000u) && sp > 0) {
        const uint2 e = stack[sp - 1];
        if (e.y & 0xFF000000u) { G = e; --sp; }
        else if (T.y == 0u) { T = e; --sp; }
      }
      if (G.y & 0xFF000000u) {
        const int pos = 31 - __clz(int(G.y));
        G.y &= ~(1u << pos);
        const uint32_t slot = uint32_t(pos - 24) ^ oct;
        const uint32_t child = G.x + __popc(G.y & 0xFFu & ((1u << slot) - 1u));
        if (G.y & 0xFF000000u) { if (sp < kStack8) stack[sp++] = G; else atomicOr(sc.error_flags, 1u); }
        uint2 Tn;
        node8_step(sc, child, O, idir, oct, tnear, hit.t, G, Tn);
        if (Tn.y != 0u) {
          if (T.y != 0u) { if (sp < kStack8) stack[sp++] = T; else atomicOr(sc.error_flags, 1u); }
          T = Tn;
        }
      }
    }
    // ---- triangle phase: lanes that hold triangles test one
    if (has_ray && T.y != 0u) {
      const uint32_t j = uint32_t(__ffs(int(T.y)) - 1);
      T.y &= T.y - 1u;
      const uint32_t tri = T.x + j;
      if (intersect_triangle(sc, tri, O, D, tnear, hit.t, hit.u, hit.v, ray_mask)) {
        hit.tri = tri;
        if (kAnyHit) { T.y = 0u; G.y = 0u; sp = 0; }
      }
    }
    // ---- a ray with nothing left hands its hit over
    if (has_ray && T.y == 0u && !(G.y & 0xFF000000u) && sp == 0) {
      io.store(tag, hit);
      has_ray = false;
    }
  }
}

// ---------------------------------------------------------------- persistent warps, free-running lanes, postponed triangles
//
// trace_dyn8 keeps trace_ray8's per-lane loop (the hardware handles the divergence, no per-round votes) and adds the two
// scheduling rules of Aila & Laine / Ylitie et al.:
//   * dynamic fetch: a warp leaves the traversal loop to refill its idle lanes from the ticket counter once fewer than
//     DeviceScene::dyn_fetch_lanes lanes are still running together;
//   * postponed triangles: a lane that reaches the triangle loop with fewer than DeviceScene::dyn_tri_lanes lanes beside it, and that
//     still has inner children of the current node to visit, pushes the triangle group and goes on with the nodes; the
//     group is popped like any other entry, by which time more lanes tend to hold triangles.  A group is never postponed
//     when the lane has no pending child of the current node, so every postponement is followed by a node step and the
//     loop terminates.
// The tests are the same ones in a different order: a closest hit can differ from trace_ray8's only where two triangles
// are hit at exactly the same distance.

template <bool kAnyHit, class Io>
__device__ __forceinline__ void trace_dyn8(const DeviceScene& sc, uint32_t n, unsigned* ticket, Io& io) {
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lane_lt = (1u << lane) - 1u;
  uint2 stack[kStack8];
  int sp = 0;
  uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);
  uint32_t tag = kInvalid, ray_mask = 0u, oct = 0u;
  bool has_ray = false, exhausted = false;
  float3 O = f3(0.f), D = f3(0.f), idir = f3(0.f);
  float tnear = 0.f;
  Hit hit; hit.t = 0.f; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
  for (;;) {
    __syncwarp();
    // ---- refill the idle lanes
    const unsigned idle = __ballot_sync(kFullMask, !has_ray);
    if (!exhausted && idle) {
      const unsigned cnt = __popc(idle);
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(ticket, cnt);
      base = __shfl_sync(kFullMask, base, 0);
      exhausted = base + cnt >= n;
      if (!has_ray) {
        const uint32_t i = base + __popc(idle & lane_lt);
        if (i < n) {
          float tfar;
          tag = io.load(i, O, D, tnear, tfar, ray_mask);
          hit.t = tfar; hit.u = 0.f; hit.v = 0.f; hit.tri = kInvalid;
          const float kHuge = 1e24f;  // node8_step adds 2^15 grid steps to the plane offsets: keep that finite
          idir = f3(fabsf(D.x) > 1e-24f ? 1.0f / D.x : copysignf(kHuge, D.x),
                    fabsf(D.y) > 1e-24f ? 1.0f / D.y : copysignf(kHuge, D.y),
                    fabsf(D.z) > 1e-24f ? 1.0f / D.z : copysignf(kHuge, D.z));
          oct = (idir.x >= 0.0f ? 1u : 0u) | (idir.y >= 0.0f ? 2u : 0u) | (idir.z >= 0.0f ? 4u : 0u);
          sp = 0;
          G = make_uint2(0u, (1u << (24u + oct)) | 1u);
          T = make_uint2(0u, 0u);
          has_ray = true;
        }
      }
    }
    if (__ballot_sync(kFullMask, has_ray) == 0u) break;  // nothing in flight and nothing fetched: the queue is drained
    // ---- traversal, lanes run freely
    if (has_ray) {
      for (;;) {
        if (G.y & 0xFF000000u) {
          const int pos = 31 - __clz(int(G.y));
          G.y &= ~(1u << pos);
          const uint32_t slot = uint32_t(pos - 24) ^ oct;
          const uint32_t child = G.x + __popc(G.y & 0xFFu & ((1u << slot) - 1u));
          if (G.y & 0xFF000000u) { if (sp < kStack8) stack[sp++] = G; else atomicOr(sc.error_flags, 1u); }
          node8_step(sc, child, O, idir, oct, tnear, hit.t, G, T);
        } else {
          T = G;  // a postponed triangle group (or nothing)
          G = make_uint2(0u, 0u);
        }
        bool done = false;
        while (T.y) {
          if ((G.y & 0xFF000000u) && __popc(__activemask()) < sc.dyn_tri_lanes) {
            if (sp < kStack8) { stack[sp++] = T; T.y = 0u; break; }
          }
          const uint32_t j = uint32_t(__ffs(int(T.y)) - 1);
          T.y &= T.y - 1u;
          const uint32_t tri = T.x + j;
          if (intersect_triangle(sc, tri, O, D, tnear, hit.t, hit.u, hit.v, ray_mask)) {
            hit.tri = tri;
            if (kAnyHit) { done = true; break; }
          }
        }
        if (!done && !(G.y & 0xFF000000u)) {
          if (sp == 0) done = true;
          else G = stack[--sp];
        }
        if (done) {
          io.store(tag, hit);
          has_ray = false;
          T.y = 0u; G.y = 0u; sp = 0;
          break;
        }
        if (!exhausted && __popc(__activemask()) < sc.dyn_fetch_lanes) break;
      }
    }
  }
}

// ---------------------------------------------------------------- persistent warps, warp-synchronous rounds, a shared triangle queue
//
// In trace_dyn8 the leaf test — Embree's arithmetic, 28 % of the stage's warp instructions — runs with 6 of 32 lanes: at any moment
// few lanes of a warp hold a triangle.  trace_queue8 separates WHO FINDS a triangle from WHO TESTS it.  Every round each lane with
// node work advances one node (as in trace_stream8); the triangles its node test finds are not tested by the lane but appended as
// (owner lane, triangle) items to a ring in shared memory, and whenever the ring holds 32 items the warp tests 32 of them at once,
// lane i taking item i: the owner's ray is read from shared memory, a hit is merged into the owner's best hit with a 64-bit
// atomicMin on (t bits, triangle).  A lane whose triangles are still queued goes on with its other nodes; its culling distance
// lags by the queue's latency (a round or two), which costs a few node visits and never a hit.  The ring is drained to fewer than
// 32 items before every node phase and a lane appends at most kEnqueuePerRound triangles per round, so it cannot overflow; when no
// lane can advance a node the remainder is tested as a partial chunk.  The tests are the same ones in another order: a closest hit can
// differ from trace_ray8's only where two triangles are hit at exactly the same distance (the smaller triangle index wins here).
// Needs 256-thread CTAs (one WarpQueue per warp) and fewer than 2^27 triangle references.
#ifndef B2PT_Q_SMEMRCP
#define B2PT_Q_SMEMRCP 0
#endif
#ifndef B2PT_Q_UNROLL2
#define B2PT_Q_UNROLL2 1
#endif
#ifndef B2PT_Q_ENQ
#define B2PT_Q_ENQ 8
#endif
constexpr int kEnqueuePerRound = B2PT_Q_ENQ;
constexpr int kQueueItems = kEnqueuePerRound <= 15 ? 512 : 1024;  // a power of two >= 32 * kEnqueuePerRound + 31
static_assert(kQueueItems >= 32 * kEnqueuePerRound + 31 && kEnqueuePerRound <= 24, "triangle ring too small");
struct WarpQueue {
  uint32_t items[kQueueItems];   // owner lane << 27 | triangle
  float ray[8][32];              // Ox Oy Oz tnear Dx Dy Dz bits(mask), [field][owner lane]
  unsigned long long best[32];   // bits(t) << 32 | triangle; low word kInvalid = no hit yet (t = the ray's tfar)
  float2 uv[32];
};

template <bool kAnyHit, class Io>
__device__ __forceinline__ void trace_queue8(const DeviceScene& sc, uint32_t n, unsigned* ticket, Io& io) {
  __shared__ WarpQueue s_wq[8];
  WarpQueue& q = s_wq[(threadIdx.x >> 5) & 7u];
#if B2PT_Q_SMEMRCP
  __shared__ uint32_t s_rcp[2048];  // the RCPPS table next to the queue: a hit's lookup is a shared-memory load
  for (unsigned i = threadIdx.x; i < 2048u; i += blockDim.x) s_rcp[i] = kRcpTable[i];
  __syncthreads();
  const uint32_t* rcp_table = s_rcp;
#else
  const uint32_t* rcp_table = kRcpTable;
#endif
  const unsigned lane = threadIdx.x & 31u;
  const unsigned lane_lt = (1u << lane) - 1u;
  uint2 stack[kStack8];
  int sp = 0;
  uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);
  uint32_t tag = kInvalid, oct = 0u;
  bool has_ray = false, exhausted = false;  // exhausted: the global counter has passed n (this warp may still hold tickets)
  float3 O = f3(0.f), idir = f3(0.f);
  float tnear = 0.f;
  uint32_t head = 0u, tail = 0u;  // items tested / appended so far (the same in every lane)
  uint32_t my_last = 0u;          // tail after this lane's last append: its items are all tested once head >= my_last
  // Tickets are reserved `chunk` at a time and handed out from the warp's own range [next, end): one atomic per refill on ONE address
  // was the any-hit stage's largest stall (28 % of its warp samples, ncu r02 s17) — 3.5 M short rays per ms are ~350 refills per us.
  uint32_t next = 0u, end = 0u;
  const unsigned chunk = max(32u, min(unsigned(sc.queue_chunk), n / (gridDim.x * 32u)));
  for (;;) {
    // ---- refill
    const unsigned idle = __ballot_sync(kFullMask, !has_ray);
    const unsigned cnt = __popc(idle);
    if (cnt >= unsigned(sc.queue_fetch_lanes) && (!exhausted || next != end)) {
      const unsigned avail = end - next;
      const unsigned r = __popc(idle & lane_lt);
      uint32_t i = next + r;
      bool mine = !has_ray && r < avail;
      if (avail < cnt && !exhausted) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(ticket, chunk);
        base = __shfl_sync(kFullMask, base, 0);
        if (r >= avail) { i = base + (r - avail); mine = !has_ray && i < n; }
        exhausted = base + chunk >= n;
        end = min(base + chunk, n);
        next = min(base + (cnt - avail), end);
        if (base >= n) next = end = 0u;
      } else {
        next += min(cnt, avail);
      }
      if (mine) {
        {
          float tfar; float3 D; uint32_t ray_mask;
          tag = io.load(i, O, D, tnear, tfar, ray_mask);
          const float kHuge = 1e24f;  // node8_step adds 2^15 grid steps to the plane offsets: keep that finite
          idir = f3(fabsf(D.x) > 1e-24f ? 1.0f / D.x : copysignf(kHuge, D.x),
                    fabsf(D.y) > 1e-24f ? 1.0f / D.y : copysignf(kHuge, D.y),
                    fabsf(D.z) > 1e-24f ? 1.0f / D.z : copysignf(kHuge, D.z));
          oct = (idir.x >= 0.0f ? 1u : 0u) | (idir.y >= 0.0f ? 2u : 0u) | (idir.z >= 0.0f ? 4u : 0u);
          sp = 0;
          G = make_uint2(0u, (1u << (24u + oct)) | 1u);
          T = make_uint2(0u, 0u);
          q.ray[0][lane] = O.x; q.ray[1][lane] = O.y; q.ray[2][lane] = O.z; q.ray[3][lane] = tnear;
          q.ray[4][lane] = D.x; q.ray[5][lane] = D.y; q.ray[6][lane] = D.z; q.ray[7][lane] = __uint_as_float(ray_mask);
          q.best[lane] = ((unsigned long long)__float_as_uint(tfar) << 32) | 0xFFFFFFFFull;
          q.uv[lane] = make_float2(0.f, 0.f);
          my_last = head;
          has_ray = true;
        }
      }
    }
    __syncwarp();
    if (__ballot_sync(kFullMask, has_ray) == 0u) {
      if (exhausted && next == end) break;
      continue;
    }
    // ---- node phase: one node per lane; its triangles go to the ring
    bool advanced = false;
    unsigned want = 0;
    if (has_ray) {
      const unsigned long long b = q.best[lane];
      if (kAnyHit && uint32_t(b) != kInvalid) { G.y = 0u; T.y = 0u; sp = 0; }  // occluded: nothing left to look at
      else {
        if (T.y == 0u) {
          if (!(G.y & 0xFF000000u) && sp > 0) G = stack[--sp];
          if (G.y & 0xFF000000u) {
            const int pos = 31 - __clz(int(G.y));
            G.y &= ~(1u << pos);
            const uint32_t slot = uint32_t(pos - 24) ^ oct;
            const uint32_t child = G.x + __popc(G.y & 0xFFu & ((1u << slot) - 1u));
            if (G.y & 0xFF000000u) { if (sp < kStack8) stack[sp++] = G; else atomicOr(sc.error_flags, 1u); }
            node8_step(sc, child, O, idir, oct, tnear, __uint_as_float(uint32_t(b >> 32)), G, T);
            advanced = true;
          }
        }
        want = min(unsigned(__popc(T.y)), unsigned(kEnqueuePerRound));
        // (asking for the lane's next node with prefetch.global.L1 at this point — CCTL.PF1, two per lane and round — was measured at
        // 58.8 instead of 101.9 Msamples/s: the instruction itself is the cost.  profiles/r02_ab_runs.md, session s34.)
      }
    }
    unsigned incl = want;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(kFullMask, incl, o);
      if (lane >= unsigned(o)) incl += t;
    }
    const unsigned total = __shfl_sync(kFullMask, incl, 31);
    if (want) {
      uint32_t at = tail + incl - want;
#if B2PT_Q_UNROLL2
      const uint32_t owner_base = (lane << 27) | T.x;  // T.x + j < 2^27: the sum never carries into the lane bits
      for (unsigned k = 0; k < want; k += 2u, at += 2u) {
        const uint32_t j0 = uint32_t(__ffs(int(T.y)) - 1);
        T.y &= T.y - 1u;
        q.items[at & uint32_t(kQueueItems - 1)] = owner_base + j0;
        if (k + 1u < want) {
          const uint32_t j1 = uint32_t(__ffs(int(T.y)) - 1);
          T.y &= T.y - 1u;
          q.items[(at + 1u) & uint32_t(kQueueItems - 1)] = owner_base + j1;
        }
      }
      my_last = tail + incl;
#else
      for (unsigned k = 0; k < want; ++k, ++at) {
        const uint32_t j = uint32_t(__ffs(int(T.y)) - 1);
        T.y &= T.y - 1u;
        q.items[at & uint32_t(kQueueItems - 1)] = (lane << 27) | (T.x + j);
      }
      my_last = at;
#endif
    }
    tail += total;
    const unsigned any_advanced = __ballot_sync(kFullMask, advanced);
    __syncwarp();
    // ---- triangle phase: 32 items at a time; what is left when no lane could advance a node
    // (a partial chunk also when so few lanes advanced a node that waiting for 32 items would leave the warp mostly idle)
    while (tail - head >= 32u || (tail != head && __popc(any_advanced) < sc.queue_drain_lanes)) {
      const uint32_t m = min(32u, tail - head);
      bool hit = false;
      unsigned long long key = 0ull;
      uint32_t owner = 0u;
      float u = 0.f, v = 0.f;
      if (lane < m) {
        const uint32_t item = q.items[(head + lane) & uint32_t(kQueueItems - 1)];
        owner = item >> 27;
        const uint32_t tri = item & 0x07FFFFFFu;
        const unsigned long long b = q.best[owner];
        if (!(kAnyHit && uint32_t(b) != kInvalid)) {
          const float3 Oo = f3(q.ray[0][owner], q.ray[1][owner], q.ray[2][owner]);
          const float3 Do = f3(q.ray[4][owner], q.ray[5][owner], q.ray[6][owner]);
          float t = __uint_as_float(uint32_t(b >> 32));
          const TriData td = tri_load(sc, tri);
          if (intersect_triangle_data(td, Oo, Do, q.ray[3][owner], t, u, v, __float_as_uint(q.ray[7][owner]), rcp_table)) {
            hit = true;
            key = ((unsigned long long)__float_as_uint(t) << 32) | tri;
            atomicMin(&q.best[owner], key);
          }
        }
      }
      __syncwarp();
      if (hit && q.best[owner] == key) q.uv[owner] = make_float2(u, v);
      head += m;
      __syncwarp();
      if (m < 32u) break;
    }
    // ---- a ray with nothing left, and all of its items tested, hands its hit over
    if (has_ray && T.y == 0u && !(G.y & 0xFF000000u) && sp == 0 && int(head - my_last) >= 0) {
      const unsigned long long b = q.best[lane];
      Hit h;
      h.t = __uint_as_float(uint32_t(b >> 32)); h.tri = uint32_t(b);
      const float2 uv = q.uv[lane];
      h.u = uv.x; h.v = uv.y;
      io.store(tag, h);
      has_ray = false;
    }
  }
}

template <bool kAnyHit>
__device__ __forceinline__ Hit trace_any_bvh(const DeviceScene& sc, float3 O, float3 D, float tnear, float tfar, uint32_t ray_mask) {
#ifdef B2PT_EXPERIMENTS
  // A/B builds (-DB2PT_EXPERIMENTS): the scene carries one tree or the other (B2PT_BVH=4 builds the 4-wide uncompressed tree)
  if (!sc.nodes8) return trace_ray_masked<kAnyHit>(sc, O, D, tnear, tfar, ray_mask);
#endif
  return trace_ray8<kAnyHit>(sc, O, D, tnear, tfar, ray_mask);
}
template <bool kAnyHit, bool kMeshOnly>
__device__ __forceinline__ Hit trace_ray(const DeviceScene& sc, float3 O, float3 D, float tnear, float tfar) {
  return trace_any_bvh<kAnyHit>(sc, O, D, tnear, tfar, kMeshOnly ? kMaskMesh : kMaskAll);
}


// blocks of a persistent traversal launch over n rays: one wave of 256-thread CTAs, 4 per SM
inline int grid_for_trace(size_t n, int block) {
  size_t want = (n + block - 1) / block;
  size_t cap = size_t(kNumSMs) * 8 * 4;
  if (want < 1) want = 1;
  return int(want < cap ? want : cap);
}
inline int persistent_grid(size_t n, int block) {
  size_t want = (n + block - 1) / block;
  size_t cap = size_t(kNumSMs) * 4;
  if (want < 1) want = 1;
  return int(want < cap ? want : cap);
}

// geomID / primID of a triangle in leaf order
__device__ __forceinline__ uint32_t tri_geom(const DeviceScene& sc, uint32_t pos) { return __float_as_uint(__ldg(sc.tris + size_t(pos) * kTriFloat4s).w); }
__device__ __forceinline__ uint32_t tri_prim(const DeviceScene& sc, uint32_t pos) { return __float_as_uint(__ldg(sc.tris + size_t(pos) * kTriFloat4s + 1).w); }

}  // namespace b2pt
