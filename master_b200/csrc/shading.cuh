// shading.cuh — surface reconstruction, BSDFs, lobe sampling, lights, cameras: the per-vertex
// arithmetic of the reference's scene services, as straight-line device functions (no virtual
// dispatch: the BSDF class becomes a switch on B2ptMaterial::type).
//
// Every function names the reference code it restates; the floating-point association follows
// the reference (glm 0.9.8 semantics, see common.cuh) because north_star's per-path level
// compares radiance at 1e-4.  Local frame convention: y is the normal; toSurface(w) = w * tangent,
// toWorld(s) = tangent * s (SurfacePoint.hpp:49-50).
#pragma once
#include "rng.cuh"
#include "scene.cuh"

namespace b2pt {

// SurfacePoint (SurfacePoint.hpp:37-64)
struct Surface {
  float3 pos;
  float3 gnormal;
  mat3 tan;  // c0 = bitangent, c1 = normal, c2 = tangent
  uint32_t material_id;

  __device__ __forceinline__ float3 normal() const { return tan.c1; }
  __device__ __forceinline__ float3 to_world(float3 s) const { return mul(tan, s); }
  __device__ __forceinline__ float3 to_surface(float3 w) const { return mul(w, tan); }
  __device__ __forceinline__ bool is_light() const { return (material_id & 3u) == B2PT_ENTITY_LIGHT; }
  __device__ __forceinline__ bool is_camera() const { return (material_id & 3u) == B2PT_ENTITY_CAMERA; }
  __device__ __forceinline__ uint32_t material_index() const { return material_id >> 2; }
};

// BSDFQuery / BSDFSample (BSDF.hpp:9-21)
struct BsdfQuery {
  float3 throughput;
  float density, density_rev;
  uint32_t finite;
};
struct BsdfSample : BsdfQuery {
  float3 omega;
};

// Edge (SurfacePoint.hpp:66-83)
struct Edge {
  float dist_sq_inv, f_cos, b_cos, f_geometry, b_geometry;
  __device__ __forceinline__ Edge(float3 fst_pos, float3 fst_normal, float3 snd_pos, float3 snd_normal, float3 omega) {
    dist_sq_inv = 1.0f / distance2(fst_pos, snd_pos);
    f_cos = fabsf(dot(omega, snd_normal));
    b_cos = fabsf(dot(omega, fst_normal));
    f_geometry = dist_sq_inv * f_cos;
    b_geometry = dist_sq_inv * b_cos;
  }
};

// MIS exponent policy: FixedBeta<0|1|2> / VariableBeta (Beta.hpp:24-42); PT uses pow(x, beta) (PT.cpp:72,112)
__device__ __forceinline__ float beta_fn(float x, float beta) {
  if (beta == 2.0f) return x * x;
  if (beta == 1.0f) return x;
  if (beta == 0.0f) return x == 0.0f ? 0.0f : 1.0f;
  return cr_powf(x, beta);
}
__device__ __forceinline__ float pow_beta(float x, float beta) {  // std::pow(x, beta) as PT.cpp calls it
  if (beta == 1.0f) return x;
  if (beta == 2.0f) return x * x;
  if (beta == 0.0f) return 1.0f;
  return cr_powf(x, beta);
}

// ---------------------------------------------------------------- Scene::querySurface (Scene.cpp:80-126)
// The words of triangle `tri` querySurface needs: its shading record (7 float4) and the two edge words of its intersection record.
// Split from the arithmetic so that a stage can ask for them as soon as it knows the triangle, next to its other loads.
struct SurfaceRaw { float4 a[7]; float4 q1, q2; };
__device__ __forceinline__ SurfaceRaw load_surface_raw(const DeviceScene& sc, uint32_t tri) {
  const float4* fr = sc.tri_frames + size_t(tri) * kFrameFloat4s;
  SurfaceRaw r;
#pragma unroll
  for (int k = 0; k < 7; ++k) r.a[k] = __ldg(fr + k);
  r.q1 = __ldg(sc.tris + size_t(tri) * kTriFloat4s + 1);
  r.q2 = __ldg(sc.tris + size_t(tri) * kTriFloat4s + 2);
  return r;
}
// The same in two steps for a stage short of registers: the first and the last word of the shading record and one edge word are asked
// for early (they pull the record's one or two 128-byte lines and the triangle's sector into L1), the other six words where they are used.
__device__ __forceinline__ void load_surface_early(const DeviceScene& sc, uint32_t tri, SurfaceRaw& r) {
  const float4* fr = sc.tri_frames + size_t(tri) * kFrameFloat4s;
  r.a[0] = __ldg(fr + 0);
  r.a[6] = __ldg(fr + 6);
  r.q1 = __ldg(sc.tris + size_t(tri) * kTriFloat4s + 1);
}
__device__ __forceinline__ void load_surface_rest(const DeviceScene& sc, uint32_t tri, SurfaceRaw& r) {
  const float4* fr = sc.tri_frames + size_t(tri) * kFrameFloat4s;
#pragma unroll
  for (int k = 1; k < 6; ++k) r.a[k] = __ldg(fr + k);
  r.q2 = __ldg(sc.tris + size_t(tri) * kTriFloat4s + 2);
}
__device__ __forceinline__ Surface query_surface_raw(const SurfaceRaw& raw, float u, float v, float t, float3 org, float3 dir) {
  const float4 a0 = raw.a[0], a1 = raw.a[1], a2 = raw.a[2], a3 = raw.a[3], a4 = raw.a[4], a5 = raw.a[5], a6 = raw.a[6];
  const float m[28] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w, a3.x, a3.y,
                       a3.z, a3.w, a4.x, a4.y, a4.z, a4.w, a5.x, a5.y, a5.z, a5.w, a6.x, a6.y, a6.z, a6.w};
  const float w = 1.f - u - v;
  Surface p;
  p.pos = org + dir * t;
  float c[9];
#pragma unroll
  for (int j = 0; j < 9; ++j) c[j] = w * m[j] + u * m[9 + j] + v * m[18 + j];
  float3 t0 = f3(c[0], c[1], c[2]), t1 = f3(c[3], c[4], c[5]), t2 = f3(c[6], c[7], c[8]);
  t1 = normalize(t1);
  t0 = t0 - dot(t0, t1) * t1;
  t0 = normalize(t0);
  t2 = t2 - dot(t2, t1) * t1 - dot(t2, t0) * t0;
  t2 = normalize(t2);
  p.tan.c0 = t0; p.tan.c1 = t1; p.tan.c2 = t2;
  // RayIsect::gnormal() = normalize(-Ng) with Embree's Ng = cross(e1, e2) (triangle.h:54)
  const float4 q1 = raw.q1, q2 = raw.q2;
  const float3 e1 = f3(q1.x, q1.y, q1.z), e2 = f3(q2.x, q2.y, q2.z);
  const float3 Ng = f3(e1.y * e2.z - e1.z * e2.y, e1.z * e2.x - e1.x * e2.z, e1.x * e2.y - e1.y * e2.x);
  float3 gn = normalize(-Ng);
  const float3 omega = normalize(-dir);
  gn = gn * (dot(omega, gn) < 0.0f ? -1.0f : 1.0f);
  p.gnormal = gn;
  p.material_id = __float_as_uint(m[27]);
  return p;
}
__device__ __forceinline__ Surface query_surface(const DeviceScene& sc, uint32_t tri, float u, float v, float t, float3 org,
                                                 float3 dir) {
  const SurfaceRaw raw = load_surface_raw(sc, tri);
  return query_surface_raw(raw, u, v, t, org, dir);
}

// Ray origin of Scene::intersect / intersectMesh (Scene.cpp:186-189): 1e-4 along +-gnormal
__device__ __forceinline__ float3 offset_origin(float3 pos, float3 gnormal, float3 direction) {
  return pos + (dot(gnormal, direction) > 0.0f ? 1.0f : -1.0f) * gnormal * 0.0001f;
}

// ---------------------------------------------------------------- Sample.inl
struct AngularBound { float theta_inf, theta_sup, phi_inf, phi_sup; };

// angular_bound (Sample.inl:5-37)
__device__ __forceinline__ AngularBound angular_bound(float3 center, float radius) {
  AngularBound b;
  b.theta_inf = 0.f; b.theta_sup = kHalfPi; b.phi_inf = 0.f; b.phi_sup = kTwoPi;
  const float lateral_distance_sq = center.x * center.x + center.z * center.z;
  const float distance_sq = lateral_distance_sq + center.y * center.y;
  const float radius_sq = radius * radius;
  if (radius_sq < distance_sq) {
    const float lateral_distance = sqrtf(lateral_distance_sq);
    const float distance = sqrtf(distance_sq);
    const float theta_center = cr_asinf(lateral_distance / distance);
    const float theta_radius = cr_asinf(radius / distance);
    if (lateral_distance_sq < radius_sq) {
      b.theta_sup = fminf(kHalfPi, theta_center + theta_radius);
    } else if (radius_sq < distance_sq) {
      b.theta_inf = theta_center - theta_radius;
      b.theta_sup = fminf(kHalfPi, theta_center + theta_radius);
      const float phi_center = cr_atan2f(center.z, center.x);
      const float phi_radius = cr_asinf(radius / lateral_distance);
      b.phi_inf = phi_center - phi_radius;
      b.phi_sup = phi_center + phi_radius;
    }
  }
  return b;
}

struct LambertRange { float theta_inf, theta_range, phi_inf, phi_range; };

// cos of a bound of angular_bound.  An emitter INSIDE the scene's bounding sphere (the usual case) has the unrestricted bounds 0 and
// pi/2 exactly, and every next-event sample evaluates this four times: the two known answers are constants —
// float(cos(double(0x3FC90FDB))) = 0xB33BBD2E, nowhere near a rounding tie, and cos(0) = 1 — the rest goes through cr_cosf.
__device__ __forceinline__ float cos_of_bound(float x) {
  if (x == kHalfPi) return __uint_as_float(0xB33BBD2Eu);
  if (x == 0.0f) return 1.0f;
  return cr_cosf(x);
}

// the general case out of line: three asin, one atan2 and two cos in double are ~850 instructions, and five inlined copies of them
// (emitter sampling, next-event sampling, emission, the LightBSDF cases of query and sample) were 40 % of the logic stage's code
__device__ B2PT_MATH_INLINE LambertRange lambert_range_general(float3 center, float radius) {
  const AngularBound bound = angular_bound(center, radius);
  LambertRange r;
  const float cos_sup = cos_of_bound(bound.theta_sup), cos_inf = cos_of_bound(bound.theta_inf);
  const float uniform_theta_inf = cos_sup * cos_sup;
  const float uniform_theta_sup = cos_inf * cos_inf;
  const float uniform_phi_inf = bound.phi_inf * kOneOverPi * 0.5f;
  const float uniform_phi_sup = bound.phi_sup * kOneOverPi * 0.5f;
  r.theta_inf = uniform_theta_inf;
  r.theta_range = uniform_theta_sup - uniform_theta_inf;
  r.phi_inf = uniform_phi_inf;
  r.phi_range = uniform_phi_sup - uniform_phi_inf;
  return r;
}
__device__ __forceinline__ LambertRange lambert_range(float3 center, float radius) {
  // angular_bound's own test (Sample.inl:13): a sphere that contains the point restricts nothing, and the same expressions on the
  // bounds (0, pi/2, 0, 2 pi) are constants
  const float lateral_distance_sq = center.x * center.x + center.z * center.z;
  const float distance_sq = lateral_distance_sq + center.y * center.y;
  const float radius_sq = radius * radius;
  if (radius_sq < distance_sq) return lambert_range_general(center, radius);
  LambertRange r;
  const float cos_sup = cos_of_bound(kHalfPi), cos_inf = cos_of_bound(0.0f);
  const float uniform_theta_inf = cos_sup * cos_sup;
  const float uniform_theta_sup = cos_inf * cos_inf;
  const float uniform_phi_inf = 0.0f * kOneOverPi * 0.5f;
  const float uniform_phi_sup = kTwoPi * kOneOverPi * 0.5f;
  r.theta_inf = uniform_theta_inf;
  r.theta_range = uniform_theta_sup - uniform_theta_inf;
  r.phi_inf = uniform_phi_inf;
  r.phi_range = uniform_phi_sup - uniform_phi_inf;
  return r;
}

// lambert_adjust (Sample.inl:119-131)
__device__ __forceinline__ float lambert_adjust(float3 center, float radius) {
  const LambertRange r = lambert_range(center, radius);
  return r.theta_range * r.phi_range;
}

// sample_lambert(generator, omega) (Sample.inl:48-56)
__device__ __forceinline__ float3 sample_lambert(Rng& rng, float3 omega) {
  const float y = sqrtf(rng.sample()) * sign_glm(omega.y);
  const float r = sqrtf(1.0f - y * y);
  const float phi = rng.sample() * 2.0f * kPi;
  return cr_polar(r, y, phi);
}

// sample_lambert(generator, sphere, omega) (Sample.inl:58-82): lobe restricted to the cone around a sphere
__device__ __forceinline__ float3 sample_lambert_bounded(Rng& rng, float3 center, float radius, float3 omega, float& adjust) {
  center.y *= sign_glm(omega.y);
  const LambertRange lr = lambert_range(center, radius);
  adjust = lr.theta_range * lr.phi_range;
  const float y = sqrtf(rng.sample() * lr.theta_range + lr.theta_inf) * sign_glm(omega.y);
  const float phi = kTwoPi * (rng.sample() * lr.phi_range + lr.phi_inf);
  const float r = sqrtf(1 - y * y);
  return cr_polar(r, y, phi);
}

// reflection_to_surface (Sample.inl:39-46)
__device__ __forceinline__ mat3 reflection_to_surface(float3 reflection) {
  mat3 m;
  m.c1 = reflection;
  m.c2 = normalize(f3(0.0f, 1.0f, 0.0f) - m.c1.y * m.c1);
  m.c0 = normalize(cross(m.c1, m.c2));
  return m;
}

// sample_phong(generator, omega, power) (Sample.inl:137-149)
__device__ __forceinline__ float3 sample_phong(Rng& rng, float3 omega, float power) {
  const mat3 refl_to_surf = reflection_to_surface(f3(-omega.x, omega.y, -omega.z));
  const float y = cr_powf(rng.sample(), 1.0f / (power + 1.0f));
  const float r = sqrtf(1.0f - y * y);
  const float phi = rng.sample() * 2.0f * kPi;
  return mul(refl_to_surf, cr_polar(r, y, phi));
}

// sample_phong(generator, omega, power, sphere) / phong_adjust (Sample.inl:151-205): Phong lobe restricted to the cone
// that subtends a sphere given in the local frame
struct PhongRange { mat3 reflection; float theta_inf, theta_range, phi_inf, phi_range; };

__device__ __forceinline__ PhongRange phong_range(float3 omega, float power, float3 center, float radius) {
  PhongRange r;
  r.reflection = reflection_to_surface(f3(-omega.x, omega.y, -omega.z));
  const AngularBound bound = angular_bound(mul(center, r.reflection), radius);  // sphere.center * reflection
  const float uniform_theta_inf = cr_powf(cr_cosf(bound.theta_sup), power + 1.0f);
  const float uniform_theta_sup = cr_powf(cr_cosf(bound.theta_inf), power + 1.0f);
  const float uniform_phi_inf = bound.phi_inf * kOneOverPi * 0.5f;
  const float uniform_phi_sup = bound.phi_sup * kOneOverPi * 0.5f;
  r.theta_inf = uniform_theta_inf;
  r.theta_range = uniform_theta_sup - uniform_theta_inf;
  r.phi_inf = uniform_phi_inf;
  r.phi_range = uniform_phi_sup - uniform_phi_inf;
  return r;
}
__device__ __forceinline__ float phong_adjust(float3 omega, float power, float3 center, float radius) {
  const PhongRange r = phong_range(omega, power, center, radius);
  return r.theta_range * r.phi_range;
}
__device__ __forceinline__ float3 sample_phong_bounded(Rng& rng, float3 omega, float power, float3 center, float radius) {
  const PhongRange pr = phong_range(omega, power, center, radius);
  const float y = cr_powf(rng.sample() * pr.theta_range + pr.theta_inf, 1.0f / (power + 1.0f));
  const float phi = kTwoPi * (rng.sample() * pr.phi_range + pr.phi_inf);
  const float r = sqrtf(1 - y * y);
  return mul(pr.reflection, cr_polar(r, y, phi));
}

// ---------------------------------------------------------------- BSDF::query (BSDF.cpp)
__device__ __forceinline__ BsdfQuery diffuse_query(const B2ptMaterial& m, float3 gnormal, float3 incident, float3 outgoing) {
  // DiffuseBSDF::_query (BSDF.cpp:291-304), arguments already in the local frame
  const float same_side = dot(incident, gnormal) * dot(outgoing, gnormal) > 0.0f ? 1.0f : 0.0f;
  BsdfQuery q;
  q.throughput = f3(m.diffuse[0], m.diffuse[1], m.diffuse[2]) * kOneOverPi * same_side;
  q.density = fabsf(outgoing.y * kOneOverPi) * same_side;
  q.density_rev = fabsf(incident.y * kOneOverPi) * same_side;
  q.finite = 1;
  return q;
}

__device__ __forceinline__ BsdfQuery phong_query(const B2ptMaterial& m, float3 incident, float3 outgoing, float same_side) {
  // PhongBSDF::_query (BSDF.cpp:349-391)
  const float diffuse_density_factor = m.diffuse_probability;
  const float specular_density_factor = 1.0f - m.diffuse_probability;
  const float diffuse_density = fabsf(outgoing.y * kOneOverPi);
  const float diffuse_density_rev = fabsf(incident.y * kOneOverPi);
  const float3 diffuse = f3(m.diffuse[0], m.diffuse[1], m.diffuse[2]) * kOneOverPi;
  const float half_over_pi = 0.5f * kOneOverPi;
  const float3 reflected = f3(-incident.x, incident.y, -incident.z);
  const float cos_alpha = clampf(dot(outgoing, reflected), 0.0f, 1.0f);
  const float cos_alpha_pow = cr_powf(cos_alpha, m.power);
  const float specular_density = (m.power + 1.0f) * half_over_pi * cos_alpha_pow;
  const float specular_density_rev = specular_density;
  const float3 specular = f3(m.specular[0], m.specular[1], m.specular[2]) * (m.power + 2.0f) * half_over_pi * cos_alpha_pow;
  BsdfQuery q;
  q.density = same_side * (specular_density * specular_density_factor + diffuse_density * diffuse_density_factor);
  q.density_rev = same_side * (specular_density_rev * specular_density_factor + diffuse_density_rev * diffuse_density_factor);
  q.throughput = same_side * (diffuse + specular);
  q.finite = 1;
  return q;
}

// The restriction factor of an emitter's lobe, lambert_adjust(sphere in the emitter's frame) (LightBSDF::query, BSDF.cpp:93-114): it
// depends on the point of the emitter only, not on the directions of a query, and it is the expensive part of the query (two asin,
// one atan2, four cos, all correctly rounded).  A stage that queries the same emitter point many times computes it once and hands it
// to bsdf_query; -1 = not an area emitter / not computed.
__device__ __forceinline__ float light_lobe_adjust(const DeviceScene& sc, const Surface& s) {
  const B2ptMaterial& m = sc.materials[s.material_index()];
  if (m.type != B2PT_BSDF_LIGHT) return -1.0f;
  const float3 center = s.to_surface(f3(m.sphere[0], m.sphere[1], m.sphere[2]) - s.pos);
  return lambert_adjust(center, m.sphere[3]);
}

// Scene::queryBSDF(surface, incident, outgoing) (Scene.cpp:142-149) -> BSDF::query of the surface's material
// light_adjust >= 0: light_lobe_adjust(sc, s) computed earlier for this surface (same function, same inputs: same bits)
__device__ __forceinline__ BsdfQuery bsdf_query(const DeviceScene& sc, const Surface& s, float3 incident, float3 outgoing, float light_adjust = -1.0f) {
  const B2ptMaterial& m = sc.materials[s.material_index()];
  BsdfQuery q;
  switch (m.type) {
    case B2PT_BSDF_DIFFUSE:  // DiffuseBSDF::query (BSDF.cpp:239-243)
      return diffuse_query(m, s.to_surface(s.gnormal), s.to_surface(incident), s.to_surface(outgoing));
    case B2PT_BSDF_PHONG: {  // PhongBSDF::query (BSDF.cpp:317-326)
      const float same_side = dot(incident, s.gnormal) * dot(outgoing, s.gnormal) > 0.0f ? 1.0f : 0.0f;
      return phong_query(m, s.to_surface(incident), s.to_surface(outgoing), same_side);
    }
    case B2PT_BSDF_LIGHT: {  // LightBSDF::query (BSDF.cpp:93-114)
      const float3 local_outgoing = s.to_surface(outgoing);
      const float adjust = light_adjust >= 0.0f ? light_adjust : lambert_adjust(s.to_surface(f3(m.sphere[0], m.sphere[1], m.sphere[2]) - s.pos), m.sphere[3]);
      q.throughput = local_outgoing.y > 0.0f ? f3(1.0f) : f3(0.0f);
      q.density = (local_outgoing.y > 0.0f ? 1.0f : 0.0f) * local_outgoing.y * kOneOverPi / adjust;
      q.density_rev = 0.0f;
      q.finite = 1;
      return q;
    }
    case B2PT_BSDF_SUN:  // sun_light_bsdf::query (BSDF.cpp:179-189)
      q.throughput = f3(0.0f); q.density = 1.0f; q.density_rev = 1.0f; q.finite = 1;
      return q;
    case B2PT_BSDF_CAMERA: {  // CameraBSDF::query (BSDF.cpp:210-222); pow(x, 1.0f) == x
      const float3 local_incident = s.to_surface(incident);
      q.throughput = f3(local_incident.y > 0.0f ? 1.0f : 0.0f) / fabsf(local_incident.y);
      q.density = 0.0f; q.density_rev = 1.0f; q.finite = 1;
      return q;
    }
    default:  // DeltaBSDF::query (BSDF.cpp:438-448): reflection / transmission
      q.throughput = f3(0.0f); q.density = 0.0f; q.density_rev = 0.0f; q.finite = 0;
      return q;
  }
}

// bsdf_query for the places a stage reaches rarely (a sun light's next-event sample, a surface that is not what its entity type says):
// one out-of-line copy instead of another inlined switch in the stage's hot code
__device__ B2PT_MATH_INLINE BsdfQuery bsdf_query_outofline(const DeviceScene& sc, Surface s, float3 incident, float3 outgoing) {
  return bsdf_query(sc, s, incident, outgoing);
}
// bsdf_query of a surface ON an emitter (entity type light: LightBSDF or sun_light_bsdf, AreaLights.cpp:25-33), the only materials a
// path that hits a light can find there
__device__ __forceinline__ BsdfQuery bsdf_query_emitter(const DeviceScene& sc, const Surface& s, float3 incident, float3 outgoing) {
  const B2ptMaterial& m = sc.materials[s.material_index()];
  BsdfQuery q;
  if (m.type == B2PT_BSDF_LIGHT) {  // LightBSDF::query (BSDF.cpp:93-114), as in bsdf_query
    const float3 local_outgoing = s.to_surface(outgoing);
    const float adjust = lambert_adjust(s.to_surface(f3(m.sphere[0], m.sphere[1], m.sphere[2]) - s.pos), m.sphere[3]);
    q.throughput = local_outgoing.y > 0.0f ? f3(1.0f) : f3(0.0f);
    q.density = (local_outgoing.y > 0.0f ? 1.0f : 0.0f) * local_outgoing.y * kOneOverPi / adjust;
    q.density_rev = 0.0f;
    q.finite = 1;
    return q;
  }
  if (m.type == B2PT_BSDF_SUN) {  // sun_light_bsdf::query (BSDF.cpp:179-189)
    q.throughput = f3(0.0f); q.density = 1.0f; q.density_rev = 1.0f; q.finite = 1;
    return q;
  }
  return bsdf_query_outofline(sc, s, incident, outgoing);
}

// CameraBSDF::sample (BSDF.cpp:197-208): no draws
__device__ __forceinline__ BsdfSample camera_sample(const Surface& s, float3 omega) {
  BsdfSample r;
  r.omega = -omega;
  r.throughput = f3(1.0f) / fabsf(dot(s.normal(), omega));
  r.density = 1.0f; r.density_rev = 0.0f; r.finite = 1;
  return r;
}

// Scene::sampleBSDF (Scene.cpp:133-140) -> BSDF::sample of the surface's material
__device__ __forceinline__ BsdfSample bsdf_sample(const DeviceScene& sc, Rng& rng, const Surface& s, float3 omega) {
  const B2ptMaterial& m = sc.materials[s.material_index()];
  BsdfSample r;
  switch (m.type) {
    case B2PT_BSDF_DIFFUSE: {  // DiffuseBSDF::sample (BSDF.cpp:245-262)
      const float3 local_omega = s.to_surface(omega);
      const float3 d = sample_lambert(rng, local_omega);
      const BsdfQuery q = diffuse_query(m, s.to_surface(s.gnormal), local_omega, d);
      r.omega = s.to_world(d);
      r.throughput = q.throughput; r.density = q.density; r.density_rev = q.density_rev; r.finite = q.finite;
      return r;
    }
    case B2PT_BSDF_PHONG: {  // PhongBSDF::sample (BSDF.cpp:328-347)
      const float3 local_omega = s.to_surface(omega);
      // sample_lambert (Sample.inl:48-56) or sample_phong (Sample.inl:137-149): both draw (u1, u2), make a height y from u1 and turn
      // (r, y, phi) into a direction — the same operations in the same order as the two functions above, but with ONE polar
      // conversion after the branch, so the double-precision sincos runs once per warp instead of once per lobe
      const bool diffuse_lobe = rng.sample() < m.diffuse_probability;
      const float u1 = rng.sample();
      const float y = diffuse_lobe ? sqrtf(u1) * sign_glm(local_omega.y) : cr_powf(u1, 1.0f / (m.power + 1.0f));
      const float rr = sqrtf(1.0f - y * y);
      const float phi = rng.sample() * 2.0f * kPi;
      float3 d = cr_polar(rr, y, phi);
      if (!diffuse_lobe) d = mul(reflection_to_surface(f3(-local_omega.x, local_omega.y, -local_omega.z)), d);
      r.omega = s.to_world(d);
      const float same_side = dot(omega, s.gnormal) * dot(r.omega, s.gnormal) > 0.0f ? 1.0f : 0.0f;
      const BsdfQuery q = phong_query(m, local_omega, d, same_side);
      r.throughput = q.throughput; r.density = q.density; r.density_rev = q.density_rev; r.finite = q.finite;
      return r;
    }
    case B2PT_BSDF_REFLECTION: {  // ReflectionBSDF::sample (BSDF.cpp:450-465)
      const float3 local_omega = s.to_surface(omega);
      r.throughput = f3(1.0f) / local_omega.y;
      r.omega = s.to_world(f3(-local_omega.x, local_omega.y, -local_omega.z));
      r.density = 1.0f; r.density_rev = 1.0f; r.finite = 0;
      return r;
    }
    case B2PT_BSDF_TRANSMISSION: {  // TransmissionBSDF::sample (BSDF.cpp:472-504)
      const float3 local_omega = s.to_surface(omega);
      float3 o;
      if (local_omega.y > 0.f) {
        const float eta = m.ext_over_int_ior;
        o = -eta * (local_omega - f3(0.0f, local_omega.y, 0.0f)) -
            f3(0.0f, sqrtf(1 - eta * eta * (1 - local_omega.y * local_omega.y)), 0.0f);
      } else {
        const float eta = 1.0f / m.ext_over_int_ior;
        o = -eta * (local_omega - f3(0.0f, local_omega.y, 0.0f)) +
            f3(0.0f, sqrtf(1 - eta * eta * (1 - local_omega.y * local_omega.y)), 0.0f);
      }
      r.throughput = f3(1.0f) / fabsf(o.y);
      r.omega = s.to_world(o);
      r.density = 1.0f; r.density_rev = 1.0f; r.finite = 0;
      return r;
    }
    case B2PT_BSDF_LIGHT: {  // LightBSDF::sample (BSDF.cpp:76-91)
      const float3 center = s.to_surface(f3(m.sphere[0], m.sphere[1], m.sphere[2]) - s.pos);
      float adjust;
      const float3 d = sample_lambert_bounded(rng, center, m.sphere[3], s.to_surface(omega), adjust);
      r.throughput = f3(1.0f);
      r.omega = s.to_world(d);
      r.density = fabsf(d.y) * kOneOverPi / adjust;  // lambert_density (Sample.inl:133-135)
      r.density_rev = 0.0f; r.finite = 1;
      return r;
    }
    case B2PT_BSDF_SUN:  // sun_light_bsdf::sample (BSDF.cpp:164-177)
      r.throughput = f3(1.0f); r.omega = omega; r.density = 1.0f; r.density_rev = 0.0f; r.finite = 1;
      return r;
    default:  // CameraBSDF::sample (BSDF.cpp:197-208)
      return camera_sample(s, omega);
  }
}

// ---------------------------------------------------------------- AreaLights (AreaLights.cpp)
struct LightSample {
  Surface surface;
  float3 radiance;
  float area_density, light_density;
  bool directional;
  __device__ __forceinline__ float combined_density() const { return area_density * light_density; }
};

// AreaLights::_sampleLight (AreaLights.cpp:216-221) on top of piecewise_sampler_t::sample (Sample.inl:245-247):
// libstdc++'s piecewise_constant_distribution<float>::operator() with the engine, one double per pick.
__device__ __forceinline__ uint32_t pick_light(const DeviceScene& sc, Rng& rng) {
  const double p = rng.sample_double();
  float sample;
  if (sc.num_lights == 1) {
    sample = float(p);  // _M_cp is empty for the single interval [0, 1]
  } else {
    uint32_t lo = 0, hi = sc.num_lights;  // std::lower_bound(cp, p)
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (sc.light_cp[mid] < p) lo = mid + 1; else hi = mid;
    }
    const uint32_t i = lo;
    const double pref = i > 0 ? sc.light_cp[i - 1] : 0.0;
    sample = float(double(sc.light_int[i]) + (p - pref) / sc.light_den[i]);
  }
  const uint32_t id = uint32_t(sample * float(sc.num_lights));
  return min(id, sc.num_lights - 1u);
}

// AreaLights::sample (AreaLights.cpp:121-140) + _samplePosition (:223-231).
// `vec2(generator.sample(), generator.sample())` is evaluated right to left by the reference build
// (g++ on x86-64; pinned by tests/test_oracle_pinning.py), so the FIRST draw is the y component.
__device__ __forceinline__ LightSample sample_light(const DeviceScene& sc, Rng& rng) {
  const uint32_t id = pick_light(sc, rng);
  const B2ptLight& l = sc.lights[id];
  const float s1 = rng.sample();
  const float s0 = rng.sample();
  const float ux = (s0 - 0.5f) * l.size[0];
  const float uy = (s1 - 0.5f) * l.size[1];
  const float3 left = f3(l.tangent[0], l.tangent[1], l.tangent[2]);
  const float3 nrm = f3(l.tangent[3], l.tangent[4], l.tangent[5]);
  const float3 up = f3(l.tangent[6], l.tangent[7], l.tangent[8]);
  LightSample r;
  r.surface.pos = f3(l.position[0], l.position[1], l.position[2]) + ux * left + uy * up;
  r.surface.tan.c0 = left; r.surface.tan.c1 = nrm; r.surface.tan.c2 = up;
  r.surface.gnormal = nrm;
  r.surface.material_id = l.material_id;
  r.radiance = f3(l.exitance[0], l.exitance[1], l.exitance[2]) * kOneOverPi;
  r.area_density = 1.0f / (l.size[0] * l.size[1]);
  r.light_density = l.weight;
  r.directional = l.diffuse == 0.0f;
  return r;
}

// Scene::queryLSDF (Scene.cpp:128-131) -> AreaLights::queryLSDF (AreaLights.cpp:142-155)
__device__ __forceinline__ void query_lsdf(const DeviceScene& sc, const Surface& s, float3 omega, float3& radiance, float& density) {
  const uint32_t light_id = sc.materials[s.material_index()].light_id;
  const B2ptLight& l = sc.lights[light_id];
  const float cos_theta = dot(omega, f3(l.tangent[3], l.tangent[4], l.tangent[5]));
  radiance = f3(l.exitance[0], l.exitance[1], l.exitance[2]) * kOneOverPi * (cos_theta > 0.0f ? 1.0f : 0.0f);
  density = l.weight / (l.size[0] * l.size[1]);
}

// ---------------------------------------------------------------- cameras (Cameras.cpp, Technique.cpp)
// Technique::_camera_surface (Technique.cpp:107-116)
__device__ __forceinline__ Surface camera_surface(const B2ptCamera& c) {
  const float3 v0 = f3(c.view_to_world[0], c.view_to_world[1], c.view_to_world[2]);
  const float3 v1 = f3(c.view_to_world[3], c.view_to_world[4], c.view_to_world[5]);
  const float3 v2 = f3(c.view_to_world[6], c.view_to_world[7], c.view_to_world[8]);
  Surface s;
  s.pos = f3(c.position[0], c.position[1], c.position[2]);
  s.tan.c0 = v1; s.tan.c1 = -v2; s.tan.c2 = v0;
  s.material_id = 0u << 2 | B2PT_ENTITY_CAMERA;
  s.gnormal = -v2;
  return s;
}

// ray_direction (Cameras.cpp:120-127) rotated to world space (Technique.cpp:330)
__device__ __forceinline__ float3 camera_ray_direction(const B2ptCamera& c, float px, float py, float width, float res_y_inv) {
  const float x = px * res_y_inv * 2.0f - width * res_y_inv;
  const float y = py * res_y_inv * 2.0f - 1.0f;
  const float3 d = normalize(f3(x, y, -c.focal_length_y));
  mat3 m;
  m.c0 = f3(c.view_to_world[0], c.view_to_world[1], c.view_to_world[2]);
  m.c1 = f3(c.view_to_world[3], c.view_to_world[4], c.view_to_world[5]);
  m.c2 = f3(c.view_to_world[6], c.view_to_world[7], c.view_to_world[8]);
  return mul(m, d);
}

// pixel_position (Cameras.cpp:134-149)
__device__ __forceinline__ float2 pixel_position(float3 direction, float width, float height, float res_y_inv, float flen) {
  const float factor = flen / -direction.z;
  float x = direction.x * factor;
  float y = direction.y * factor;
  y = (y + 1.0f) * height * 0.5f;
  x = (x + width * res_y_inv) * height * 0.5f;
  return make_float2(x, y);
}

}  // namespace b2pt
