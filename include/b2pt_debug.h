/*
 * b2pt_debug.h — parity probes of libb2pt.so.  NOT part of the drop-in boundary (include/b2pt.h):
 * these entry points expose single device functions of the hot path so that tests/ can compare them
 * with the reference's own functions on the same inputs (oracle/_ref/ref_harness density | pinning | camera).
 * file:line cite /root/reference.
 */
#ifndef B2PT_DEBUG_H
#define B2PT_DEBUG_H

#include "b2pt.h"

#ifdef __cplusplus
extern "C" {
#endif

/*
 * One density-estimation loop: BSDF::gathering_density (BSDF.cpp:42-66; DiffuseBSDF :264-289, PhongBSDF :393-434) with
 * Intersector::intersectFast (Intersector.cpp:17-46) as the trial.  The surface of a job is the closest MESH hit of its
 * ray (Scene::intersectMesh's RTCRay, Scene.cpp:204-226), omega = -dir; the merge sphere is (target, radius).
 * Trial t of the loop draws from the stream (key, t) — upg_density_* in csrc/vcm.cu run exactly this function.
 */
typedef struct { float org[3]; float dir[3]; float target[3]; float radius; uint32_t key; } B2ptDensityJob;
/* trials: the N at which the loop returned (0: the ray missed or the surface has no lobe); density = N / adjust;
   kind: 0 none, 1 DiffuseBSDF, 2 PhongBSDF; t: hit distance of the job's ray */
typedef struct { uint32_t trials; float density; uint32_t kind; float t; } B2ptDensityResult;
/* jobs / out in HOST memory */
int b2pt_debug_density(b2pt_scene_t scene, const B2ptDensityJob* jobs, size_t n, B2ptDensityResult* out);

/*
 * Primary rays as Technique::_for_each_ray shoots them (Technique.cpp:321-331): for pixel[i], sample[i] the two jitter
 * draws of stream (params->seed, pixel, sample) and ray_direction (Cameras.cpp:120-127) rotated to world space.
 * directions: n * 3 floats (HOST).
 */
int b2pt_debug_primary(const B2ptRenderParams* params, const uint32_t* pixels, const uint32_t* samples, size_t n,
                       float* directions);

/*
 * ray_direction (Cameras.cpp:120-127) at explicit image positions (n * 2 floats, HOST) and pixel_position
 * (Cameras.cpp:134-149) of world_to_view * direction: directions n * 3, pixels n * 2 floats (HOST).
 */
int b2pt_debug_camera(const B2ptRenderParams* params, const float* positions, size_t n, float* directions, float* pixels);

/*
 * AreaLights::sample (AreaLights.cpp:121-140, 216-231) with the stream keys[i]: positions n * 3 floats, material_ids and
 * draws (engine calls consumed) n uint32 each (HOST).
 */
int b2pt_debug_light_samples(b2pt_scene_t scene, const uint32_t* keys, size_t n, float* positions, uint32_t* material_ids,
                             uint32_t* draws);

/*
 * The engine itself: raw[i] = draw i of stream `key`, floats[i] = the i-th uniform_real_distribution<float> draw,
 * doubles[i] = the i-th generate_canonical<double, 53> (two engine calls each); n values each (HOST).
 */
int b2pt_debug_draws(uint32_t key, size_t n, uint32_t* raw, float* floats, double* doubles);

/*
 * Tree statistics of the scene's BVH and, optionally, per-ray traversal counts.
 * stats[0] = SAH cost of the 8-wide tree (sum over nodes of area(node) / area(root) * (1 per child box test * 8 ... see DESIGN §4),
 * stats[1] = inner nodes, stats[2] = leaf children, stats[3] = triangle references, stats[4] = sum over leaves of
 * area(leaf) / area(root) * triangles, stats[5] = sum over inner nodes of area(node) / area(root).
 */
int b2pt_debug_tree_stats(b2pt_scene_t scene, double* stats6);
/* rays in HOST memory; counts: n * 2 uint32 (node visits, triangle tests) of the one-ray-per-thread traversal */
int b2pt_debug_trace_counts(b2pt_scene_t scene, const B2ptRay* rays, size_t n, unsigned mode, uint32_t* counts);

/* The persisting-L2 window over the BVH block (csrc/scene.cu): out[0] = bytes of the BVH block (nodes + triangle records),
   out[1] = window bytes attached to the render streams (0: off), out[2] = hit ratio, out[3] = the device's
   persistingL2CacheMaxSize, out[4] = its accessPolicyMaxWindowSize, out[5] = its l2CacheSize. */
int b2pt_debug_l2_window(b2pt_scene_t scene, double* out6);

/* The shading stages' powf (csrc/common.cuh: cr_powf — Phong lobes, PhongBSDF::_query BSDF.cpp:349-391, sample_phong Sample.inl:137-149)
 * on n host (x, y) pairs: `table` = the table-driven evaluation that runs (csrc/pow_core.h), `lib` = float(exp2(y log2 x)) through the CUDA
 * math library, which it replaced.  tests/test_pow_gpu.py sets both beside the host build of the same header and libquadmath. */
int b2pt_debug_pow(const float* x, const float* y, size_t n, float* table, float* lib);

#ifdef __cplusplus
}
#endif
#endif /* B2PT_DEBUG_H */
