/*
 * b2pt.h — C ABI of the B200 wavefront path-tracing core.
 *
 * This is the drop-in boundary for the per-pixel sampling hot path of the
 * ciechowoj/master thesis renderer.  Every entry point replaces one piece of the
 * reference's host interface (file:line cite /root/reference):
 *
 *   b2pt_scene_create      <- Scene::Scene + Scene::buildAccelStructs (Scene.cpp:9-73):
 *                             takes the already-loaded meshes / materials / lights,
 *                             copies them to the device and builds the BVH there
 *                             (replaces rtcNewTriangleMesh / rtcCommit).
 *   b2pt_trace[_device]    <- Scene::intersect / intersectMesh / occluded's Embree
 *                             calls rtcIntersect / rtcOccluded (Scene.cpp:151-226).
 *   b2pt_render            <- Technique::render (Technique.cpp:15-77): one or more
 *                             1-spp passes of PT (PT.cpp:15-120), BPT (BPT.cpp:14-324)
 *                             or VCM/UPG (UPG.cpp:40-831) over the whole window,
 *                             accumulated as (sum r, sum g, sum b, sum 1) per pixel
 *                             like Technique::_commit_images (Technique.cpp:194-244).
 *   b2pt_film_*            <- subimage_view_t's dvec4 buffer (ImageView.hpp:10-46),
 *                             kept resident on the device between passes.
 *   b2pt_comm_* /
 *   b2pt_reduce_accum      <- merge_exr (Options.cpp:1340-1409): the sum of independently
 *                             rendered accumulators, as ONE NCCL reduce of the device films.
 *   b2pt_grid_*            <- v3::HashGrid3D<T>::build / rQuery (HashGrid3D.hpp:181-325).
 *   b2pt_last_error        <- the C++ exceptions the reference throws
 *                             (runtime_assert.cpp:7-11, Scene.cpp:55-57).
 *
 * Plain pointers and sizes only; no C++ or torch types.  All functions return 0
 * on success, non-zero on failure (message via b2pt_last_error()).  A handle is
 * single-caller (the reference calls Technique::render from one thread,
 * Application.cpp:41-79).  There is no CPU fallback: without a CUDA device every
 * compute entry point fails with B2PT_ERR_NO_DEVICE.
 */
#ifndef B2PT_H
#define B2PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2PT_ABI_VERSION 3

enum {
  B2PT_OK = 0,
  B2PT_ERR_INVALID = 1,   /* bad argument (the reference's runtime_assert) */
  B2PT_ERR_NO_DEVICE = 2, /* no CUDA device / driver */
  B2PT_ERR_CUDA = 3,      /* a CUDA call failed; see b2pt_last_error() */
  B2PT_ERR_CAPACITY = 4   /* a stated cap was exceeded (device pool, traversal stack, > 2^28 triangles) */
};

/* entity type in the low two bits of material_id (SurfacePoint.hpp:8-17) */
enum { B2PT_ENTITY_CAMERA = 0, B2PT_ENTITY_MESH = 1, B2PT_ENTITY_LIGHT = 2, B2PT_ENTITY_EMPTY = 3 };

/* BSDF classes of BSDF.hpp:54-174 */
enum {
  B2PT_BSDF_CAMERA = 0,
  B2PT_BSDF_DIFFUSE = 1,
  B2PT_BSDF_PHONG = 2,
  B2PT_BSDF_REFLECTION = 3,
  B2PT_BSDF_TRANSMISSION = 4,
  B2PT_BSDF_LIGHT = 5, /* LightBSDF: diffuse area emitter */
  B2PT_BSDF_SUN = 6    /* sun_light_bsdf: directional emitter */
};

/* One entry of Materials::bsdfs (Materials.hpp:14-27), flattened. 64 bytes. */
typedef struct {
  uint32_t type;              /* B2PT_BSDF_* */
  float diffuse[3];           /* DiffuseBSDF::_diffuse / PhongBSDF::_diffuse */
  float specular[3];          /* PhongBSDF::_specular */
  float power;                /* PhongBSDF::_power */
  float diffuse_probability;  /* PhongBSDF::_diffuse_probability (BSDF.cpp:306-315) */
  float ext_over_int_ior;     /* TransmissionBSDF::externalOverInternalIOR */
  float internal_ior;         /* TransmissionBSDF::internalIOR */
  uint32_t light_id;          /* LightBSDF::_light_id / sun_light_bsdf::_light_id */
  float sphere[4];            /* LightBSDF::_sphere: centre xyz, radius */
} B2ptMaterial;

/* AreaLight (AreaLights.hpp:44-60) + its pick weight (AreaLights.cpp:203-214). 80 bytes. */
typedef struct {
  float position[3];
  float tangent[9];   /* column-major mat3: col0 = left, col1 = normal, col2 = up */
  float size[2];
  float exitance[3];
  float diffuse;      /* 1 = area light, 0 = directional */
  uint32_t material_id; /* encoded: index << 2 | B2PT_ENTITY_LIGHT */
  float weight;       /* power / total power */
} B2ptLight;

/*
 * The scene as haste::Scene holds it after loadScene (loader.cpp:458-487):
 * per-mesh vertex / tangent-frame / index arrays (AreaLights.hpp:14-20), light
 * quads already appended as 2-triangle meshes (loader.cpp:434-456).
 * Mesh m owns vertices [vertex_offset[m], vertex_offset[m+1]) and indices
 * [index_offset[m], index_offset[m+1]) (mesh-local vertex numbers, 3 per
 * triangle).  geomID == mesh number, primID == triangle number inside the mesh,
 * as Scene.cpp:56-63 asserts.
 */
typedef struct {
  uint32_t num_meshes;
  const uint32_t* mesh_material_id;   /* [num_meshes] encoded material ids */
  const uint32_t* mesh_vertex_offset; /* [num_meshes + 1] */
  const uint32_t* mesh_index_offset;  /* [num_meshes + 1] */
  const float* vertices;              /* [num_vertices * 3] */
  const float* tangents;              /* [num_vertices * 9] column-major mat3 per vertex:
                                         col0 = bitangent, col1 = normal, col2 = tangent */
  const int32_t* indices;             /* [num_indices] */
  uint32_t num_materials;
  const B2ptMaterial* materials;
  uint32_t num_lights;
  const B2ptLight* lights;
} B2ptSceneDesc;

/* render_context_t (Technique.hpp:14-27) for one camera at one resolution. */
typedef struct {
  float view_to_world[9]; /* column-major mat3 */
  float world_to_view[9];
  float position[3];
  float focal_length_y;
} B2ptCamera;

enum { B2PT_TECHNIQUE_PT = 0, B2PT_TECHNIQUE_BPT = 1, B2PT_TECHNIQUE_VCM = 2, B2PT_TECHNIQUE_UPG = 3 };

/* Options fields the techniques read (Options.hpp:19-71, make_technique.cpp:86-170). */
typedef struct {
  uint32_t technique;     /* B2PT_TECHNIQUE_* */
  float beta;             /* MIS exponent; 0, 1, 2 use FixedBeta (Beta.hpp:24-36), else pow() */
  float roulette;         /* Options::roulette, default 0.9 */
  float lights;           /* Options::lights, default 1 */
  uint32_t max_path;      /* PT only (PT.hpp:24); 0 = unlimited */
  float radius;           /* VCM/UPG initial radius */
  float alpha;            /* VCM radius shrink exponent */
  uint32_t num_photons;   /* VCM/UPG light paths per pass; 0 = width*height */
  uint32_t unbiased, enable_vc, enable_vm, from_light;
  uint32_t width, height;
  B2ptCamera camera;
  uint32_t seed;          /* base seed of the per-(pixel, sample) RNG streams */
  uint32_t first_sample;  /* index of the first pass rendered by this call */
  uint32_t sample_stride; /* pass s of this call has index first_sample + s*sample_stride
                             (multi-GPU: rank g of G uses first = g, stride = G); 0 means 1 */
  uint32_t num_samples;   /* passes (samples per pixel) to add */
  float sky_horizon[3], sky_zenith[3]; /* Technique::set_sky_gradient */
  uint32_t max_subpath;   /* cap on stored light-subpath vertices per path; 0 = the reference's 1024
                             (BPT.hpp:29, UPG.hpp:54) */
} B2ptRenderParams;

/* statistics_t fields the GPU technique can fill (statistics.hpp, Technique.cpp:52-76) */
typedef struct {
  uint64_t num_samples;      /* passes done by the call */
  uint64_t num_basic_rays;   /* closest-hit rays (Scene::numNormalRays) */
  uint64_t num_shadow_rays;  /* occlusion rays (Scene::numShadowRays) */
  uint64_t num_paths;        /* pixel-samples */
  uint64_t numeric_errors;   /* non-finite pixel samples dropped (Technique.cpp:224-230) */
  uint64_t capacity_events;  /* light vertices / connections that did not fit the device pools; always 0 in a
                                call that returned B2PT_OK (a non-zero count fails the call with B2PT_ERR_CAPACITY) */
  uint64_t kernel_launches;  /* CUDA kernels launched by the call */
  float device_ms;           /* CUDA-event time of the call on its stream */
  /* Per-stage device time, CUDA events around every launch of the stage on the call's stream.
     Filled only when the environment variable B2PT_PROFILE_STAGES=1 is set (the extra events
     cost a little); 0 otherwise. */
  float trace_ms;            /* closest-hit traversal kernels (Scene::intersect / intersectMesh) */
  float shadow_ms;           /* occlusion traversal kernels (Scene::occluded) */
  float shade_ms;            /* everything else: raygen, shading, connection, splat, grid stages */
  uint64_t trace_launches;   /* launches behind trace_ms */
  uint64_t shadow_launches;  /* launches behind shadow_ms */
} B2ptStats;

typedef struct b2pt_scene_s* b2pt_scene_t;

/* RTCRay's fields the reference sets (Scene.cpp:162-171). 32 bytes. */
typedef struct { float org[3]; float tnear; float dir[3]; float tfar; } B2ptRay;
/* RTCRay's outputs: tfar, u, v, geomID, primID (RayIsect.hpp:8-27). geom == 0xFFFFFFFF on a miss. */
typedef struct { float t, u, v; uint32_t geom, prim; } B2ptHit;

enum {
  B2PT_TRACE_CLOSEST_ALL = 0,  /* rtcIntersect, mask 0xFFFFFFFF (Scene::intersect) */
  B2PT_TRACE_CLOSEST_MESH = 1, /* rtcIntersect, mask 1<<mesh   (Scene::intersectMesh) */
  B2PT_TRACE_OCCLUDED_MESH = 2 /* rtcOccluded,  mask 1<<mesh   (Scene::occluded); geom = 0 if blocked */
};

const char* b2pt_last_error(void);
int b2pt_abi_version(void);
/* number of visible CUDA devices, 0 if none */
int b2pt_device_count(void);
/* makes `device` current for the CALLING host thread (cudaSetDevice): a multi-GPU host uses one thread per device, and every
   handle (scene, film, communicator) belongs to the device that was current when it was created */
int b2pt_set_device(int device);
/* waits for everything queued on the current device */
int b2pt_device_synchronize(void);

int b2pt_scene_create(const B2ptSceneDesc* desc, b2pt_scene_t* out);
void b2pt_scene_destroy(b2pt_scene_t scene);
/* triangles, BVH nodes, BVH bytes, build milliseconds (device time) */
int b2pt_scene_info(b2pt_scene_t scene, uint64_t* num_triangles, uint64_t* num_nodes,
                    uint64_t* bvh_bytes, float* build_ms);

/* rays / hits in HOST memory; copies both ways */
int b2pt_trace(b2pt_scene_t scene, const B2ptRay* rays, size_t n, unsigned mode, B2ptHit* hits);
/* rays / hits already in DEVICE memory; asynchronous on `stream` (a cudaStream_t, may be NULL).  Batches of >= 2^18 rays run
   as persistent warps that fetch rays through ONE counter owned by the scene: calls on the same scene must be ordered with
   each other (same stream, or events between streams).  When two triangles are hit at exactly the same distance such a batch
   reports the one with the smaller position in the scene's leaf order; a small batch reports the one traversal met first. */
int b2pt_trace_device(b2pt_scene_t scene, const B2ptRay* rays_dev, size_t n, unsigned mode,
                      B2ptHit* hits_dev, void* stream);

/*
 * Adds params->num_samples passes into accum_dev: width*height dvec4
 * (sum r, sum g, sum b, sum 1), row-major, y up (ImageView.hpp:10-46), DEVICE
 * memory.  path_radiance_dev (nullable, device, num_samples*width*height*3
 * floats) receives the per-pixel-sample eye-path radiance that
 * Technique::_for_each_ray adds to _eye_image (Technique.cpp:338).
 * Launches on `stream` (a cudaStream_t, may be NULL) and returns when the passes are complete: the call
 * synchronises with the stream (the integrators read their device counters between bursts of launches).
 * Fails with B2PT_ERR_CAPACITY — film incomplete — if a BVH traversal stack or a vertex pool overflowed
 * during THIS call (the flags are reset per call).
 */
int b2pt_render(b2pt_scene_t scene, const B2ptRenderParams* params, double* accum_dev,
                float* path_radiance_dev, void* stream, B2ptStats* stats);

/*
 * Same, through HOST memory: view_host (width*height dvec4) += the passes.  This is what
 * GpuTechnique::render calls — one pass per call in the reference (Application.cpp:66); the
 * device->host copy is inside.  Passes are pure functions of (seed, sample index), so the
 * library renders AHEAD of a caller that asks for consecutive sample indices with otherwise
 * unchanged parameters (csrc/host_pipeline.cu): what a call adds is exactly the passes it names.
 * num_samples = 0 waits for passes being rendered ahead and drops them (a flush);
 * B2PT_HOST_LOOKAHEAD=0 in the environment turns rendering ahead off.
 */
int b2pt_render_host(b2pt_scene_t scene, const B2ptRenderParams* params, double* view_host,
                     B2ptStats* stats);

/*
 * The error record of a pass (Technique.cpp:68-74) on the DEVICE film: rms_abs_errors (ImageView.cpp:60-85) of
 * accum_dev (width*height dvec4) against reference_dev (width*height vec3, floats, same row order), or, with
 * window = {center_x, center_y, radius}, rms_abs_errors_windowed (ImageView.cpp:87-117) over the pixels
 * max(0, c - r + 1) <= p < min(c + r, size) — what Technique::_make_measurements evaluates per trace point.
 * Per pixel the arithmetic is the reference's (d = |vec3(a.xyz / a.w) - b| in float, d.x + d.y + d.z, dot(d, d));
 * the sums over pixels are taken in double (the reference adds them in float, in row order), so the results agree
 * with the reference's to float summation error, not bit for bit.  out (HOST): out[0] = rms, out[1] = abs.
 * The whole image divides by width*height*3, a window by its pixel count, as the reference does.
 */
int b2pt_film_errors(const double* accum_dev, const float* reference_dev, uint32_t width, uint32_t height,
                     const int32_t* window, float* out, void* stream);

/*
 * b2pt_film_errors for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  It has no counterpart in the reference.  Same per-pixel
 * arithmetic and the same (rms, abs) as b2pt_film_errors, with three differences:
 *   - it calls no cudaMalloc / cudaFree / allocating routine and never synchronises: one cudaMemsetAsync of 256 bytes and one kernel on
 *     `stream`; out_dev is a DEVICE buffer of two floats (out_dev[0] = rms, out_dev[1] = abs), valid once the stream reaches that point;
 *   - its scratch (one partial sum per 4096-pixel tile) is a file-scope __device__ array of 98 560 bytes inside libb2pt.so: every GPU
 *     that loads the library holds it until the process exits.  The number of pixels SUMMED (the window's, else width * height) must not
 *     exceed B2PT_FILM_ERRORS_STATIC_MAX_PIXELS; beyond it the call fails with B2PT_ERR_CAPACITY and launches nothing;
 *   - every call on a device shares that array: calls that overlap on the GPU (different streams with no event between them) race.
 *     Calls ordered on one stream, or by events, are safe; what must start at zero is cleared by every call.
 * The partial sums are added in a fixed order, so the result is deterministic; it agrees with b2pt_film_errors to the rounding of
 * a double sum (1e-6 relative in the float results is the tests' bound).
 */
#define B2PT_FILM_ERRORS_STATIC_MAX_PIXELS (1u << 24)
int b2pt_film_errors_static(const double* accum_dev, const float* reference_dev, uint32_t width, uint32_t height,
                            const int32_t* window, float* out_dev, void* stream);

/* ---- the film on the device: subimage_view_t's dvec4 accumulator (ImageView.hpp:10-46), width*height*(sum r, sum g, sum b, sum 1),
 * resident in HBM between passes; it crosses PCIe only when the host asks. ---- */
typedef struct b2pt_film_s* b2pt_film_t;
int b2pt_film_create(uint32_t width, uint32_t height, b2pt_film_t* out); /* zeroed, on the current CUDA device */
void b2pt_film_destroy(b2pt_film_t film);
double* b2pt_film_data(b2pt_film_t film);                /* the DEVICE pointer b2pt_render / b2pt_reduce_accum / b2pt_film_errors take */
int b2pt_film_clear(b2pt_film_t film, void* stream);
/* device -> host: view_host = film (add == 0) or view_host += film (add != 0); synchronises `stream` */
int b2pt_film_read(b2pt_film_t film, double* view_host, int add, void* stream);

/* ---- multi-GPU: samples split over a replicated scene, ONE reduce of the films.
 * merge_exr (Options.cpp:1340-1409) adds the RGBA accumulators of independently rendered images, count channel included;
 * b2pt_reduce_accum is that sum over the ranks of a communicator, on the device, over NVLink (ncclAllReduce / ncclReduce, float64).
 * Rank g of G renders the sample indices g, g + G, ... (B2ptRenderParams::first_sample = g, sample_stride = G).
 * NCCL is loaded at run time (libnccl.so.2, or the file named by B2PT_NCCL_LIB); nranks == 1 needs no NCCL at all. ---- */
typedef struct b2pt_comm_s* b2pt_comm_t;
#define B2PT_COMM_ID_BYTES 128
/* rank 0 makes the id (ncclGetUniqueId) and hands the 128 bytes to the other ranks by any means (file, socket, launcher) */
int b2pt_comm_unique_id(void* id);
/* one communicator per rank, on the CURRENT CUDA device; collective: returns when all nranks ranks have called it */
int b2pt_comm_init(int nranks, int rank, const void* id, b2pt_comm_t* out);
/* accum_dev (num_doubles float64, device) = sum over ranks, in place; root = -1: every rank gets the sum, else only `root`.
   Asynchronous on `stream`. */
int b2pt_reduce_accum(b2pt_comm_t comm, double* accum_dev, size_t num_doubles, int root, void* stream);
void b2pt_comm_destroy(b2pt_comm_t comm);

/* ---- photon index: v3::HashGrid3D<T>::build / rQuery (HashGrid3D.hpp:181-325) ---- */
typedef struct b2pt_grid_s* b2pt_grid_t;
/* points: n * 3 floats (HOST).  Builds the sorted cell grid on the device. */
int b2pt_grid_build(const float* points, size_t n, float radius, b2pt_grid_t* out);
void b2pt_grid_destroy(b2pt_grid_t grid);
/*
 * For each of nq query points (HOST, nq*3 floats): the indices i of all points with
 * distance2(point_i, q) < radius^2 (strict, HashGrid3D.hpp:196).  counts[q] receives the
 * number found; indices[q*max_per_query ...] the first max_per_query of them (unordered).
 */
int b2pt_grid_query(b2pt_grid_t grid, const float* queries, size_t nq, uint32_t max_per_query,
                    uint32_t* counts, uint32_t* indices);
/* the same with queries / counts / indices already in DEVICE memory; asynchronous on `stream` */
int b2pt_grid_query_device(b2pt_grid_t grid, const float* queries_dev, size_t nq, uint32_t max_per_query,
                           uint32_t* counts_dev, uint32_t* indices_dev, void* stream);
/* points indexed (light-origin vertices left out, HashGrid3D.hpp:253), cell size, device milliseconds of the build */
int b2pt_grid_info(b2pt_grid_t grid, uint64_t* num_points, float* radius, float* build_ms);

#ifdef __cplusplus
}
#endif
#endif /* B2PT_H */
