// ref_harness.cpp — TEST INFRASTRUCTURE: headless driver around the UNMODIFIED reference
// (ciechowoj/master @ /root/reference).  Built only by oracle/build_ref.sh into
// oracle/_ref/ref_harness; never linked into the product library.
//
// It links the reference's own translation units (Technique, PT, BPT, UPG, Scene, BSDF,
// AreaLights, Cameras, loader, ... — see build_ref.sh) against the vendored Embree 2.12 and
// Assimp fork, compiled with -fno-access-control so the harness can own the pixel loop and
// read private state, and with -include rng_shim.h so every (pixel, sample) has its own
// random stream.  All light transport, intersection and sampling code executed below is the
// reference's.
//
// Sub-commands:
//   dump   <scene.blend> <out.b2s> [--light px py pz  dx dy dz  ux uy uz  er eg eb  sx sy]
//   rays   <scene> <mode 0|1|2> <rays.bin> <hits.bin>       (B2ptRay in, B2ptHit out)
//   render <scene> key=value ...   per-(pixel,sample) streams, 1 thread, golden outputs
//   time   <scene> key=value ...   the reference's own Technique::render loop, N threads
//   grid   <in.bin> <out.bin>      v3::HashGrid3D build + rQuery on explicit points / queries
//   density <scene> <jobs.bin> <out.bin>   BSDF::gathering_density + Intersector::intersectFast per job (trial counts)
//   errors <accum.f64> <ref.f32> W H [cx cy r]   rms_abs_errors / rms_abs_errors_windowed (ImageView.cpp:60-117)
//   pinning <scene> key=value ...  raw draws, float / piecewise draws, primary rays of _for_each_ray, light samples
//   camera                         ray_direction / pixel_position / lookAt frames at the in-source unittest() inputs
//   rawdump <scene.blend> <out.b2i>  what the Assimp fork hands to loadScene (loader.cpp:458-487), before the loader's own math
// <scene> is a .blend (needs the reference's models/) or a .b2s written by `dump`.
//
// -DB2PT_STOCK_TIMING (oracle/build_ref.sh, second binary oracle/_ref/ref_harness_stock): the same driver WITHOUT the RNG shim
// (std::mt19937 as the reference ships), compiled like the reference's own Makefile:27 (-O2, AVX2 + FMA) and linked with the AVX2
// Embree (Makefile:30-37) — the timing-only CPU baseline; commands that need per-(pixel, sample) streams are left out.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <string>
#include <thread>

#include <xmmintrin.h>
#include <pmmintrin.h>

#include <assimp/postprocess.h>
#include <assimp/scene.h>
#include <assimp/Importer.hpp>

#include <BPT.hpp>
#include <PT.hpp>
#include <UPG.hpp>
#include <HashGrid3D.hpp>
#include <Scene.hpp>
#include <loader.hpp>


using namespace haste;
using std::string;

#include "../master_b200/host/b2s_io.hpp"
using namespace b2pt_host;

static bool ends_with(const string& s, const string& e) {
  return s.size() >= e.size() && s.compare(s.size() - e.size(), e.size(), e) == 0;
}

static shared<Scene> open_scene(const string& path) {
  shared<Scene> scene = ends_with(path, ".b2s") ? read_b2s(path) : loadScene(path);
  static RTCDevice device = rtcNewDevice(NULL);
  scene->buildAccelStructs(device);
  return scene;
}

// Adds one area light the way loader.cpp:434-456 does (for scenes that ship without one).
static shared<Scene> with_extra_light(const string& path, const float* a) {
  // load_lights needs mutable pieces, so redo loadScene's steps around the extra light
  shared<Scene> base = loadScene(path);
  Cameras cameras = base->_cameras;
  vector<Mesh> meshes = base->meshes;
  Materials materials;
  for (size_t i = 0; i < base->materials.bsdfs.size(); ++i) materials.names.push_back(base->materials.names[i]);
  // BSDFs are move-only; steal them from the (about to be dropped) base scene
  Materials& src = const_cast<Materials&>(base->materials);
  for (auto& b : src.bsdfs) materials.bsdfs.push_back(std::move(b));
  AreaLights lights = base->lights;
  bounding_sphere_t sphere = {vec3(0.0f), 0.0f};
  {  // compute_bounding_sphere (loader.cpp:408-432) over the meshes present before the light
    size_t n = 0;
    for (auto& m : meshes) { for (auto& v : m.vertices) sphere.center += v; n += m.vertices.size(); }
    sphere.center /= float(n);
    for (auto& m : meshes) for (auto& v : m.vertices) sphere.radius = glm::max(sphere.radius, glm::distance2(sphere.center, v));
    sphere.radius = glm::sqrt(sphere.radius);
  }
  size_t i = lights.num_lights();
  lights.addLight("extra_light", materials.names.size(), vec3(a[0], a[1], a[2]), normalize(vec3(a[3], a[4], a[5])),
                  normalize(vec3(a[6], a[7], a[8])), vec3(a[9], a[10], a[11]), vec2(a[12], a[13]), true);
  meshes.push_back(lights.light(i).create_mesh(materials.names.size(), "extra_light"));
  materials.names.push_back("extra_light");
  materials.bsdfs.push_back(lights.light(i).create_bsdf(sphere, i, true));
  return make_shared<Scene>(std::move(cameras), std::move(materials), std::move(meshes), std::move(lights));
}

// ---------------------------------------------------------------- helpers
typedef std::map<string, string> Args;
static Args parse_kv(int argc, char** argv, int first) {
  Args a;
  for (int i = first; i < argc; ++i) {
    string s = argv[i];
    size_t e = s.find('=');
    if (e == string::npos) throw std::runtime_error("expected key=value, got " + s);
    a[s.substr(0, e)] = s.substr(e + 1);
  }
  return a;
}
static string arg(const Args& a, const string& k, const string& d) { auto i = a.find(k); return i == a.end() ? d : i->second; }
static double argd(const Args& a, const string& k, double d) { auto i = a.find(k); return i == a.end() ? d : atof(i->second.c_str()); }

static shared<Technique> make(const shared<Scene>& scene, const Args& a, size_t threads) {
  string t = arg(a, "technique", "PT");
  float lights = argd(a, "lights", 1.0), roulette = argd(a, "roulette", 0.9), beta = argd(a, "beta", t == "PT" ? 1.0 : 2.0);
  size_t W = argd(a, "width", 512), H = argd(a, "height", 512);
  if (t == "PT") {
    size_t max_path = argd(a, "max_path", 0);
    return make_shared<PathTracing>(scene, lights, roulette, beta, max_path ? max_path : size_t(PTRDIFF_MAX), threads);
  }
  if (t == "BPT") {
    if (beta == 0.0f) return make_shared<BPT0>(scene, lights, roulette, beta, threads);
    if (beta == 1.0f) return make_shared<BPT1>(scene, lights, roulette, beta, threads);
    if (beta == 2.0f) return make_shared<BPT2>(scene, lights, roulette, beta, threads);
    return make_shared<BPTb>(scene, lights, roulette, beta, threads);
  }
  if (t == "VCM" || t == "UPG") {
    bool unbiased = t == "UPG";
    bool vc = argd(a, "enable_vc", 1) != 0, vm = argd(a, "enable_vm", 1) != 0, from_light = argd(a, "from_light", 1) != 0;
    size_t photons = argd(a, "num_photons", 0);
    if (!photons) photons = W * H;  // Options.cpp:914-916
    float radius = argd(a, "radius", 0.01), alpha = argd(a, "alpha", 0.75);
    if (beta == 0.0f) return make_shared<UPG0>(scene, unbiased, vc, vm, from_light, lights, roulette, photons, radius, alpha, beta, threads);
    if (beta == 1.0f) return make_shared<UPG1>(scene, unbiased, vc, vm, from_light, lights, roulette, photons, radius, alpha, beta, threads);
    if (beta == 2.0f) return make_shared<UPG2>(scene, unbiased, vc, vm, from_light, lights, roulette, photons, radius, alpha, beta, threads);
    return make_shared<UPGb>(scene, unbiased, vc, vm, from_light, lights, roulette, photons, radius, alpha, beta, threads);
  }
  throw std::runtime_error("unknown technique " + t);
}

// sky=hr,hg,hb,zr,zg,zb  ->  Technique::set_sky_gradient (Technique.cpp:91-93)
static void apply_sky(Technique& T, const Args& a) {
  string v = arg(a, "sky", "");
  if (v.empty()) return;
  float f[6] = {0, 0, 0, 0, 0, 0};
  if (sscanf(v.c_str(), "%f,%f,%f,%f,%f,%f", f, f + 1, f + 2, f + 3, f + 4, f + 5) != 6) throw std::runtime_error("sky= needs six numbers");
  T.set_sky_gradient(vec3(f[0], f[1], f[2]), vec3(f[3], f[4], f[5]));
}

static render_context_t make_context(const Scene& scene, size_t cameraId, size_t W, size_t H, random_generator_t& gen) {
  // Technique::render, Technique.cpp:37-45
  auto& cameras = scene.cameras();
  render_context_t context;
  context.view_to_world_mat3 = cameras.view_to_world_mat3(cameraId);
  context.world_to_view_mat3 = cameras.world_to_view_mat3(cameraId);
  context.camera_position = cameras.position(cameraId);
  context.resolution = vec2(W, H);
  context.resolution_y_inv = 1.0f / context.resolution.y;
  context.focal_length_y = cameras.focal_length_y(cameraId, context.resolution.x / context.resolution.y);
  context.focal_factor_y = context.focal_length_y * context.focal_length_y * 0.25f;
  context.generator = &gen;
  context.camera_id = int32_t(cameraId);
  return context;
}

template <class T> static void write_vec(const string& path, const std::vector<T>& v) {
  std::ofstream f(path, std::ios::binary);
  f.write((const char*)v.data(), v.size() * sizeof(T));
}

#ifndef B2PT_STOCK_TIMING  // everything that needs the counter-based engine of rng_shim.h
// UPGBase::_preprocess + _scatter (UPG.cpp:177-186, 599-672) with ONE stream per photon path instead of one
// sequential generator: photon j of pass `sample` draws from stream (seed, domain 2, j, sample).  Everything the
// photon does — _traceLight, the vertex records, the v3::HashGrid3D build — is the reference's own code; only the
// seeding between two _traceLight calls is the harness's.  Returns false if T is not this UPGBase instantiation.
template <class U>
static bool upg_photon_pass(Technique& T, random_generator_t& gen, uint32_t seed, uint32_t sample, const string& dump) {
  U* u = dynamic_cast<U*>(&T);
  if (!u) return false;
  if (!u->_unbiased) {  // UPG.cpp:180-183 with num_samples = the pass index
    u->_radius = u->_initial_radius * pow((double(sample) + 1.0f), u->_alpha * 0.5f - 0.5f);
    u->_circle = pi<float>() * u->_radius * u->_radius;
  }
  size_t size = 0;
  u->_light_offsets.assign(1, 0);
  for (size_t j = 0; j < u->_num_photons; ++j) {
    gen.seed(std::b2pt_counter_engine::stream_key(seed, 2, uint32_t(j), sample));
    u->_traceLight(gen, u->_light_paths, size);
    u->_light_offsets.push_back(uint32_t(size));
  }
  u->_light_paths.resize(size);
  u->_num_scattered = u->_num_photons;
  u->_num_scattered_inv = 1.0f / float(u->_num_scattered);
  u->_statistics.num_scattered += u->_num_scattered;
  u->_vertices = v3::HashGrid3D<typename U::LightVertex>(&u->_light_paths, u->_radius);
  if (!dump.empty()) {
    std::vector<float> v;
    for (auto& x : u->_light_paths) {
      v.push_back(x.surface.position().x); v.push_back(x.surface.position().y); v.push_back(x.surface.position().z);
      v.push_back(x.throughput.x); v.push_back(x.throughput.y); v.push_back(x.throughput.z);
      v.push_back(x.a); v.push_back(x.A); v.push_back(x.B); v.push_back(x.bGeometry);
    }
    write_vec(dump + ".photons.f32", v);
    write_vec(dump + ".offsets.u32", u->_light_offsets);
    std::vector<float> r(1, u->_radius);
    write_vec(dump + ".radius.f32", r);
  }
  return true;
}

static bool photon_pass(Technique& T, random_generator_t& gen, uint32_t seed, uint32_t sample, const string& dump) {
  return upg_photon_pass<UPG0>(T, gen, seed, sample, dump) || upg_photon_pass<UPG1>(T, gen, seed, sample, dump) ||
         upg_photon_pass<UPG2>(T, gen, seed, sample, dump) || upg_photon_pass<UPGb>(T, gen, seed, sample, dump);
}

#endif  // B2PT_STOCK_TIMING

// ---------------------------------------------------------------- sub-commands
static int cmd_dump(int argc, char** argv) {
  if (argc < 4) return 2;
  shared<Scene> scene;
  if (argc >= 5 && string(argv[4]) == "--light") {
    if (argc < 19) throw std::runtime_error("--light needs 14 numbers");
    float a[14];
    for (int i = 0; i < 14; ++i) a[i] = atof(argv[5 + i]);
    scene = with_extra_light(argv[2], a);
  } else {
    scene = ends_with(argv[2], ".b2s") ? read_b2s(argv[2]) : loadScene(argv[2]);
  }
  write_b2s(*scene, argv[3]);
  size_t tris = 0;
  for (auto& m : scene->meshes) tris += m.indices.size() / 3;
  printf("{\"meshes\": %zu, \"triangles\": %zu, \"materials\": %zu, \"lights\": %zu, \"cameras\": %zu}\n",
         scene->meshes.size(), tris, scene->materials.bsdfs.size(), scene->lights.num_lights(), scene->cameras().numCameras());
  return 0;
}

static int cmd_rays(int argc, char** argv) {
  if (argc < 6) return 2;
  shared<Scene> scene = open_scene(argv[2]);
  int mode = atoi(argv[3]);
  std::ifstream f(argv[4], std::ios::binary | std::ios::ate);
  size_t n = size_t(f.tellg()) / sizeof(B2ptRay);
  f.seekg(0);
  std::vector<B2ptRay> rays(n);
  f.read((char*)rays.data(), n * sizeof(B2ptRay));
  std::vector<B2ptHit> hits(n);
  auto t0 = std::chrono::steady_clock::now();
  for (size_t i = 0; i < n; ++i) {
    // the RTCRay set-up of Scene.cpp:162-171 / 186-196 / 209-219
    RTCRay r;
    std::memcpy(r.org, rays[i].org, 12);
    std::memcpy(r.dir, rays[i].dir, 12);
    r.tnear = rays[i].tnear; r.tfar = rays[i].tfar;
    r.geomID = r.primID = r.instID = RTC_INVALID_GEOMETRY_ID;
    r.mask = mode == B2PT_TRACE_CLOSEST_ALL ? 0xFFFFFFFFu : 1u << uint32_t(entity_type::mesh);
    r.time = 0.f;
    if (mode == B2PT_TRACE_OCCLUDED_MESH) rtcOccluded(scene->rtcScene, r); else rtcIntersect(scene->rtcScene, r);
    hits[i].t = r.tfar; hits[i].u = r.u; hits[i].v = r.v; hits[i].geom = r.geomID; hits[i].prim = r.primID;
  }
  double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  write_vec(argv[5], hits);
  printf("{\"rays\": %zu, \"seconds\": %.6f, \"mrays_per_s\": %.4f}\n", n, s, n / s * 1e-6);
  return 0;
}

#ifndef B2PT_STOCK_TIMING
// Per-(pixel, sample) streams.  The harness owns only the pixel loop and the seeding; the jitter
// draws, _traceEye, the splats and the commit are the reference's (Technique.cpp:194-352).
static int cmd_render(int argc, char** argv) {
  if (argc < 3) return 2;
  Args a = parse_kv(argc, argv, 3);
  shared<Scene> scene = open_scene(argv[2]);
  size_t W = argd(a, "width", 512), H = argd(a, "height", 512), cam = argd(a, "camera", 0);
  uint32_t seed = argd(a, "seed", 1234), first = argd(a, "first_sample", 0), stride = argd(a, "sample_stride", 1);
  size_t spp = argd(a, "num_samples", 1);
  bool per_path = argd(a, "per_path", 1) != 0;
  string out = arg(a, "out", "/tmp/ref_render");
  shared<Technique> tech = make(scene, a, 1);
  Technique& T = *tech;
  apply_sky(T, a);

  std::vector<dvec4> accum(W * H, dvec4(0.0));
  subimage_view_t view(accum.data(), W, H);
  random_generator_t gen(size_t(0));
  render_context_t context = make_context(*scene, cam, W, H, gen);
  std::vector<float> path_radiance;  // [spp][H][W][3], eye image only
  std::vector<double> light_image(W * H * 3, 0.0);  // sum over passes of the light-tracing splats
  size_t rays0 = scene->numNormalRays(), shadow0 = scene->numShadowRays();
  auto t0 = std::chrono::steady_clock::now();
  for (size_t s = 0; s < spp; ++s) {
    uint32_t sample = first + uint32_t(s) * stride;
    T._adjust_helper_image(view);
    // UPG/VCM: the photon pass with one stream per photon (domain 2); PT/BPT: _preprocess is a no-op
    if (!photon_pass(T, gen, seed, sample, s + 1 == spp ? arg(a, "dump_photons", "") : string())) {
      gen.seed(std::b2pt_counter_engine::stream_key(seed, 1, 0, sample));
      T._preprocess(gen, double(sample));
    }
    for (size_t y = 0; y < H; ++y) {
      for (size_t x = 0; x < W; ++x) {
        gen.seed(std::b2pt_counter_engine::stream_key(seed, 0, uint32_t(y * W + x), sample));
        subimage_view_t px = view;
        px._xOffset = x; px._xWindow = 1; px._yOffset = y; px._yWindow = 1;
        T._for_each_ray(px, context);
        if (per_path) {
          const dvec3& e = T._eye_image[y * W + x];
          path_radiance.push_back(float(e.x)); path_radiance.push_back(float(e.y)); path_radiance.push_back(float(e.z));
        }
      }
    }
    for (size_t i = 0; i < W * H; ++i) {
      light_image[i * 3 + 0] += T._light_image[i].x; light_image[i * 3 + 1] += T._light_image[i].y; light_image[i * 3 + 2] += T._light_image[i].z;
    }
    T._commit_images(view);
    ++T._statistics.num_samples;
  }
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  write_vec(out + ".accum.f64", accum);
  write_vec(out + ".light.f64", light_image);
  if (per_path) write_vec(out + ".paths.f32", path_radiance);
  size_t nb = scene->numNormalRays() - rays0, ns = scene->numShadowRays() - shadow0;
  printf("{\"width\": %zu, \"height\": %zu, \"num_samples\": %zu, \"seconds\": %.6f, \"basic_rays\": %zu, \"shadow_rays\": %zu, "
         "\"focal_length_y\": %.9g, \"msamples_per_s\": %.6f, \"mrays_per_s\": %.6f}\n",
         W, H, spp, sec, nb, ns, context.focal_length_y, W * H * spp / sec * 1e-6, (nb + ns) / sec * 1e-6);
  return 0;
}

#endif  // B2PT_STOCK_TIMING

// The reference's own pass loop (Application::render -> Technique::render), as it ships:
// num_threads pool threads, 32x32 tiles, per-tile std::random_device seeding when threads > 1.
static int cmd_time(int argc, char** argv) {
  if (argc < 3) return 2;
  Args a = parse_kv(argc, argv, 3);
  shared<Scene> scene = open_scene(argv[2]);
  size_t W = argd(a, "width", 512), H = argd(a, "height", 512), cam = argd(a, "camera", 0);
  size_t threads = argd(a, "threads", 0);
  if (!threads) threads = std::thread::hardware_concurrency();
  size_t spp = argd(a, "num_samples", 1), warm = argd(a, "warmup", 0);
  double budget = argd(a, "seconds", 0.0);  // stop after this much time (0 = run all num_samples)
  shared<Technique> tech = make(scene, a, threads);
  apply_sky(*tech, a);
  std::vector<dvec4> accum(W * H, dvec4(0.0));
  subimage_view_t view(accum.data(), W, H);
  random_generator_t gen(size_t(argd(a, "seed", 1234)));
  for (size_t s = 0; s < warm; ++s) tech->render(view, gen, cam, {}, {});
  size_t rays0 = scene->numNormalRays(), shadow0 = scene->numShadowRays();
  auto t0 = std::chrono::steady_clock::now();
  size_t done = 0;
  double sec = 0;
  for (size_t s = 0; s < spp; ++s) {
    tech->render(view, gen, cam, {}, {});
    ++done;
    sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (budget > 0 && sec >= budget) break;
  }
  size_t nb = scene->numNormalRays() - rays0, ns = scene->numShadowRays() - shadow0;
  string out = arg(a, "out", "");
  if (!out.empty()) write_vec(out + ".accum.f64", accum);
  printf("{\"width\": %zu, \"height\": %zu, \"num_samples\": %zu, \"threads\": %zu, \"seconds\": %.6f, \"basic_rays\": %zu, "
         "\"shadow_rays\": %zu, \"msamples_per_s\": %.6f, \"mrays_per_s\": %.6f}\n",
         W, H, done, threads, sec, nb, ns, W * H * done / sec * 1e-6, (nb + ns) / sec * 1e-6);
  return 0;
}

// v3::HashGrid3D<T>::build / rQuery (HashGrid3D.hpp:181-325) on explicit points and queries.
//   in:  u32 n, f32 points[3n], u32 nq, f32 queries[3nq], f32 radius
//   out: per query u32 count, u32 indices[count]   (the callback order of rQuery)
struct GridPoint {
  vec3 p;
  const vec3& position() const { return p; }
  bool is_light() const { return false; }
};

static int cmd_grid(int argc, char** argv) {
  if (argc < 4) return 2;
  std::ifstream f(argv[2], std::ios::binary);
  uint32_t n = 0, nq = 0; float radius = 0;
  f.read((char*)&n, 4);
  std::vector<GridPoint> pts(n);
  f.read((char*)pts.data(), size_t(n) * 12);
  f.read((char*)&nq, 4);
  std::vector<vec3> q(nq);
  f.read((char*)q.data(), size_t(nq) * 12);
  f.read((char*)&radius, 4);
  auto t0 = std::chrono::steady_clock::now();
  v3::HashGrid3D<GridPoint> grid(&pts, radius);
  double build = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::vector<uint32_t> out;
  size_t total = 0;
  t0 = std::chrono::steady_clock::now();
  for (uint32_t i = 0; i < nq; ++i) {
    size_t at = out.size();
    out.push_back(0);
    grid.rQuery([&](uint32_t index) { out.push_back(index); }, q[i], radius);
    out[at] = uint32_t(out.size() - at - 1);
    total += out[at];
  }
  double query = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  write_vec(argv[3], out);
  printf("{\"points\": %u, \"queries\": %u, \"radius\": %.9g, \"found\": %zu, \"build_s\": %.6f, \"query_s\": %.6f}\n", n, nq, radius, total,
         build, query);
  return 0;
}


#ifndef B2PT_STOCK_TIMING
// BSDF::gathering_density (BSDF.cpp:42-66, 264-289, 393-434) + Intersector::intersectFast (Intersector.cpp:17-46) on explicit jobs.
//   in:  u32 n, then n x { f32 ray_org[3], ray_dir[3], target[3], radius; u32 key }   (44 bytes)
//        the surface of a job is the closest MESH hit of its ray (Scene::intersectMesh's RTCRay), omega = -ray_dir
//   out: n x { u32 trials, f32 density, u32 bsdf_kind (0 none / not a lobe, 1 diffuse, 2 phong), f32 t }
// The engine runs in trial mode (rng_shim.h) so that trial t of a loop sees the numbers the device gives trial t;
// trials = engine draws / draws-per-trial, i.e. the N at which the reference's own loop returned.
struct DensityJobIn { float org[3], dir[3], target[3], radius; uint32_t key; };
struct DensityJobOut { uint32_t trials; float density; uint32_t kind; float t; };

static int cmd_density(int argc, char** argv) {
  if (argc < 5) return 2;
  shared<Scene> scene = open_scene(argv[2]);
  std::ifstream f(argv[3], std::ios::binary);
  uint32_t n = 0;
  f.read((char*)&n, 4);
  std::vector<DensityJobIn> jobs(n);
  f.read((char*)jobs.data(), size_t(n) * sizeof(DensityJobIn));
  std::vector<DensityJobOut> out(n);
  random_generator_t gen(size_t(0));
  size_t rays0 = scene->numNormalRays();
  auto t0 = std::chrono::steady_clock::now();
  for (uint32_t i = 0; i < n; ++i) {
    const DensityJobIn& j = jobs[i];
    RayIsect r;
    std::memcpy(r.org, j.org, 12);
    std::memcpy(r.dir, j.dir, 12);
    r.tnear = 0.f; r.tfar = INFINITY;
    r.geomID = r.primID = r.instID = RTC_INVALID_GEOMETRY_ID;
    r.mask = 1u << uint32_t(entity_type::mesh);
    r.time = 0.f;
    rtcIntersect(scene->rtcScene, r);
    out[i] = DensityJobOut{0u, INFINITY, 0u, r.tfar};
    if (!r.isPresent()) continue;
    SurfacePoint surface = scene->querySurface(r);
    const BSDF& bsdf = scene->queryBSDF(surface);
    uint32_t draws = 0;
    if (dynamic_cast<const DiffuseBSDF*>(&bsdf)) { draws = 2; out[i].kind = 1; }
    else if (dynamic_cast<const PhongBSDF*>(&bsdf)) { draws = 3; out[i].kind = 2; }
    else continue;  // delta lobes never reach the estimate; BSDF::sample_bounded would assert
    gen.seed(j.key);
    gen.engine.trial_draws = draws;
    bounding_sphere_t target = {vec3(j.target[0], j.target[1], j.target[2]), j.radius};
    vec3 omega = -vec3(j.dir[0], j.dir[1], j.dir[2]);
    out[i].density = bsdf.gathering_density(gen, scene.get(), surface, target, omega);
    out[i].trials = gen.engine.ctr / draws;
    gen.engine.trial_draws = 0;
  }
  double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  write_vec(argv[4], out);
  printf("{\"jobs\": %u, \"seconds\": %.6f, \"rays\": %zu}\n", n, s, scene->numNormalRays() - rays0);
  return 0;
}

#endif  // B2PT_STOCK_TIMING

// rms_abs_errors / rms_abs_errors_windowed (ImageView.cpp:60-117) — the reference's own functions — on a dvec4 film and a vec3 reference.
static int cmd_errors(int argc, char** argv) {
  if (argc < 6) return 2;
  size_t W = atoi(argv[4]), H = atoi(argv[5]);
  std::vector<dvec4> a(W * H);
  std::vector<vec3> b(W * H);
  { std::ifstream f(argv[2], std::ios::binary); f.read((char*)a.data(), W * H * sizeof(dvec4)); if (!f) throw std::runtime_error("short accum file"); }
  { std::ifstream f(argv[3], std::ios::binary); f.read((char*)b.data(), W * H * sizeof(vec3)); if (!f) throw std::runtime_error("short reference file"); }
  float rms = 0, abs = 0;
  if (argc >= 9) rms_abs_errors_windowed(rms, abs, image_view_t<dvec4>(a, W, H), image_view_t<vec3>(b, W, H), ivec2(atoi(argv[6]), atoi(argv[7])), atoi(argv[8]));
  else rms_abs_errors(rms, abs, image_view_t<dvec4>(a, W, H), image_view_t<vec3>(b, W, H));
  printf("{\"rms\": %.9g, \"abs\": %.9g}\n", rms, abs);
  return 0;
}

#ifndef B2PT_STOCK_TIMING
// A technique that only records the primary ray _for_each_ray hands to _traceEye (Technique.cpp:321-338).
struct ProbeTechnique : Technique {
  ProbeTechnique(const shared<const Scene>& scene) : Technique(scene, 1) {}
  Ray last;
  vec3 _traceEye(render_context_t& context, Ray ray) override { last = ray; return vec3(0.0f); }
};

// What the parity contract rests on but the C++ standard leaves open (SURVEY.md §7 "hard parts"): the order in which the two
// draws of `vec2(x + g.sample(), y + g.sample())` (Technique.cpp:322) and `vec2(g.sample(), g.sample())` (AreaLights.cpp:224)
// are evaluated in THIS build of the reference, plus libstdc++'s float / piecewise draws from the engine.
static int cmd_pinning(int argc, char** argv) {
  if (argc < 3) return 2;
  Args a = parse_kv(argc, argv, 3);
  shared<Scene> scene = open_scene(argv[2]);
  size_t W = argd(a, "width", 64), H = argd(a, "height", 48), cam = argd(a, "camera", 0);
  uint32_t seed = argd(a, "seed", 1234), count = argd(a, "count", 16);
  random_generator_t gen(size_t(0));
  printf("{\"width\": %zu, \"height\": %zu, \"seed\": %u", W, H, seed);
  uint32_t key = std::b2pt_counter_engine::stream_key(seed, 0, 7, 3);
  printf(", \"key\": %u, \"raw\": [", key);
  gen.seed(key);
  for (int i = 0; i < 8; ++i) printf("%s%u", i ? ", " : "", (unsigned)gen());
  printf("], \"floats\": [");
  gen.seed(key);
  for (int i = 0; i < 8; ++i) printf("%s%.9g", i ? ", " : "", gen.sample());
  // piecewise_sampler_t over three weights (Sample.inl:233-247): one double per pick = two engine calls
  const float weights[3] = {0.2f, 0.5f, 0.3f};
  piecewise_sampler_t sampler(weights, weights + 3);
  printf("], \"piecewise_weights\": [0.2, 0.5, 0.3], \"piecewise\": [");
  gen.seed(key);
  for (int i = 0; i < 8; ++i) printf("%s%.9g", i ? ", " : "", sampler.sample(gen));
  printf("], \"piecewise_draws_used\": %u", gen.engine.ctr);
  // primary rays: pixel p = (x, y), sample s, through the reference's own _for_each_ray
  ProbeTechnique probe(scene);
  std::vector<dvec4> accum(W * H, dvec4(0.0));
  subimage_view_t view(accum.data(), W, H);
  render_context_t context = make_context(*scene, cam, W, H, gen);
  probe._adjust_helper_image(view);
  printf(", \"focal_length_y\": %.9g, \"primary\": [", context.focal_length_y);
  for (uint32_t i = 0; i < count; ++i) {
    uint32_t pixel = (i * 2654435761u) % uint32_t(W * H), sample = i % 5;
    gen.seed(std::b2pt_counter_engine::stream_key(seed, 0, pixel, sample));
    subimage_view_t px = view;
    px._xOffset = pixel % W; px._xWindow = 1; px._yOffset = pixel / W; px._yWindow = 1;
    probe._for_each_ray(px, context);
    printf("%s{\"pixel\": %u, \"sample\": %u, \"origin\": [%.9g, %.9g, %.9g], \"direction\": [%.9g, %.9g, %.9g]}", i ? ", " : "", pixel, sample,
           probe.last.origin.x, probe.last.origin.y, probe.last.origin.z, probe.last.direction.x, probe.last.direction.y, probe.last.direction.z);
  }
  // light samples: AreaLights::sample (AreaLights.cpp:121-140) with a known stream
  printf("], \"lights\": [");
  for (uint32_t i = 0; i < count; ++i) {
    uint32_t k = std::b2pt_counter_engine::stream_key(seed, 3, i, 0);
    gen.seed(k);
    LightSample ls = scene->lights.sample(gen);
    printf("%s{\"key\": %u, \"position\": [%.9g, %.9g, %.9g], \"material_id\": %u, \"draws\": %u}", i ? ", " : "", k, ls.surface.position().x,
           ls.surface.position().y, ls.surface.position().z, (unsigned)ls.surface.material_id, gen.engine.ctr);
  }
  printf("]}\n");
  return 0;
}

#endif  // B2PT_STOCK_TIMING

// The known answers the reference keeps in its own (never compiled) unittest() blocks and stale gtest file:
// Cameras.cpp:164-189, Technique.cpp:118-152, unit_tests/Cameras.test.cpp:38-47 — evaluated by the reference's functions.
static int cmd_camera(int, char**) {
  auto v3 = [](const char* name, vec3 v) { printf("\"%s\": [%.9g, %.9g, %.9g], ", name, v.x, v.y, v.z); };
  printf("{");
  {  // Cameras.cpp:164-173: ray_direction <-> pixel_position round trip
    vec2 resolution(800.0f, 600.0f);
    float fov_y = half_pi<float>();
    vec3 d = ray_direction(vec2(123.4f, 345.0f), resolution, fov_y);
    vec2 p = pixel_position(d, resolution, fov_y);
    v3("roundtrip_direction", d);
    printf("\"roundtrip_pixel\": [%.9g, %.9g], \"flength_90\": %.9g, ", p.x, p.y, normalized_flength_y(fov_y));
    // unit_tests/Cameras.test.cpp:38-47: directions at the image centre and corners, 800 x 600, fov 90 degrees
    v3("dir_400_300", ray_direction(vec2(400.0f, 300.0f), resolution, fov_y));
    v3("dir_0_0", ray_direction(vec2(0.0f, 0.0f), resolution, fov_y));
    v3("dir_800_600", ray_direction(vec2(800.0f, 600.0f), resolution, fov_y));
  }
  {  // Cameras.cpp:175-189
    mat4 view = glm::lookAt(vec3(1.f, 3.f, 2.f), vec3(1.f, 3.f, 2.f) + vec3(0.0f, 0.0f, -1.0f), vec3(0.0f, 1.0f, 0.0f));
    v3("lookat_position", camera_position(view));
    mat3 t = camera_tangent_space(view);
    v3("lookat_tangent0", t[0]); v3("lookat_tangent1", t[1]); v3("lookat_tangent2", t[2]);
  }
  {  // Technique.cpp:118-152
    vec3 position(1.f, 3.f, 2.f), direction(1.0f, 0.0f, 10.0f), up(0.0f, 1.0f, 0.0f);
    mat3 view_to_world = inverse(glm::lookAt(position, position + direction, up));
    v3("frame_tangent0", view_to_world[1]); v3("frame_tangent1", view_to_world[2]); v3("frame_tangent2", view_to_world[0]);
    v3("frame_expected1", -normalize(direction)); v3("frame_expected2", normalize(cross(direction, up)));
    mat3 v2 = inverse(glm::lookAt(vec3(1.f), vec3(1.f) + vec3(1.0f, 2.0f, 3.0f), up));
    v3("frame2_col2", v2[2]); v3("frame2_expected", -normalize(vec3(1.0f, 2.0f, 3.0f)));
  }
  printf("\"ok\": true}\n");
  return 0;
}

// The importer's output as loadScene sees it (loader.cpp:458-487: same flags), written before any of the loader's own arithmetic:
// cameras, material properties, per-mesh positions / normals / optional tangents / faces, area lights.  Format: raw_import.hpp.
static int cmd_rawdump(int argc, char** argv) {
  if (argc < 4) return 2;
  Assimp::Importer importer;
  auto flags = aiProcess_Triangulate | aiProcess_GenNormals | aiProcess_JoinIdenticalVertices | aiProcess_PreTransformVertices;
  const aiScene* scene = importer.ReadFile(argv[2], flags);
  if (!scene) throw std::runtime_error(string("Cannot load \"") + argv[2] + "\" scene.");
  std::ofstream f(argv[3], std::ios::binary);
  auto u32 = [&](uint32_t v) { f.write((const char*)&v, 4); };
  auto f32 = [&](float v) { f.write((const char*)&v, 4); };
  auto str = [&](const char* s) { u32(uint32_t(strlen(s))); f.write(s, strlen(s)); };
  auto v3 = [&](const aiVector3D& v) { f32(v.x); f32(v.y); f32(v.z); };
  auto c3 = [&](const aiColor3D& v) { f32(v.r); f32(v.g); f32(v.b); };
  f.write("B2IMPORT", 8);
  uint32_t area_lights = 0;
  for (unsigned i = 0; i < scene->mNumLights; ++i) area_lights += scene->mLights[i]->mType == aiLightSource_AREA;
  u32(scene->mNumCameras); u32(scene->mNumMaterials); u32(scene->mNumMeshes); u32(scene->mNumLights);
  for (unsigned i = 0; i < scene->mNumCameras; ++i) {
    auto c = scene->mCameras[i];
    str(c->mName.C_Str()); v3(c->mPosition); v3(c->mLookAt); v3(c->mUp); f32(c->mHorizontalFOV); f32(c->mClipPlaneNear); f32(c->mClipPlaneFar);
  }
  for (unsigned i = 0; i < scene->mNumMaterials; ++i) {
    const aiMaterial* m = scene->mMaterials[i];
    aiString name; m->Get(AI_MATKEY_NAME, name);
    int transparency = 0, mirror = 0; float ior = 0, shin = 0;
    m->Get("$mat.blend.transparency.use", 0, 0, transparency); m->Get("$mat.blend.transparency.ior", 0, 0, ior);
    m->Get("$mat.blend.mirror.use", 0, 0, mirror); m->Get(AI_MATKEY_SHININESS, shin);
    aiColor3D diffuse, specular;
    m->Get(AI_MATKEY_COLOR_DIFFUSE, diffuse); m->Get(AI_MATKEY_COLOR_SPECULAR, specular);
    str(name.C_Str()); u32(transparency != 0); f32(ior); u32(mirror != 0); c3(diffuse); c3(specular); f32(shin);
  }
  for (unsigned i = 0; i < scene->mNumMeshes; ++i) {
    const aiMesh* m = scene->mMeshes[i];
    const bool has_tangents = m->mBitangents != nullptr && m->mTangents != nullptr;
    str(m->mName.C_Str()); u32(m->mMaterialIndex); u32(m->mNumVertices); u32(m->mNumFaces); u32(has_tangents);
    for (unsigned j = 0; j < m->mNumVertices; ++j) v3(m->mVertices[j]);
    for (unsigned j = 0; j < m->mNumVertices; ++j) v3(m->mNormals[j]);
    if (has_tangents) {
      for (unsigned j = 0; j < m->mNumVertices; ++j) v3(m->mTangents[j]);
      for (unsigned j = 0; j < m->mNumVertices; ++j) v3(m->mBitangents[j]);
    }
    for (unsigned j = 0; j < m->mNumFaces; ++j) {
      if (m->mFaces[j].mNumIndices != 3) throw std::runtime_error("non-triangle face after aiProcess_Triangulate");
      for (int k = 0; k < 3; ++k) u32(m->mFaces[j].mIndices[k]);
    }
  }
  for (unsigned i = 0; i < scene->mNumLights; ++i) {
    const aiLight* l = scene->mLights[i];
    str(l->mName.C_Str()); u32(l->mType == aiLightSource_AREA); v3(l->mPosition); v3(l->mDirection); v3(l->mUp); c3(l->mColorDiffuse);
    f32(l->mSize.x); f32(l->mSize.y); u32(l->mDiffuse ? 1u : 0u);
  }
  printf("{\"cameras\": %u, \"materials\": %u, \"meshes\": %u, \"lights\": %u, \"area_lights\": %u}\n", scene->mNumCameras, scene->mNumMaterials,
         scene->mNumMeshes, scene->mNumLights, area_lights);
  return 0;
}

int main(int argc, char** argv) {
  _MM_SET_FLUSH_ZERO_MODE(_MM_FLUSH_ZERO_ON);          // main.cpp:70-71
  _MM_SET_DENORMALS_ZERO_MODE(_MM_DENORMALS_ZERO_ON);
  try {
    string c = argc > 1 ? argv[1] : "";
    if (c == "dump") return cmd_dump(argc, argv);
    if (c == "rays") return cmd_rays(argc, argv);
    if (c == "time") return cmd_time(argc, argv);
    if (c == "grid") return cmd_grid(argc, argv);
    if (c == "errors") return cmd_errors(argc, argv);
    if (c == "camera") return cmd_camera(argc, argv);
    if (c == "rawdump") return cmd_rawdump(argc, argv);
#ifndef B2PT_STOCK_TIMING
    if (c == "render") return cmd_render(argc, argv);
    if (c == "density") return cmd_density(argc, argv);
    if (c == "pinning") return cmd_pinning(argc, argv);
#endif
    fprintf(stderr, "usage: ref_harness dump|rays|render|time|grid|density|errors|pinning|camera ...\n");
    return 2;
  } catch (const std::exception& e) {
    fprintf(stderr, "ref_harness: %s\n", e.what());
    return 1;
  }
}
