// b2s_io.hpp — haste::Scene <-> flattened scene (.b2s file or B2ptSceneDesc in memory).
//
// Host side of the drop-in, compiled TOGETHER WITH the reference's own headers (Scene.hpp, BSDF.hpp,
// AreaLights.hpp, Cameras.hpp): this is the code a maintainer of ciechowoj/master adds next to
// loader.cpp.  It needs -fno-access-control (or friend declarations) because the BSDF classes keep
// their parameters private (BSDF.hpp:54-174).
//
//   flatten_bsdf      BSDF subclass -> B2ptMaterial (the switch that replaces virtual dispatch on the GPU)
//   write_b2s         Scene -> file         (scene ingestion without Assimp)
//   read_b2s          file -> Scene through the reference's own constructors
//   FlatSceneArrays   Scene -> the arrays B2ptSceneDesc points at (what GpuTechnique uploads)
#pragma once
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <Scene.hpp>

#include "../../include/b2pt.h"

namespace b2pt_host {

using namespace haste;
using std::string;

// magic "B2SCENE1"; u32 counts[8] = {cameras, meshes, vertices, indices, materials, lights, 0, 0};
// cameras * CameraRec; u32 mesh_material_id[M]; u32 vertex_offset[M+1]; u32 index_offset[M+1];
// f32 vertices[3V]; f32 tangents[9V]; i32 indices[I]; B2ptMaterial[materials]; B2ptLight[lights].
struct CameraRec {
  float position[3], direction[3], up[3];
  float fovx, near, far, focal;
  float world_to_view[9], view_to_world[9];
};

inline void put(std::ofstream& f, const void* p, size_t n) { f.write((const char*)p, n); }
inline void get(std::ifstream& f, void* p, size_t n) {
  f.read((char*)p, n);
  if (!f) throw std::runtime_error("short read in .b2s");
}

inline B2ptMaterial flatten_bsdf(const BSDF* bsdf) {
  B2ptMaterial m;
  std::memset(&m, 0, sizeof m);
  m.light_id = UINT32_MAX;
  if (auto b = dynamic_cast<const DiffuseBSDF*>(bsdf)) {
    m.type = B2PT_BSDF_DIFFUSE;
    std::memcpy(m.diffuse, &b->_diffuse, 12);
  } else if (auto b = dynamic_cast<const PhongBSDF*>(bsdf)) {
    m.type = B2PT_BSDF_PHONG;
    std::memcpy(m.diffuse, &b->_diffuse, 12);
    std::memcpy(m.specular, &b->_specular, 12);
    m.power = b->_power;
    m.diffuse_probability = b->_diffuse_probability;
  } else if (dynamic_cast<const ReflectionBSDF*>(bsdf)) {
    m.type = B2PT_BSDF_REFLECTION;
  } else if (auto b = dynamic_cast<const TransmissionBSDF*>(bsdf)) {
    m.type = B2PT_BSDF_TRANSMISSION;
    m.ext_over_int_ior = b->externalOverInternalIOR;
    m.internal_ior = b->internalIOR;
  } else if (auto b = dynamic_cast<const LightBSDF*>(bsdf)) {
    m.type = B2PT_BSDF_LIGHT;
    m.light_id = b->_light_id;
    std::memcpy(m.sphere, &b->_sphere, 16);
  } else if (auto b = dynamic_cast<const sun_light_bsdf*>(bsdf)) {
    m.type = B2PT_BSDF_SUN;
    m.light_id = b->_light_id;
    std::memcpy(m.sphere, &b->_sphere, 16);
  } else if (dynamic_cast<const CameraBSDF*>(bsdf)) {
    m.type = B2PT_BSDF_CAMERA;
  } else {
    throw std::runtime_error("unknown BSDF class");
  }
  return m;
}

inline void write_b2s(const Scene& scene, const string& path) {
  std::ofstream f(path, std::ios::binary);
  uint32_t nv = 0, ni = 0;
  for (auto& m : scene.meshes) { nv += m.vertices.size(); ni += m.indices.size(); }
  const Cameras& cams = scene.cameras();
  uint32_t counts[8] = {uint32_t(cams.numCameras()), uint32_t(scene.meshes.size()), nv, ni,
                        uint32_t(scene.materials.bsdfs.size()), uint32_t(scene.lights.num_lights()), 0, 0};
  put(f, "B2SCENE1", 8);
  put(f, counts, sizeof counts);
  for (size_t i = 0; i < cams.numCameras(); ++i) {
    CameraRec c;
    std::memcpy(c.position, &cams._descs[i].position, 12);
    std::memcpy(c.direction, &cams._descs[i].direction, 12);
    std::memcpy(c.up, &cams._descs[i].up, 12);
    c.fovx = cams._descs[i].fovx; c.near = cams._descs[i].near; c.far = cams._descs[i].far;
    c.focal = cams._focals[i];
    mat3 w2v = cams.world_to_view_mat3(i), v2w = cams.view_to_world_mat3(i);
    std::memcpy(c.world_to_view, &w2v, 36);
    std::memcpy(c.view_to_world, &v2w, 36);
    put(f, &c, sizeof c);
  }
  std::vector<uint32_t> mat, voff(1, 0), ioff(1, 0);
  for (auto& m : scene.meshes) {
    mat.push_back(m.material_id);
    voff.push_back(voff.back() + m.vertices.size());
    ioff.push_back(ioff.back() + m.indices.size());
  }
  put(f, mat.data(), mat.size() * 4);
  put(f, voff.data(), voff.size() * 4);
  put(f, ioff.data(), ioff.size() * 4);
  for (auto& m : scene.meshes) put(f, m.vertices.data(), m.vertices.size() * 12);
  for (auto& m : scene.meshes) put(f, m.tangents.data(), m.tangents.size() * 36);
  for (auto& m : scene.meshes) put(f, m.indices.data(), m.indices.size() * 4);
  for (auto& b : scene.materials.bsdfs) { B2ptMaterial m = flatten_bsdf(b.get()); put(f, &m, sizeof m); }
  for (size_t i = 0; i < scene.lights.num_lights(); ++i) {
    const AreaLight& l = scene.lights.light(i);
    B2ptLight r;
    std::memcpy(r.position, &l.position, 12);
    std::memcpy(r.tangent, &l.tangent, 36);
    std::memcpy(r.size, &l.size, 8);
    std::memcpy(r.exitance, &l.exitance, 12);
    r.diffuse = l.diffuse; r.material_id = l.material_id; r.weight = scene.lights._weights[i];
    put(f, &r, sizeof r);
  }
}

// Rebuilds a haste::Scene from a .b2s through the reference's own constructors
// (Cameras::addCameraFovX, AreaLights::addLight, the BSDF classes, Scene::Scene).
inline shared<Scene> read_b2s(const string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("cannot open " + path);
  char magic[8]; get(f, magic, 8);
  if (std::memcmp(magic, "B2SCENE1", 8)) throw std::runtime_error("not a .b2s file");
  uint32_t counts[8]; get(f, counts, sizeof counts);
  uint32_t nc = counts[0], nm = counts[1], nv = counts[2], ni = counts[3], nmat = counts[4], nl = counts[5];
  Cameras cameras;
  for (uint32_t i = 0; i < nc; ++i) {
    CameraRec c; get(f, &c, sizeof c);
    cameras.addCameraFovX("camera" + std::to_string(i), *(vec3*)c.position, *(vec3*)c.direction,
                          *(vec3*)c.up, c.fovx, c.near, c.far);
  }
  std::vector<uint32_t> mat(nm), voff(nm + 1), ioff(nm + 1);
  get(f, mat.data(), nm * 4); get(f, voff.data(), (nm + 1) * 4); get(f, ioff.data(), (nm + 1) * 4);
  std::vector<vec3> verts(nv); std::vector<mat3> tans(nv); std::vector<int> idx(ni);
  get(f, verts.data(), size_t(nv) * 12); get(f, tans.data(), size_t(nv) * 36); get(f, idx.data(), size_t(ni) * 4);
  std::vector<B2ptMaterial> mats(nmat); get(f, mats.data(), nmat * sizeof(B2ptMaterial));
  std::vector<B2ptLight> lights(nl); get(f, lights.data(), nl * sizeof(B2ptLight));

  vector<Mesh> meshes(nm);
  for (uint32_t m = 0; m < nm; ++m) {
    meshes[m].name = "mesh" + std::to_string(m);
    meshes[m].material_id = mat[m];
    meshes[m].vertices.assign(verts.begin() + voff[m], verts.begin() + voff[m + 1]);
    meshes[m].tangents.assign(tans.begin() + voff[m], tans.begin() + voff[m + 1]);
    meshes[m].indices.assign(idx.begin() + ioff[m], idx.begin() + ioff[m + 1]);
  }
  AreaLights area;
  for (uint32_t i = 0; i < nl; ++i) {
    const B2ptLight& l = lights[i];
    const mat3& t = *(const mat3*)l.tangent;
    area.addLight("light" + std::to_string(i), l.material_id >> 2, *(const vec3*)l.position, t[1], t[2],
                  *(const vec3*)l.exitance, *(const vec2*)l.size, l.diffuse != 0.0f);
    // addLight recomputes tangent[0] = normalize(cross(up, direction)); keep the stored frame bit-exact
    area._lights[i].tangent = t;
  }
  Materials materials;
  for (uint32_t i = 0; i < nmat; ++i) {
    const B2ptMaterial& m = mats[i];
    materials.names.push_back("material" + std::to_string(i));
    bounding_sphere_t sphere = {vec3(m.sphere[0], m.sphere[1], m.sphere[2]), m.sphere[3]};
    switch (m.type) {
      case B2PT_BSDF_CAMERA: materials.bsdfs.push_back(unique<BSDF>(new CameraBSDF())); break;
      case B2PT_BSDF_DIFFUSE: materials.bsdfs.push_back(unique<BSDF>(new DiffuseBSDF(*(const vec3*)m.diffuse))); break;
      case B2PT_BSDF_PHONG:
        materials.bsdfs.push_back(unique<BSDF>(new PhongBSDF(*(const vec3*)m.diffuse, *(const vec3*)m.specular, m.power)));
        break;
      case B2PT_BSDF_REFLECTION: materials.bsdfs.push_back(unique<BSDF>(new ReflectionBSDF())); break;
      case B2PT_BSDF_TRANSMISSION: {
        auto t = new TransmissionBSDF(m.internal_ior, 1.0f);
        t->externalOverInternalIOR = m.ext_over_int_ior;
        materials.bsdfs.push_back(unique<BSDF>(t));
        break;
      }
      case B2PT_BSDF_LIGHT: materials.bsdfs.push_back(unique<BSDF>(new LightBSDF(sphere, m.light_id))); break;
      case B2PT_BSDF_SUN: materials.bsdfs.push_back(unique<BSDF>(new sun_light_bsdf(sphere, m.light_id))); break;
      default: throw std::runtime_error("bad material type in .b2s");
    }
  }
  return make_shared<Scene>(std::move(cameras), std::move(materials), std::move(meshes), std::move(area));
}


inline B2ptLight flatten_light(const Scene& scene, size_t i) {
  const AreaLight& l = scene.lights.light(i);
  B2ptLight r;
  std::memcpy(r.position, &l.position, 12);
  std::memcpy(r.tangent, &l.tangent, 36);
  std::memcpy(r.size, &l.size, 8);
  std::memcpy(r.exitance, &l.exitance, 12);
  r.diffuse = l.diffuse; r.material_id = l.material_id; r.weight = scene.lights._weights[i];
  return r;
}

// The arrays a B2ptSceneDesc points at, owned.
struct FlatSceneArrays {
  std::vector<uint32_t> mesh_material_id, mesh_vertex_offset, mesh_index_offset;
  std::vector<float> vertices, tangents;
  std::vector<int32_t> indices;
  std::vector<B2ptMaterial> materials;
  std::vector<B2ptLight> lights;

  explicit FlatSceneArrays(const Scene& scene) {
    mesh_vertex_offset.push_back(0); mesh_index_offset.push_back(0);
    for (auto& m : scene.meshes) {
      mesh_material_id.push_back(m.material_id);
      mesh_vertex_offset.push_back(mesh_vertex_offset.back() + uint32_t(m.vertices.size()));
      mesh_index_offset.push_back(mesh_index_offset.back() + uint32_t(m.indices.size()));
      const float* v = (const float*)m.vertices.data();
      vertices.insert(vertices.end(), v, v + m.vertices.size() * 3);
      const float* t = (const float*)m.tangents.data();
      tangents.insert(tangents.end(), t, t + m.tangents.size() * 9);
      indices.insert(indices.end(), m.indices.begin(), m.indices.end());
    }
    for (auto& b : scene.materials.bsdfs) materials.push_back(flatten_bsdf(b.get()));
    for (size_t i = 0; i < scene.lights.num_lights(); ++i) lights.push_back(flatten_light(scene, i));
  }

  B2ptSceneDesc desc() const {
    B2ptSceneDesc d;
    d.num_meshes = uint32_t(mesh_material_id.size());
    d.mesh_material_id = mesh_material_id.data();
    d.mesh_vertex_offset = mesh_vertex_offset.data();
    d.mesh_index_offset = mesh_index_offset.data();
    d.vertices = vertices.data(); d.tangents = tangents.data(); d.indices = indices.data();
    d.num_materials = uint32_t(materials.size()); d.materials = materials.data();
    d.num_lights = uint32_t(lights.size()); d.lights = lights.empty() ? nullptr : lights.data();
    return d;
  }
};

}  // namespace b2pt_host
