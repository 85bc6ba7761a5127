// raw_import.hpp — scene ingestion without Assimp: the loader's OWN arithmetic, restated on the host.
//
// loadScene (loader.cpp:458-487) is two things: the Assimp fork's .blend importer, and ~150 lines of the reference's own code that
// turn the importer's arrays into haste::Scene — per-vertex tangent frames (aiMeshToMesh, loader.cpp:309-369), the material ->
// BSDF mapping (load_meshes, :371-406), the scene bounding sphere the light lobes are bounded by (:408-432), the light quads
// and their BSDFs (load_lights, :434-456).  A `.b2i` file holds what the importer hands over (cameras, material properties,
// positions / normals / optional tangents / faces per mesh, area lights) BEFORE any of that arithmetic — any exporter can write
// one — and scene_from_import() below is the second half, against the reference's own classes (Cameras::addCameraFovX,
// AreaLights::addLight / create_mesh / create_bsdf, the BSDF constructors).  tests/test_import.py checks it byte for byte
// against the .b2s the Assimp-linked oracle dumps (`ref_harness rawdump` writes the .b2i from the same importer run).
//
//   "B2IMPORT"  u32 cameras, materials, meshes, lights      strings: u32 length + bytes
//   camera:   name, f32 position[3], look_at[3], up[3], horizontal_fov, clip_near, clip_far
//   material: name, u32 transparency_use, f32 ior, u32 mirror_use, f32 diffuse[3], specular[3], shininess
//   mesh:     name, u32 material_index, num_vertices, num_faces, has_tangents, f32 positions[3V], normals[3V],
//             (tangents[3V], bitangents[3V] if has_tangents), u32 faces[3F]
//   light:    name, u32 is_area, f32 position[3], direction[3], up[3], color[3], size[2], u32 diffuse
#pragma once
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include <BSDF.hpp>
#include <Scene.hpp>

namespace b2pt_host {

using namespace haste;

struct RawCamera { string name; vec3 position, look_at, up; float hfov, clip_near, clip_far; };
struct RawMaterial { string name; bool transparency_use; float ior; bool mirror_use; vec3 diffuse, specular; float shininess; };
struct RawMesh { string name; uint32_t material_index; bool has_tangents; std::vector<vec3> positions, normals, tangents, bitangents; std::vector<uint32_t> faces; };
struct RawLight { string name; bool is_area; vec3 position, direction, up, color; vec2 size; bool diffuse; };
struct RawScene { std::vector<RawCamera> cameras; std::vector<RawMaterial> materials; std::vector<RawMesh> meshes; std::vector<RawLight> lights; };

inline RawScene read_b2i(const string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("cannot open " + path);
  auto raw = [&](void* p, size_t n) { f.read((char*)p, n); if (!f) throw std::runtime_error("truncated .b2i file"); };
  auto u32 = [&] { uint32_t v; raw(&v, 4); return v; };
  auto f32 = [&] { float v; raw(&v, 4); return v; };
  auto v3 = [&] { vec3 v; raw(&v, 12); return v; };
  auto str = [&] { uint32_t n = u32(); if (n > (1u << 20)) throw std::runtime_error("bad string in .b2i"); string s(n, '\0'); if (n) raw(&s[0], n); return s; };
  char magic[8]; raw(magic, 8);
  if (std::memcmp(magic, "B2IMPORT", 8)) throw std::runtime_error("not a .b2i file");
  RawScene s;
  const uint32_t nc = u32(), nmat = u32(), nmesh = u32(), nl = u32();
  for (uint32_t i = 0; i < nc; ++i) { RawCamera c; c.name = str(); c.position = v3(); c.look_at = v3(); c.up = v3(); c.hfov = f32(); c.clip_near = f32(); c.clip_far = f32(); s.cameras.push_back(c); }
  for (uint32_t i = 0; i < nmat; ++i) {
    RawMaterial m; m.name = str(); m.transparency_use = u32() != 0; m.ior = f32(); m.mirror_use = u32() != 0; m.diffuse = v3(); m.specular = v3(); m.shininess = f32();
    s.materials.push_back(m);
  }
  for (uint32_t i = 0; i < nmesh; ++i) {
    RawMesh m; m.name = str(); m.material_index = u32();
    const uint32_t nv = u32(), nf = u32(); m.has_tangents = u32() != 0;
    m.positions.resize(nv); m.normals.resize(nv); raw(m.positions.data(), size_t(nv) * 12); raw(m.normals.data(), size_t(nv) * 12);
    if (m.has_tangents) { m.tangents.resize(nv); m.bitangents.resize(nv); raw(m.tangents.data(), size_t(nv) * 12); raw(m.bitangents.data(), size_t(nv) * 12); }
    m.faces.resize(size_t(nf) * 3); raw(m.faces.data(), size_t(nf) * 12);
    s.meshes.push_back(std::move(m));
  }
  for (uint32_t i = 0; i < nl; ++i) {
    RawLight l; l.name = str(); l.is_area = u32() != 0; l.position = v3(); l.direction = v3(); l.up = v3(); l.color = v3();
    l.size.x = f32(); l.size.y = f32(); l.diffuse = u32() != 0;
    s.lights.push_back(l);
  }
  return s;
}

// aiMeshToMesh (loader.cpp:309-369)
inline Mesh mesh_from_import(const RawMesh& mesh, uint32_t materials_base, entity_type type) {
  Mesh result;
  const size_t num_faces = mesh.faces.size() / 3;
  if (!mesh.has_tangents) {
    // no tangents from the importer: the mesh is unshared (three vertices per face) and every corner gets a frame built from
    // its normal and the face's first edge (Gram-Schmidt; a normal parallel to that edge gives NaN, as in the reference)
    result.indices.resize(num_faces * 3);
    result.tangents.resize(num_faces * 3);
    result.vertices.resize(num_faces * 3);
    for (size_t j = 0; j < num_faces; ++j) {
      for (size_t k = 0; k < 3; ++k) {
        const unsigned index = mesh.faces[j * 3 + k];
        result.indices[j * 3 + k] = int(j * 3 + k);
        result.tangents[j * 3 + k][1] = mesh.normals[index];
        result.vertices[j * 3 + k] = mesh.positions[index];
      }
      const vec3 edge = result.vertices[j * 3 + 1] - result.vertices[j * 3 + 0];
      for (size_t k = 0; k < 3; ++k) {
        const vec3 normal = result.tangents[j * 3 + k][1];
        const vec3 tangent = normalize(edge - dot(normal, edge) * normal);
        const vec3 bitangent = normalize(cross(normal, tangent));
        result.tangents[j * 3 + k][2] = bitangent;
        result.tangents[j * 3 + k][0] = tangent;
      }
    }
  } else {
    const size_t nv = mesh.positions.size();
    result.tangents.resize(nv);
    result.vertices.resize(nv);
    for (size_t j = 0; j < nv; ++j) {
      result.tangents[j][0] = mesh.tangents[j];
      result.tangents[j][1] = mesh.normals[j];
      result.tangents[j][2] = mesh.bitangents[j];
      result.vertices[j] = mesh.positions[j];
    }
    result.indices.resize(num_faces * 3);
    for (size_t j = 0; j < num_faces * 3; ++j) result.indices[j] = int(mesh.faces[j]);
  }
  result.name = mesh.name;
  result.material_id = encode_material(mesh.material_index + materials_base, type);
  return result;
}

// compute_bounding_sphere (loader.cpp:408-432): centroid of all vertices, radius to the farthest one
inline bounding_sphere_t bounding_sphere_of(const std::vector<Mesh>& meshes) {
  bounding_sphere_t result = {vec3(0.0f), 0.0f};
  size_t num_vertices = 0;
  for (auto&& mesh : meshes) {
    for (auto&& vertex : mesh.vertices) result.center += vertex;
    num_vertices += mesh.vertices.size();
  }
  result.center /= static_cast<float>(num_vertices);
  for (auto&& mesh : meshes)
    for (auto&& vertex : mesh.vertices) result.radius = glm::max(result.radius, glm::distance2(result.center, vertex));
  result.radius = glm::sqrt(result.radius);
  return result;
}

// loadScene (loader.cpp:458-487) from load_cameras on (:293-307), given the importer's output
inline shared<Scene> scene_from_import(const RawScene& raw) {
  Materials materials;
  Cameras cameras;
  for (auto& c : raw.cameras) {  // load_cameras
    cameras.addCameraFovX(c.name, c.position, normalize(c.look_at), normalize(c.up), c.hfov * 2.0f, c.clip_near, c.clip_far);
    materials.names.push_back("camera");
    materials.bsdfs.push_back(unique<BSDF>(new CameraBSDF()));
  }
  std::vector<Mesh> meshes;
  const uint32_t materials_base = uint32_t(materials.names.size());
  for (auto& m : raw.materials) {  // load_meshes: material -> BSDF (loader.cpp:374-401)
    materials.names.push_back(m.name);
    if (m.transparency_use) {
      materials.bsdfs.push_back(unique<BSDF>(new TransmissionBSDF(m.ior, 1.0f)));
    } else if (m.mirror_use) {
      materials.bsdfs.push_back(unique<BSDF>(new ReflectionBSDF()));
    } else if (m.specular == vec3(0.0f)) {
      materials.bsdfs.push_back(unique<BSDF>(new DiffuseBSDF(m.diffuse)));
    } else {
      const vec3 total = m.specular + m.diffuse;
      if (total.x > 1.0f || total.y > 1.0f || total.z > 1.0f)
        std::cerr << "Invalid material: " << m.name << " [" << total.x << ", " << total.y << ", " << total.z << "]" << std::endl;
      materials.bsdfs.push_back(unique<BSDF>(new PhongBSDF(m.diffuse, m.specular, m.shininess)));
    }
  }
  for (auto& m : raw.meshes) meshes.push_back(mesh_from_import(m, materials_base, entity_type::mesh));
  AreaLights lights;
  const bounding_sphere_t sphere = bounding_sphere_of(meshes);  // load_lights: before the light quads are appended
  for (size_t i = 0; i < raw.lights.size(); ++i) {
    const RawLight& l = raw.lights[i];
    if (!l.is_area) continue;
    lights.addLight(l.name, materials.names.size(), l.position, normalize(l.direction), normalize(l.up), l.color, l.size, l.diffuse);
    meshes.push_back(lights.light(i).create_mesh(materials.names.size(), l.name));  // indexed by i like loader.cpp:448 (area lights first)
    materials.names.push_back(l.name);
    materials.bsdfs.push_back(lights.light(i).create_bsdf(sphere, i, l.diffuse));
  }
  return make_shared<Scene>(std::move(cameras), std::move(materials), std::move(meshes), std::move(lights));
}

}  // namespace b2pt_host
