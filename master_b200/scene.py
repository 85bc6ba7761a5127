"""Flattened scenes (.b2s) — the data format either side of the hot path.

A .b2s file is what haste::Scene holds after loadScene (loader.cpp:458-487) and before
Scene::buildAccelStructs: cameras, per-mesh vertex / tangent-frame / index arrays with the light
quads already appended as 2-triangle meshes (loader.cpp:434-456), the BSDF table and the light
table.  oracle/ref_harness `dump` writes it from a .blend through the reference's own loader;
`procedural.py` writes synthetic ones.  Layout (little endian):

    "B2SCENE1"  u32 counts[8] = cameras, meshes, vertices, indices, materials, lights, 0, 0
    cameras  x CameraRec (31 f32: position, direction, up, fovx, near, far, focal,
                          world_to_view mat3, view_to_world mat3 — column-major)
    u32 mesh_material_id[M]   u32 mesh_vertex_offset[M+1]   u32 mesh_index_offset[M+1]
    f32 vertices[3V]   f32 tangents[9V]   i32 indices[I]
    materials x B2ptMaterial (64 B)   lights x B2ptLight (80 B)
"""
import ctypes
import ctypes.util

import numpy as np

# The host side of the reference computes the camera constants with glibc's float functions
# (Cameras.cpp:77-83,112-118); numpy's float32 atan2/tan differ from them in the last bit, so call
# the same libm entry points.
_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
_libm.atan2f.restype = ctypes.c_float
_libm.atan2f.argtypes = [ctypes.c_float, ctypes.c_float]
_libm.tanf.restype = ctypes.c_float
_libm.tanf.argtypes = [ctypes.c_float]

MATERIAL_DTYPE = np.dtype([
    ("type", "<u4"), ("diffuse", "<f4", 3), ("specular", "<f4", 3), ("power", "<f4"),
    ("diffuse_probability", "<f4"), ("ext_over_int_ior", "<f4"), ("internal_ior", "<f4"),
    ("light_id", "<u4"), ("sphere", "<f4", 4),
])
LIGHT_DTYPE = np.dtype([
    ("position", "<f4", 3), ("tangent", "<f4", 9), ("size", "<f4", 2), ("exitance", "<f4", 3),
    ("diffuse", "<f4"), ("material_id", "<u4"), ("weight", "<f4"),
])
CAMERA_DTYPE = np.dtype([
    ("position", "<f4", 3), ("direction", "<f4", 3), ("up", "<f4", 3), ("fovx", "<f4"), ("near", "<f4"),
    ("far", "<f4"), ("focal", "<f4"), ("world_to_view", "<f4", 9), ("view_to_world", "<f4", 9),
])
assert MATERIAL_DTYPE.itemsize == 64 and LIGHT_DTYPE.itemsize == 80 and CAMERA_DTYPE.itemsize == 124

BSDF_CAMERA, BSDF_DIFFUSE, BSDF_PHONG, BSDF_REFLECTION, BSDF_TRANSMISSION, BSDF_LIGHT, BSDF_SUN = range(7)
ENTITY_CAMERA, ENTITY_MESH, ENTITY_LIGHT, ENTITY_EMPTY = range(4)


class FlatScene:
    """Host copy of a flattened scene; every array is C-contiguous and ABI-typed."""

    def __init__(self, cameras, mesh_material_id, mesh_vertex_offset, mesh_index_offset, vertices, tangents,
                 indices, materials, lights):
        self.cameras = np.ascontiguousarray(cameras, dtype=CAMERA_DTYPE)
        self.mesh_material_id = np.ascontiguousarray(mesh_material_id, dtype=np.uint32)
        self.mesh_vertex_offset = np.ascontiguousarray(mesh_vertex_offset, dtype=np.uint32)
        self.mesh_index_offset = np.ascontiguousarray(mesh_index_offset, dtype=np.uint32)
        self.vertices = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 3)
        self.tangents = np.ascontiguousarray(tangents, dtype=np.float32).reshape(-1, 9)
        self.indices = np.ascontiguousarray(indices, dtype=np.int32)
        self.materials = np.ascontiguousarray(materials, dtype=MATERIAL_DTYPE)
        self.lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE)
        m = len(self.mesh_material_id)
        if len(self.mesh_vertex_offset) != m + 1 or len(self.mesh_index_offset) != m + 1:
            raise ValueError("mesh offset tables must have num_meshes + 1 entries")
        if self.mesh_vertex_offset[-1] != len(self.vertices) or len(self.tangents) != len(self.vertices):
            raise ValueError("vertex / tangent arrays do not match mesh_vertex_offset")
        if self.mesh_index_offset[-1] != len(self.indices) or len(self.indices) % 3:
            raise ValueError("index array does not match mesh_index_offset")

    @property
    def num_meshes(self):
        return len(self.mesh_material_id)

    @property
    def num_triangles(self):
        return len(self.indices) // 3

    def triangle_vertices(self):
        """(T, 3, 3) float32 corner positions in (mesh, prim) order."""
        counts = np.diff(self.mesh_index_offset.astype(np.int64))
        base = np.repeat(self.mesh_vertex_offset[:-1].astype(np.int64), counts)
        return self.vertices[self.indices.astype(np.int64) + base].reshape(-1, 3, 3)

    def camera_context(self, camera_id, width, height):
        """render_context_t of Technique::render (Technique.cpp:37-45) for one camera and resolution.

        focal_length_y = 1 / tan(fovy / 2) with fovy = 2 * atan2(1 / aspect, focal)
        (Cameras.cpp:77-83,112-118), evaluated in float32 like the reference.
        """
        cam = self.cameras[camera_id]
        f32 = np.float32
        aspect = f32(width) / f32(height)
        fovy = f32(2.0) * f32(_libm.atan2f(float(f32(1.0) / aspect), float(cam["focal"])))
        focal_length_y = f32(1.0) / f32(_libm.tanf(float(fovy * f32(0.5))))
        return {
            "view_to_world": cam["view_to_world"].copy(),
            "world_to_view": cam["world_to_view"].copy(),
            "position": cam["position"].copy(),
            "focal_length_y": f32(focal_length_y),
        }

    def save(self, path):
        with open(path, "wb") as f:
            f.write(b"B2SCENE1")
            counts = np.array([len(self.cameras), self.num_meshes, len(self.vertices), len(self.indices),
                               len(self.materials), len(self.lights), 0, 0], dtype="<u4")
            f.write(counts.tobytes())
            for a in (self.cameras, self.mesh_material_id, self.mesh_vertex_offset, self.mesh_index_offset,
                      self.vertices, self.tangents, self.indices, self.materials, self.lights):
                f.write(a.tobytes())


def load_b2s(path):
    """Reads a .b2s scene; a .b2s.xz (the large stand-in scene, tests/golden/make_large_scene.py) is unpacked in memory."""
    if path.endswith(".xz"):
        import lzma
        with lzma.open(path, "rb") as f:
            buf = f.read()
    else:
        with open(path, "rb") as f:
            buf = f.read()
    if buf[:8] != b"B2SCENE1":
        raise ValueError("%s is not a .b2s scene" % path)
    nc, nm, nv, ni, nmat, nl, _, _ = np.frombuffer(buf, "<u4", 8, 8)
    off = 8 + 32

    def take(dtype, count):
        nonlocal off
        a = np.frombuffer(buf, dtype, count, off)
        off += a.nbytes
        return a.copy()

    cameras = take(CAMERA_DTYPE, nc)
    mat = take("<u4", nm)
    voff = take("<u4", nm + 1)
    ioff = take("<u4", nm + 1)
    verts = take("<f4", nv * 3)
    tans = take("<f4", nv * 9)
    idx = take("<i4", ni)
    materials = take(MATERIAL_DTYPE, nmat)
    lights = take(LIGHT_DTYPE, nl)
    if off != len(buf):
        raise ValueError("%s: trailing bytes" % path)
    return FlatScene(cameras, mat, voff, ioff, verts, tans, idx, materials, lights)
