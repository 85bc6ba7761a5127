"""ctypes binding of include/b2pt.h (master_b200/libb2pt.so).

Thin by design: structures mirror the header field for field, functions raise B2ptError carrying
b2pt_last_error() the way the reference throws std::runtime_error / std::logic_error
(runtime_assert.cpp:7-11).  There is no CPU fallback: if the library is missing or no CUDA device
is visible, the compute calls raise.
"""
import ctypes as C
import os

import numpy as np

from .scene import FlatScene, LIGHT_DTYPE, MATERIAL_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb2pt.so")

TRACE_CLOSEST_ALL, TRACE_CLOSEST_MESH, TRACE_OCCLUDED_MESH = 0, 1, 2
TECHNIQUE_PT, TECHNIQUE_BPT, TECHNIQUE_VCM, TECHNIQUE_UPG = 0, 1, 2, 3
TECHNIQUES = {"PT": TECHNIQUE_PT, "BPT": TECHNIQUE_BPT, "VCM": TECHNIQUE_VCM, "UPG": TECHNIQUE_UPG}

RAY_DTYPE = np.dtype([("org", "<f4", 3), ("tnear", "<f4"), ("dir", "<f4", 3), ("tfar", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("geom", "<u4"), ("prim", "<u4")])
assert RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 20


class B2ptError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("b2pt error %d: %s" % (code, message))
        self.code = code


class SceneDesc(C.Structure):
    _fields_ = [
        ("num_meshes", C.c_uint32),
        ("mesh_material_id", C.c_void_p),
        ("mesh_vertex_offset", C.c_void_p),
        ("mesh_index_offset", C.c_void_p),
        ("vertices", C.c_void_p),
        ("tangents", C.c_void_p),
        ("indices", C.c_void_p),
        ("num_materials", C.c_uint32),
        ("materials", C.c_void_p),
        ("num_lights", C.c_uint32),
        ("lights", C.c_void_p),
    ]


class Camera(C.Structure):
    _fields_ = [
        ("view_to_world", C.c_float * 9),
        ("world_to_view", C.c_float * 9),
        ("position", C.c_float * 3),
        ("focal_length_y", C.c_float),
    ]


class RenderParams(C.Structure):
    _fields_ = [
        ("technique", C.c_uint32),
        ("beta", C.c_float),
        ("roulette", C.c_float),
        ("lights", C.c_float),
        ("max_path", C.c_uint32),
        ("radius", C.c_float),
        ("alpha", C.c_float),
        ("num_photons", C.c_uint32),
        ("unbiased", C.c_uint32),
        ("enable_vc", C.c_uint32),
        ("enable_vm", C.c_uint32),
        ("from_light", C.c_uint32),
        ("width", C.c_uint32),
        ("height", C.c_uint32),
        ("camera", Camera),
        ("seed", C.c_uint32),
        ("first_sample", C.c_uint32),
        ("sample_stride", C.c_uint32),
        ("num_samples", C.c_uint32),
        ("sky_horizon", C.c_float * 3),
        ("sky_zenith", C.c_float * 3),
        ("max_subpath", C.c_uint32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("num_samples", C.c_uint64),
        ("num_basic_rays", C.c_uint64),
        ("num_shadow_rays", C.c_uint64),
        ("num_paths", C.c_uint64),
        ("numeric_errors", C.c_uint64),
        ("capacity_events", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("device_ms", C.c_float),
        ("trace_ms", C.c_float),
        ("shadow_ms", C.c_float),
        ("shade_ms", C.c_float),
        ("trace_launches", C.c_uint64),
        ("shadow_launches", C.c_uint64),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


# every symbol include/b2pt.h declares; tests/test_abi.py checks the list against the header
EXPORTS = [
    "b2pt_last_error", "b2pt_abi_version", "b2pt_device_count", "b2pt_set_device", "b2pt_device_synchronize", "b2pt_scene_create", "b2pt_scene_destroy",
    "b2pt_scene_info", "b2pt_trace", "b2pt_trace_device", "b2pt_render", "b2pt_render_host", "b2pt_grid_build",
    "b2pt_grid_destroy", "b2pt_grid_query", "b2pt_grid_query_device", "b2pt_grid_info", "b2pt_film_errors", "b2pt_film_errors_static", "b2pt_film_create", "b2pt_film_destroy", "b2pt_film_data",
    "b2pt_film_clear", "b2pt_film_read", "b2pt_comm_unique_id", "b2pt_comm_init", "b2pt_reduce_accum", "b2pt_comm_destroy",
]
# parity probes of include/b2pt_debug.h (not part of the drop-in boundary)
DEBUG_EXPORTS = [
    "b2pt_debug_density", "b2pt_debug_primary", "b2pt_debug_camera", "b2pt_debug_light_samples", "b2pt_debug_draws",
    "b2pt_debug_tree_stats", "b2pt_debug_trace_counts", "b2pt_debug_l2_window", "b2pt_debug_pow",
]
DENSITY_JOB_DTYPE = np.dtype([("org", "<f4", 3), ("dir", "<f4", 3), ("target", "<f4", 3), ("radius", "<f4"), ("key", "<u4")])
DENSITY_RESULT_DTYPE = np.dtype([("trials", "<u4"), ("density", "<f4"), ("kind", "<u4"), ("t", "<f4")])
assert DENSITY_JOB_DTYPE.itemsize == 44 and DENSITY_RESULT_DTYPE.itemsize == 16

_lib = None


def load_library():
    """dlopen libb2pt.so; raises if it has not been built (python -m master_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("B2PT_LIB", LIB_PATH)  # A/B builds (master_b200/build.py --variant=...)
    if not os.path.exists(path):
        raise B2ptError(-1, "%s not found: build it with `python -m master_b200.build` "
                            "(there is no CPU fallback)" % path)
    lib = C.CDLL(path)
    lib.b2pt_last_error.restype = C.c_char_p
    lib.b2pt_abi_version.restype = C.c_int
    lib.b2pt_device_count.restype = C.c_int
    lib.b2pt_scene_create.argtypes = [C.POINTER(SceneDesc), C.POINTER(C.c_void_p)]
    lib.b2pt_scene_destroy.argtypes = [C.c_void_p]
    lib.b2pt_scene_destroy.restype = None
    lib.b2pt_scene_info.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                    C.POINTER(C.c_float)]
    lib.b2pt_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint, C.c_void_p]
    lib.b2pt_trace_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint, C.c_void_p, C.c_void_p]
    lib.b2pt_render.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(Stats)]
    lib.b2pt_render_host.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.POINTER(Stats)]
    lib.b2pt_grid_build.argtypes = [C.c_void_p, C.c_size_t, C.c_float, C.POINTER(C.c_void_p)]
    lib.b2pt_grid_destroy.argtypes = [C.c_void_p]
    lib.b2pt_grid_destroy.restype = None
    lib.b2pt_grid_query.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p, C.c_void_p]
    lib.b2pt_grid_query_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.b2pt_grid_info.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.b2pt_film_errors.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(C.c_float), C.c_void_p]
    lib.b2pt_film_errors_static.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.b2pt_film_create.argtypes = [C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p)]
    lib.b2pt_film_destroy.argtypes = [C.c_void_p]
    lib.b2pt_film_destroy.restype = None
    lib.b2pt_film_data.argtypes = [C.c_void_p]
    lib.b2pt_film_data.restype = C.c_void_p
    lib.b2pt_film_clear.argtypes = [C.c_void_p, C.c_void_p]
    lib.b2pt_film_read.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.b2pt_comm_unique_id.argtypes = [C.c_void_p]
    lib.b2pt_comm_init.argtypes = [C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
    lib.b2pt_reduce_accum.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    lib.b2pt_comm_destroy.argtypes = [C.c_void_p]
    lib.b2pt_comm_destroy.restype = None
    lib.b2pt_debug_density.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.b2pt_debug_primary.argtypes = [C.POINTER(RenderParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.b2pt_debug_camera.argtypes = [C.POINTER(RenderParams), C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    lib.b2pt_debug_light_samples.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.b2pt_debug_draws.argtypes = [C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.b2pt_debug_pow.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    lib.b2pt_debug_tree_stats.argtypes = [C.c_void_p, C.c_void_p]
    lib.b2pt_debug_trace_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint, C.c_void_p]
    lib.b2pt_debug_l2_window.argtypes = [C.c_void_p, C.c_void_p]
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise B2ptError(rc, load_library().b2pt_last_error().decode("utf-8", "replace"))


def device_count():
    return load_library().b2pt_device_count()


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class DeviceSceneHandle:
    """Owns a b2pt_scene_t: the scene uploaded to HBM with its BVH built there."""

    def __init__(self, flat: FlatScene):
        lib = load_library()
        self.flat = flat
        d = SceneDesc()
        d.num_meshes = flat.num_meshes
        d.mesh_material_id = _ptr(flat.mesh_material_id)
        d.mesh_vertex_offset = _ptr(flat.mesh_vertex_offset)
        d.mesh_index_offset = _ptr(flat.mesh_index_offset)
        d.vertices = _ptr(flat.vertices)
        d.tangents = _ptr(flat.tangents)
        d.indices = _ptr(flat.indices)
        d.num_materials = len(flat.materials)
        d.materials = _ptr(flat.materials)
        d.num_lights = len(flat.lights)
        d.lights = _ptr(flat.lights) if len(flat.lights) else None
        assert flat.materials.dtype == MATERIAL_DTYPE and flat.lights.dtype == LIGHT_DTYPE
        h = C.c_void_p()
        check(lib.b2pt_scene_create(C.byref(d), C.byref(h)))
        self.handle = h

    def info(self):
        nt, nn, nb, ms = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_float()
        check(load_library().b2pt_scene_info(self.handle, C.byref(nt), C.byref(nn), C.byref(nb), C.byref(ms)))
        return {"triangles": nt.value, "nodes": nn.value, "bvh_bytes": nb.value, "build_ms": ms.value}

    def trace(self, rays, mode):
        """rays: RAY_DTYPE array in host memory -> HIT_DTYPE array (copies both ways)."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.empty(len(rays), dtype=HIT_DTYPE)
        check(load_library().b2pt_trace(self.handle, _ptr(rays), len(rays), mode, _ptr(hits)))
        return hits

    def trace_device(self, rays_ptr, n, mode, hits_ptr, stream=0):
        check(load_library().b2pt_trace_device(self.handle, C.c_void_p(rays_ptr), n, mode, C.c_void_p(hits_ptr),
                                               C.c_void_p(stream)))

    # ---- parity probes (include/b2pt_debug.h)
    def debug_density(self, jobs):
        jobs = np.ascontiguousarray(jobs, dtype=DENSITY_JOB_DTYPE)
        out = np.zeros(len(jobs), dtype=DENSITY_RESULT_DTYPE)
        check(load_library().b2pt_debug_density(self.handle, _ptr(jobs), len(jobs), _ptr(out)))
        return out

    def debug_light_samples(self, keys):
        keys = np.ascontiguousarray(keys, dtype=np.uint32)
        pos = np.zeros((len(keys), 3), dtype=np.float32)
        mat, draws = np.zeros(len(keys), dtype=np.uint32), np.zeros(len(keys), dtype=np.uint32)
        check(load_library().b2pt_debug_light_samples(self.handle, _ptr(keys), len(keys), _ptr(pos), _ptr(mat), _ptr(draws)))
        return pos, mat, draws

    def tree_stats(self):
        s = np.zeros(6, dtype=np.float64)
        check(load_library().b2pt_debug_tree_stats(self.handle, _ptr(s)))
        return {"sah_cost": s[0], "inner_nodes": int(s[1]), "leaf_children": int(s[2]), "triangle_refs": int(s[3]),
                "expected_triangle_tests": s[4], "expected_node_visits": s[5]}

    def l2_window(self):
        s = np.zeros(6, dtype=np.float64)
        check(load_library().b2pt_debug_l2_window(self.handle, _ptr(s)))
        return {"bvh_block_bytes": int(s[0]), "window_bytes": int(s[1]), "hit_ratio": float(s[2]), "persisting_l2_max": int(s[3]),
                "window_max": int(s[4]), "l2_bytes": int(s[5])}

    def trace_counts(self, rays, mode):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        counts = np.zeros((len(rays), 2), dtype=np.uint32)
        check(load_library().b2pt_debug_trace_counts(self.handle, _ptr(rays), len(rays), mode, _ptr(counts)))
        return counts

    def close(self):
        if getattr(self, "handle", None):
            load_library().b2pt_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_params(flat, technique="PT", width=512, height=512, camera=0, beta=None, roulette=0.9, lights=1.0, max_path=0,
                radius=0.01, alpha=0.75, num_photons=0, enable_vc=True, enable_vm=True, from_light=True, seed=1234,
                first_sample=0, sample_stride=1, num_samples=1, sky_horizon=(0, 0, 0), sky_zenith=(0, 0, 0),
                max_subpath=0, focal_length_y=None):
    """Options (Options.hpp:19-71) + render_context_t (Technique.hpp:14-27) -> B2ptRenderParams."""
    p = RenderParams()
    p.technique = TECHNIQUES[technique]
    p.beta = (1.0 if technique == "PT" else 2.0) if beta is None else beta
    p.roulette, p.lights, p.max_path = roulette, lights, max_path
    p.radius, p.alpha, p.num_photons = radius, alpha, num_photons
    p.unbiased = 1 if technique == "UPG" else 0
    p.enable_vc, p.enable_vm, p.from_light = int(enable_vc), int(enable_vm), int(from_light)
    p.width, p.height = width, height
    ctx = flat.camera_context(camera, width, height)
    p.camera.view_to_world[:] = ctx["view_to_world"].tolist()
    p.camera.world_to_view[:] = ctx["world_to_view"].tolist()
    p.camera.position[:] = ctx["position"].tolist()
    p.camera.focal_length_y = float(ctx["focal_length_y"] if focal_length_y is None else focal_length_y)
    p.seed, p.first_sample, p.sample_stride, p.num_samples = seed, first_sample, sample_stride, num_samples
    p.sky_horizon[:] = list(sky_horizon)
    p.sky_zenith[:] = list(sky_zenith)
    p.max_subpath = max_subpath
    return p


class PhotonGrid:
    """b2pt_grid_*: the photon index (v3::HashGrid3D<T>::build / rQuery, HashGrid3D.hpp:181-325) on the device."""

    def __init__(self, points, radius):
        pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        h = C.c_void_p()
        check(load_library().b2pt_grid_build(_ptr(pts), len(pts), C.c_float(radius), C.byref(h)))
        self.handle, self.radius, self.num_points = h, float(radius), len(pts)

    def info(self):
        n, r, ms = C.c_uint64(), C.c_float(), C.c_float()
        check(load_library().b2pt_grid_info(self.handle, C.byref(n), C.byref(r), C.byref(ms)))
        return {"points": n.value, "radius": r.value, "build_ms": ms.value}

    def query_device(self, queries_ptr, nq, max_per_query, counts_ptr, indices_ptr, stream=0):
        check(load_library().b2pt_grid_query_device(self.handle, C.c_void_p(queries_ptr), nq, max_per_query, C.c_void_p(counts_ptr),
                                                    C.c_void_p(indices_ptr), C.c_void_p(stream)))

    def query(self, queries, max_per_query=64):
        """-> (counts[nq], indices[nq, max_per_query]); indices[q, :min(counts[q], max)] are valid, unordered."""
        q = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, 3)
        counts = np.zeros(len(q), dtype=np.uint32)
        indices = np.zeros((len(q), max(max_per_query, 1)), dtype=np.uint32)
        check(load_library().b2pt_grid_query(self.handle, _ptr(q), len(q), max_per_query, _ptr(counts), _ptr(indices)))
        return counts, indices

    def close(self):
        if getattr(self, "handle", None):
            load_library().b2pt_grid_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def debug_primary(params, pixels, samples):
    pixels, samples = np.ascontiguousarray(pixels, dtype=np.uint32), np.ascontiguousarray(samples, dtype=np.uint32)
    out = np.zeros((len(pixels), 3), dtype=np.float32)
    check(load_library().b2pt_debug_primary(C.byref(params), _ptr(pixels), _ptr(samples), len(pixels), _ptr(out)))
    return out


def debug_camera(params, positions):
    pos = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 2)
    dirs, pix = np.zeros((len(pos), 3), dtype=np.float32), np.zeros((len(pos), 2), dtype=np.float32)
    check(load_library().b2pt_debug_camera(C.byref(params), _ptr(pos), len(pos), _ptr(dirs), _ptr(pix)))
    return dirs, pix


def debug_draws(key, n):
    raw, fl, db = np.zeros(n, dtype=np.uint32), np.zeros(n, dtype=np.float32), np.zeros(n, dtype=np.float64)
    check(load_library().b2pt_debug_draws(C.c_uint32(key), n, _ptr(raw), _ptr(fl), _ptr(db)))
    return raw, fl, db


def debug_pow(x, y):
    """(table-driven cr_powf, math-library form) of the device on float arrays x, y (b2pt_debug_pow)."""
    x = np.ascontiguousarray(x, dtype=np.float32).ravel()
    y = np.ascontiguousarray(y, dtype=np.float32).ravel()
    assert x.shape == y.shape
    tab, lib = np.zeros_like(x), np.zeros_like(x)
    check(load_library().b2pt_debug_pow(_ptr(x), _ptr(y), len(x), _ptr(tab), _ptr(lib)))
    return tab, lib


class Comm:
    """b2pt_comm_*: the communicator of the film reduce (NCCL behind the C ABI).  `id_bytes` comes from Comm.unique_id() on rank 0."""

    def __init__(self, nranks, rank, id_bytes=None):
        h = C.c_void_p()
        buf = C.create_string_buffer(bytes(id_bytes), 128) if id_bytes is not None else None
        check(load_library().b2pt_comm_init(nranks, rank, buf, C.byref(h)))
        self.handle, self.nranks, self.rank = h, nranks, rank

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        check(load_library().b2pt_comm_unique_id(buf))
        return buf.raw

    def reduce_accum(self, accum_ptr, num_doubles, root=-1, stream=0):
        check(load_library().b2pt_reduce_accum(self.handle, C.c_void_p(accum_ptr), num_doubles, root, C.c_void_p(stream)))

    def close(self):
        if getattr(self, "handle", None):
            load_library().b2pt_comm_destroy(self.handle)
            self.handle = None
