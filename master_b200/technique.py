"""Host-side mirror of the reference's technique interface for scripts and tests.

`GpuTechnique.render(view, camera_id)` has the contract of `haste::Technique::render`
(Technique.hpp:34-39, Technique.cpp:15-77): it adds one pass (or `passes` of them) of one sample
per pixel into the caller's `dvec4` accumulator — (sum r, sum g, sum b, sum 1), row-major, y up
(ImageView.hpp:10-46) — and bumps the statistics the reference keeps (statistics.hpp:32-51).
`make_technique(scene, options)` is `makeTechnique` (make_technique.cpp:112-170): the technique is
picked from `options["technique"]` in {"PT", "BPT", "VCM", "UPG"}.

The C++ drop-in that a maintainer of the reference would compile is master_b200/host/GpuTechnique.hpp
(see INTEGRATION.md); this module is the same boundary for Python callers, over the same C ABI.
"""
import ctypes as C

import numpy as np

from . import api


class GpuTechnique:
    def __init__(self, scene, technique="PT", first_sample=0, sample_stride=1, **options):
        """scene: api.DeviceSceneHandle; options: the Options fields the techniques read (Options.hpp:30-54).

        first_sample / sample_stride (GpuOptions of master_b200/host/GpuTechnique.hpp): rank g of G passes (g, G) and the
        passes this object renders are the global sample indices g, g + G, g + 2G, ... — disjoint across ranks and across
        successive calls."""
        self.scene = scene
        self.technique = technique
        self.options = dict(options)
        self.first_sample = int(first_sample)
        self.sample_stride = max(int(sample_stride), 1)
        self.statistics = {"num_samples": 0, "num_basic_rays": 0, "num_shadow_rays": 0, "total_time": 0.0,
                           "numeric_errors": 0, "records": []}

    def sample_indices(self, passes, first_sample=None, sample_stride=None):
        """(first, stride) of the next `passes` passes, as GpuTechnique::render computes them (GpuTechnique.cpp):
        first = first_sample + num_samples * stride.  Explicit arguments override (bench.py's per-step split)."""
        stride = self.sample_stride if sample_stride is None else max(int(sample_stride), 1)
        first = self.first_sample + self.statistics["num_samples"] * stride if first_sample is None else int(first_sample)
        return first, stride

    def _params(self, width, height, camera_id, passes, first_sample=None, sample_stride=None):
        first, stride = self.sample_indices(passes, first_sample, sample_stride)
        return api.make_params(self.scene.flat, technique=self.technique, width=width, height=height, camera=camera_id,
                               first_sample=first, sample_stride=stride, num_samples=passes, **self.options)

    def _account(self, stats, passes):
        s = self.statistics
        s["num_samples"] += passes
        s["num_basic_rays"] += stats.num_basic_rays
        s["num_shadow_rays"] += stats.num_shadow_rays
        s["numeric_errors"] += stats.numeric_errors
        s["total_time"] += stats.device_ms * 1e-3
        s["records"].append({"sample_index": s["num_samples"] - 1, "frame_duration": stats.device_ms * 1e-3,
                             "clock_time": s["total_time"], "numeric_errors": stats.numeric_errors})

    def render(self, view, camera_id=0, passes=1, **kw):
        """view: float64 array (height, width, 4) in HOST memory; view += the passes (device->host copy inside)."""
        if view.dtype != np.float64 or view.ndim != 3 or view.shape[2] != 4 or not view.flags["C_CONTIGUOUS"]:
            raise ValueError("view must be a C-contiguous float64 (height, width, 4) array")
        height, width = view.shape[:2]
        p = self._params(width, height, camera_id, passes, **kw)
        stats = api.Stats()
        api.check(api.load_library().b2pt_render_host(self.scene.handle, C.byref(p), view.ctypes.data_as(C.c_void_p),
                                                      C.byref(stats)))
        self._account(stats, passes)
        return stats

    def render_device(self, accum_ptr, width, height, camera_id=0, passes=1, path_radiance_ptr=0, stream=0, **kw):
        """Same, into a DEVICE dvec4 buffer (e.g. a torch.float64 CUDA tensor's data_ptr())."""
        p = self._params(width, height, camera_id, passes, **kw)
        stats = api.Stats()
        api.check(api.load_library().b2pt_render(self.scene.handle, C.byref(p), C.c_void_p(accum_ptr),
                                                 C.c_void_p(path_radiance_ptr), C.c_void_p(stream), C.byref(stats)))
        self._account(stats, passes)
        return stats


def film_errors_device(accum_ptr, reference_ptr, width, height, window=None, stream=0):
    """(rms, abs) of a DEVICE film against a DEVICE reference image (width*height*3 floats): rms_abs_errors, or with
    window=(center_x, center_y, radius) rms_abs_errors_windowed (ImageView.cpp:60-117) — the per-pass record of
    Technique::render (Technique.cpp:68-74) without bringing the film to the host."""
    out = (C.c_float * 2)()
    win = None
    if window is not None:
        win = (C.c_int32 * 3)(*[int(v) for v in window])
    api.check(api.load_library().b2pt_film_errors(C.c_void_p(accum_ptr), C.c_void_p(reference_ptr), width, height, win, out,
                                                  C.c_void_p(stream)))
    return float(out[0]), float(out[1])


def make_technique(scene, options):
    """makeTechnique(scene, options) (make_technique.cpp:112-170) for the GPU techniques."""
    options = dict(options)
    technique = options.pop("technique", "PT")
    if technique not in api.TECHNIQUES:
        raise ValueError("unknown technique %r" % (technique,))
    return GpuTechnique(scene, technique, **options)
