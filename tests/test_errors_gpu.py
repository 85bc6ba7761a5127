"""SURVEY.md §8(f) N3 on the device: the per-pass error record (Technique.cpp:68-74) = rms_abs_errors /
rms_abs_errors_windowed (ImageView.cpp:60-117) of the device film against a reference image.

The checker is a numpy restatement of the reference loop: per pixel d = |vec3(a.xyz / a.w) - b| in float,
abs += d.x + d.y + d.z and rms += dot(d, d) added in FLOAT in row order (np.cumsum is sequential), then
sqrt(rms / num), abs / num.  The device sums the same per-pixel terms in double, so it must agree with the double sum of
those terms to float rounding (1e-6) and with the reference's float loop to that loop's own summation error (1e-3)."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import REF_HARNESS

pytestmark = pytest.mark.gpu


def reference_errors(film, ref, tmp_path, window=None):
    """rms_abs_errors / rms_abs_errors_windowed as COMPILED FROM THE REFERENCE (ImageView.cpp:60-117 inside oracle/_ref/ref_harness)."""
    h, w = film.shape[:2]
    a, b = str(tmp_path / "film.f64"), str(tmp_path / "ref.f32")
    np.ascontiguousarray(film, dtype=np.float64).tofile(a)
    np.ascontiguousarray(ref, dtype=np.float32).tofile(b)
    cmd = [REF_HARNESS, "errors", a, b, str(w), str(h)] + ([str(x) for x in window] if window else [])
    r = json.loads(subprocess.run(cmd, check=True, capture_output=True, text=True).stdout.strip().splitlines()[-1])
    return r["rms"], r["abs"]


needs_harness = pytest.mark.skipif(not os.path.exists(REF_HARNESS), reason="oracle/_ref/ref_harness not built")


@needs_harness
@pytest.mark.parametrize("width,height,window", [(37, 21, None), (512, 512, None), (37, 21, (10, 10, 4)), (37, 21, (0, 0, 3)), (37, 21, (36, 20, 5)),
                                                 (37, 21, (18, 10, 100)), (1024, 1024, (512, 100, 16))])
def test_against_the_references_compiled_function(width, height, window, tmp_path):
    """b2pt_film_errors vs the reference's own rms_abs_errors(_windowed): the per-pixel terms are the reference's arithmetic, the
    device adds them in double where the reference adds in float — agreement to the float loop's summation error."""
    import torch
    from master_b200.technique import film_errors_device
    film, ref = make(width, height, width * 3 + height)
    f, r = torch.from_numpy(film).cuda(), torch.from_numpy(ref).cuda()
    rms, ab = film_errors_device(f.data_ptr(), r.data_ptr(), width, height, window=window)
    want_rms, want_abs = reference_errors(film, ref, tmp_path, window)
    tol = 1e-3 if window is None else 2e-5
    assert rms == pytest.approx(want_rms, rel=tol) and ab == pytest.approx(want_abs, rel=tol)
    # and the numpy restatement used by the other tests IS the reference loop, to the bit
    if window is None:
        sq, a = per_pixel_terms(film, ref)
        assert reference_loop(sq, a, width * height * 3) == (pytest.approx(want_rms, rel=1e-7), pytest.approx(want_abs, rel=1e-7))


def per_pixel_terms(film, ref):
    rgb = (film[:, :, :3] / film[:, :, 3:4]).astype(np.float32)
    d = np.abs(rgb - ref)
    ab = (d[:, :, 0] + d[:, :, 1]) + d[:, :, 2]
    sq = (d[:, :, 0] * d[:, :, 0] + d[:, :, 1] * d[:, :, 1]) + d[:, :, 2] * d[:, :, 2]
    return sq.astype(np.float32), ab.astype(np.float32)


def reference_loop(sq, ab, num):
    """ImageView.cpp:70-84 / 103-116: float accumulators, row order."""
    rms = np.cumsum(sq.reshape(-1), dtype=np.float32)[-1]
    a = np.cumsum(ab.reshape(-1), dtype=np.float32)[-1]
    return float(np.sqrt(np.float32(rms) / np.float32(num))), float(np.float32(a) / np.float32(num))


def make(width, height, seed):
    rng = np.random.default_rng(seed)
    film = np.empty((height, width, 4))
    n = rng.integers(1, 64, size=(height, width)).astype(np.float64)
    film[:, :, :3] = rng.random((height, width, 3)) * 4.0 * n[:, :, None]
    film[:, :, 3] = n
    ref = (rng.random((height, width, 3)) * 4.0).astype(np.float32)
    return film, ref


@pytest.mark.parametrize("width,height", [(1, 1), (37, 21), (512, 512), (1024, 1024)])
def test_whole_image(width, height):
    import torch
    from master_b200.technique import film_errors_device
    film, ref = make(width, height, width * 7 + height)
    f, r = torch.from_numpy(film).cuda(), torch.from_numpy(ref).cuda()
    rms, ab = film_errors_device(f.data_ptr(), r.data_ptr(), width, height)
    sq, a = per_pixel_terms(film, ref)
    num = width * height * 3
    assert rms == pytest.approx(np.sqrt(sq.astype(np.float64).sum() / num), rel=1e-6)
    assert ab == pytest.approx(a.astype(np.float64).sum() / num, rel=1e-6)
    want_rms, want_abs = reference_loop(sq, a, num)
    assert rms == pytest.approx(want_rms, rel=1e-3) and ab == pytest.approx(want_abs, rel=1e-3)


@pytest.mark.parametrize("window", [(10, 10, 4), (0, 0, 3), (36, 20, 5), (18, 10, 100), (5, 5, 1)])
def test_window_around_a_trace_point(window):
    """Technique::_make_measurements: pixels max(0, c - r + 1) <= p < min(c + r, size), divided by the pixel count."""
    import torch
    from master_b200.technique import film_errors_device
    width, height = 37, 21
    film, ref = make(width, height, 11)
    f, r = torch.from_numpy(film).cuda(), torch.from_numpy(ref).cuda()
    rms, ab = film_errors_device(f.data_ptr(), r.data_ptr(), width, height, window=window)
    cx, cy, rad = window
    xs = slice(max(0, cx - rad + 1), min(cx + rad, width))
    ys = slice(max(0, cy - rad + 1), min(cy + rad, height))
    sq, a = per_pixel_terms(film[ys, xs], ref[ys, xs])
    want_rms, want_abs = reference_loop(sq, a, sq.size)
    assert rms == pytest.approx(want_rms, rel=1e-5) and ab == pytest.approx(want_abs, rel=1e-5)


def test_record_of_a_rendered_pass(flat_scenes):
    """The film of a PT pass against itself is zero error; against a shifted copy the error is the shift."""
    import torch
    from master_b200 import api
    from master_b200.technique import GpuTechnique, film_errors_device
    scene = api.DeviceSceneHandle(flat_scenes("CornellBoxDiffuse"))
    w = h = 64
    film = torch.zeros((h, w, 4), dtype=torch.float64, device="cuda")
    GpuTechnique(scene, "PT", seed=3).render_device(film.data_ptr(), w, h, passes=4)
    image = (film[:, :, :3] / film[:, :, 3:4]).float().contiguous()
    assert film_errors_device(film.data_ptr(), image.data_ptr(), w, h) == (0.0, 0.0)
    shifted = (image + 0.25).contiguous()
    rms, ab = film_errors_device(film.data_ptr(), shifted.data_ptr(), w, h)
    assert rms == pytest.approx(0.25, rel=1e-5) and ab == pytest.approx(0.25, rel=1e-5)


def test_bad_arguments():
    from master_b200 import api
    from master_b200.technique import film_errors_device
    with pytest.raises(api.B2ptError) as e:
        film_errors_device(0, 0, 4, 4)
    assert e.value.code == 1
