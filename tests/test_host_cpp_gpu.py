"""The C++ drop-in: the reference's own Technique interface (haste::Technique::render, driven like
Application::render / Framework::runBatch do) with master_b200/host/GpuTechnique plugged in.  The binary is the
reference's classes + GpuTechnique.cpp + libb2pt.so (master_b200/host/build_host.sh)."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_path

HOST = os.path.join(ROOT, "master_b200", "host", "_build", "b2pt_master")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(HOST), reason="master_b200/host/_build/b2pt_master not built")]


@pytest.mark.parametrize("name,tech", [("CornellBoxDiffuse", "PT"), ("CornellBoxDiffuse", "BPT"), ("CornellBoxSpecular", "BPT")])
def test_gpu_technique_behind_the_reference_interface(name, tech, tmp_path):
    g = np.load(golden_path("render_%s_%s.npz" % (name, tech)))
    w, h, spp = int(g["width"]), int(g["height"]), g["paths"].shape[0]
    out = str(tmp_path / "accum.f64")
    r = subprocess.run([HOST, golden_path("scenes", name + ".b2s"), "technique=" + tech, "width=%d" % w, "height=%d" % h,
                        "num_samples=%d" % spp, "seed=%d" % int(g["seed"]), "beta=%d" % (1 if tech == "PT" else 2), "out=" + out],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    accum = np.fromfile(out, dtype=np.float64).reshape(-1, 4)
    ref = g["accum"]
    assert info["num_samples"] == spp and np.all(accum[:, 3] == ref[:, 3])
    # same streams as the golden: pixels whose paths took the same decisions agree to 1e-4
    scale = np.maximum(np.abs(ref[:, :3]).max(axis=1), 1e-6)
    close = (np.abs(accum[:, :3] - ref[:, :3]).max(axis=1) / scale <= 1e-4).mean()
    print(json.dumps(dict(scene=name, technique=tech, pixels_within_1e4=float(close), info=info)))
    assert close >= 0.999  # measured 100 %: the same paths as the golden, summed per pixel
    np.testing.assert_allclose(accum[:, :3].mean(axis=0), ref[:, :3].mean(axis=0), rtol=2e-3)


def test_missing_scene_is_an_exception_not_a_crash():
    r = subprocess.run([HOST, "/nonexistent.b2s"], capture_output=True, text=True)
    assert r.returncode == 1 and "cannot open" in r.stderr


def test_exr_checkpoint_and_continue(tmp_path):
    """the host driver writes the reference's EXR checkpoint (R, G, B, denom + options / statistics /
    records attributes) and resumes from it like `master continue`: 2 passes + 2 continued passes == 4 passes."""
    from master_b200 import exr
    scene = golden_path("scenes", "CornellBoxDiffuse.b2s")
    # the reference's command line: master <in> --PT --resolution=WxH --num-samples=n --seed=s --output=file; master continue <file>
    a, b, c = (str(tmp_path / n) for n in ("two.exr", "four_continued.exr", "four.exr"))
    base = [scene, "--PT", "--resolution=48x40", "--seed=77", "--batch"]
    for args in (base + ["--num-samples=2", "--output=" + a], ["continue", a, "--num-samples=4", "--output=" + b],
                 base + ["--num-samples=4", "--output=" + c]):
        r = subprocess.run([HOST] + args, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    m2, d2, has_denom = exr.load_exr(a)
    mb, db, _ = exr.load_exr(b)
    mc, dc, _ = exr.load_exr(c)
    assert has_denom and d2.shape == (40, 48, 4) and np.all(d2[:, :, 3] == 2) and np.all(db[:, :, 3] == 4) and np.all(dc[:, :, 3] == 4)
    assert m2["statistics.num_samples"] == "2" and mb["statistics.num_samples"] == "4" and mc["options.technique"] == "PT"
    assert "records[1].frame_duration" in m2 and "records[3].frame_duration" in mb
    # the same paths were traced; b2pt_render_host attributes the rays of a batch rendered ahead to its passes in proportion
    assert int(mb["statistics.num_basic_rays"]) == pytest.approx(int(mc["statistics.num_basic_rays"]), rel=0.01)
    np.testing.assert_allclose(db, dc, rtol=1e-5, atol=1e-6)  # the checkpoint holds float32 sums
    assert np.abs(dc[:, :, :3]).sum() > 0


def test_max_path_reaches_the_technique(tmp_path):
    """--max-path (PT.cpp:23-40) is parsed and forwarded: the C++ drop-in reproduces the oracle's max_path = 3 golden."""
    g = np.load(golden_path("render_CornellBoxDiffuse_PT_maxpath3.npz"))
    w, h, spp = int(g["width"]), int(g["height"]), int(g["spp"])
    out = str(tmp_path / "accum.f64")
    r = subprocess.run([HOST, golden_path("scenes", "CornellBoxDiffuse.b2s"), "--PT", "--resolution=%dx%d" % (w, h), "--num-samples=%d" % spp,
                        "--seed=%d" % int(g["seed"]), "--max-path=3", "out=" + out], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    accum = np.fromfile(out, dtype=np.float64).reshape(-1, 4)
    np.testing.assert_allclose(accum[:, :3].mean(axis=0), g["accum"][:, :3].mean(axis=0), rtol=1e-5)


def test_multi_gpu_driver_on_one_device_equals_the_single_gpu_film(tmp_path):
    """`--gpus=1` takes the resident-film path (b2pt_film_* + b2pt_reduce_accum with one rank): same film as the host-view path."""
    scene = golden_path("scenes", "CornellBoxDiffuse.b2s")
    a, b = str(tmp_path / "a.f64"), str(tmp_path / "b.f64")
    base = [HOST, scene, "--BPT", "--beta=2", "--resolution=40x32", "--num-samples=6", "--seed=5"]
    ra = subprocess.run(base + ["out=" + a], capture_output=True, text=True)
    assert ra.returncode == 0, ra.stderr
    import ctypes
    os.environ.setdefault("B2PT_FORCE_MULTI", "1")
    rb = subprocess.run(base + ["--gpus=1", "out=" + b], capture_output=True, text=True, env=dict(os.environ, B2PT_FORCE_MULTI="1"))
    assert rb.returncode == 0, rb.stderr
    fa, fb = np.fromfile(a, dtype=np.float64).reshape(-1, 4), np.fromfile(b, dtype=np.float64).reshape(-1, 4)
    assert np.all(fa[:, 3] == 6) and np.all(fb[:, 3] == 6)
    np.testing.assert_allclose(fb, fa, rtol=1e-6, atol=1e-9)  # atomics order only
    assert json.loads(rb.stdout.strip().splitlines()[-1])["gpus"] == 1
