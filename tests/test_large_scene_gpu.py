"""Parity on the workload that is benchmarked: scenes/_large/LivingRoom.b2s.xz (567 304 triangles; the stand-in for the stripped
CrytekSponza.blend; built from the reference's models/ by tests/golden/make_large_scene.py).

* ray level: >= 1 M rays per class (camera / cosine bounce / shadow) against Embree itself, run on this host by the shipped
  oracle/_ref/ref_harness — ids >= 99.99 %, t / u / v bit-exact on the same triangle (Scene.cpp:151-226);
* per path: PT and BPT at 128 x 128 x 2 spp against goldens the oracle rendered from the same scene file with the same streams.

On this scene (Phong materials everywhere, 567 k triangles) the reference's OWN per-path result depends on the last bit of the host's
float libm: the unmodified reference linked against glibc and the same objects linked against a correctly rounded float libm
(oracle/cr_libm.cpp -> ref_harness_crlibm; sinf, cosf, powf, asinf, atan2f evaluated in double and rounded once) disagree on 48 of
32 768 PT paths and 82 BPT paths — a 1-ulp difference in a sampled direction picks another triangle or falls on the other side of
a roulette threshold.  The device rounds those functions correctly (csrc/common.cuh: cr_*), so it is held to >= 99.99 % against
the correctly rounded oracle and to the oracle-vs-oracle level (>= 99.7 %) against the glibc one."""
import json
import os
import sys

import numpy as np
import pytest

from conftest import REF_HARNESS, ROOT, golden_path
from test_pt_gpu import gpu_render, image_rel_rmse, path_agreement
from test_trace_gpu import compare_closest

pytestmark = pytest.mark.gpu

LARGE = os.path.join(ROOT, "scenes", "_large", "LivingRoom.b2s.xz")
needs_scene = pytest.mark.skipif(not os.path.exists(LARGE), reason="scenes/_large/LivingRoom.b2s.xz absent (built from the reference's models/)")


@pytest.fixture(scope="module")
def large():
    from master_b200 import api
    from master_b200.scene import load_b2s
    flat = load_b2s(LARGE)
    scene = api.DeviceSceneHandle(flat)
    yield flat, scene
    scene.close()


@needs_scene
@pytest.mark.skipif(not os.path.exists(REF_HARNESS), reason="oracle/_ref/ref_harness not built")
def test_livingroom_live_embree_parity_1m(large, tmp_path):
    sys.path.insert(0, golden_path())
    sys.path.insert(0, ROOT)
    import bench
    import make_golden as mg
    from master_b200 import api
    flat, scene = large
    mg.TMP = str(tmp_path)
    plain = bench.plain_scene_file(LARGE)
    sets = mg.make_ray_sets(flat, plain, 1024, 1024, seed=11)
    assert len(sets["camera_rays"]) >= 1 << 20 and len(sets["bounce_rays"]) >= 1000000 and len(sets["shadow_rays"]) >= 1000000
    compare_closest(scene.trace(sets["camera_rays"], api.TRACE_CLOSEST_ALL), sets["camera_hits"], "LivingRoom 1M camera")
    compare_closest(scene.trace(sets["bounce_rays"], api.TRACE_CLOSEST_ALL), sets["bounce_hits"], "LivingRoom 1M bounce/all")
    compare_closest(scene.trace(sets["bounce_rays"], api.TRACE_CLOSEST_MESH), sets["bounce_mesh_hits"], "LivingRoom 1M bounce/mesh")
    occ = scene.trace(sets["shadow_rays"], api.TRACE_OCCLUDED_MESH)
    agree = (occ["geom"] == sets["shadow_hits"]["geom"]).mean()
    print(json.dumps(dict(what="LivingRoom 1M shadow", rays=len(occ), agree=float(agree))))
    assert agree >= 0.9999
    # batches of this size run through trace_queue8 (the integrators' traversal); one ray per thread and trace_dyn8 return the same hits
    try:
        for form, label in (("0", "one ray per thread"), ("2", "dyn8")):
            os.environ["B2PT_TRACE_STREAM"] = form
            compare_closest(scene.trace(sets["bounce_rays"], api.TRACE_CLOSEST_ALL), sets["bounce_hits"], "LivingRoom 1M bounce/all " + label)
    finally:
        del os.environ["B2PT_TRACE_STREAM"]


@needs_scene
@pytest.mark.parametrize("tech,libm", [("PT", "crlibm"), ("BPT", "crlibm"), ("PT", "glibc"), ("BPT", "glibc")])
def test_livingroom_per_path(tech, libm, large):
    flat, scene = large
    g = np.load(golden_path("render_LivingRoom_%s%s.npz" % (tech, "_crlibm" if libm == "crlibm" else "")))
    other = np.load(golden_path("render_LivingRoom_%s%s.npz" % (tech, "" if libm == "crlibm" else "_crlibm")))
    w, h, spp = int(g["width"]), int(g["height"]), int(g["spp"])
    info = json.loads(str(g["info"]))
    accum, paths, stats = gpu_render(scene, flat, tech, w, h, spp, int(g["seed"]), beta=float(g["beta"]))
    frac, err = path_agreement(paths, g["paths"])
    mean_ref, mean_ours = g["accum"][:, :3].mean(axis=0), accum[:, :3].mean(axis=0)
    oracle_vs_oracle, _ = path_agreement(other["paths"], g["paths"])
    out = dict(scene="LivingRoom", tech=tech, oracle_libm=libm, paths=int(err.size), within_1e4=frac, flips=int((err > 1e-4).sum()),
               glibc_oracle_vs_crlibm_oracle=oracle_vs_oracle,
               rays=[stats["num_basic_rays"], stats["num_shadow_rays"]], ref_rays=[info["basic_rays"], info["shadow_rays"]],
               mean_ref=mean_ref.tolist(), mean_ours=mean_ours.tolist(), rel_rmse=image_rel_rmse(accum, g["accum"]))
    if tech == "BPT":
        light_ours = accum[:, :3] - paths.astype(np.float64).sum(axis=0)
        scale = max(np.abs(g["light"]).max(), 1e-6)
        out["light_pixels_off"] = float((np.abs(light_ours - g["light"]).max(axis=1) > 1e-4 * scale + 1e-4 * np.abs(g["light"]).max(axis=1)).mean())
    print(json.dumps(out))
    if libm == "crlibm":
        assert int((err > 1e-4).sum()) <= 3  # >= 99.99 %: every path takes the decisions of the reference on a correctly rounded libm
    else:
        assert frac >= 0.997 and frac >= oracle_vs_oracle - 5e-4  # no worse than the reference against itself on another libm
    assert np.array_equal(accum[:, 3], g["accum"][:, 3])
    np.testing.assert_allclose(mean_ours, mean_ref, rtol=5e-3, atol=1e-6)
    if tech == "BPT":
        assert out["light_pixels_off"] <= 0.01
    else:  # a ray whose sampled BSDF throughput is already zero (a Phong sample below the surface) is not shot here: fewer rays, same paths
        assert 0.85 * info["basic_rays"] <= stats["num_basic_rays"] <= info["basic_rays"]


@needs_scene
@pytest.mark.parametrize("libm", ["crlibm", "glibc"])
def test_livingroom_vcm_per_path(libm, large):
    """VCM (r0 = 0.025, the radius of the bench's config 5 lines) on the bench scene: one flipped photon path changes every eye path that
    merges with it, so the bar is set beside what the reference does against ITSELF when only its libm changes."""
    flat, scene = large
    g = np.load(golden_path("render_LivingRoom_VCM%s.npz" % ("_crlibm" if libm == "crlibm" else "")))
    other = np.load(golden_path("render_LivingRoom_VCM%s.npz" % ("" if libm == "crlibm" else "_crlibm")))
    w, h, spp = int(g["width"]), int(g["height"]), int(g["spp"])
    accum, paths, stats = gpu_render(scene, flat, "VCM", w, h, spp, int(g["seed"]), beta=float(g["beta"]), radius=float(g["radius"]))
    frac, err = path_agreement(paths, g["paths"])
    oracle_vs_oracle, _ = path_agreement(other["paths"], g["paths"])
    mean_ref, mean_ours = g["accum"][:, :3].mean(axis=0), accum[:, :3].mean(axis=0)
    print(json.dumps(dict(scene="LivingRoom", tech="VCM", oracle_libm=libm, paths=int(err.size), within_1e4=frac, flips=int((err > 1e-4).sum()),
                          glibc_oracle_vs_crlibm_oracle=oracle_vs_oracle, mean_ref=mean_ref.tolist(), mean_ours=mean_ours.tolist(),
                          rel_rmse=image_rel_rmse(accum, g["accum"]), capacity_events=stats["capacity_events"])))
    assert stats["capacity_events"] == 0
    assert np.array_equal(accum[:, 3], g["accum"][:, 3])
    if libm == "crlibm":
        assert frac >= 0.999  # measured 99.98 % (7 of 32 768); VCM's bar on the small scenes is 99.5 % (tests/test_vcm_gpu.py)
    else:
        assert frac >= min(0.995, oracle_vs_oracle - 2e-3)  # no worse than the reference against itself on another libm
    np.testing.assert_allclose(mean_ours, mean_ref, rtol=1e-2, atol=1e-6)
