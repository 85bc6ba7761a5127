"""north_star level 3 on a converged image: CornellBoxDiffuse PT at 512 x 512, 256 spp on the GPU against
the image the oracle rendered at the same spp with the same streams (the stand-in for the stripped reference/CornellBoxDiffuse.PT.exr;
tests/golden/image_CornellBoxDiffuse_PT_512_256spp.npz, 214 s of the reference on one host thread), and the reference's own
consistency check across techniques (PT = BPT = VCM on CornellBoxDiffuse to three digits)."""
import json

import numpy as np
import pytest

from conftest import golden_path
from test_pt_gpu import gpu_render

pytestmark = pytest.mark.gpu


def test_converged_pt_image_matches_the_oracles(device_scenes, flat_scenes):
    g = np.load(golden_path("image_CornellBoxDiffuse_PT_512_256spp.npz"))
    w, h, spp = int(g["width"]), int(g["height"]), int(g["spp"])
    accum, _, stats = gpu_render(device_scenes("CornellBoxDiffuse"), flat_scenes("CornellBoxDiffuse"), "PT", w, h, spp, int(g["seed"]),
                                 per_path=False, beta=float(g["beta"]))
    assert np.all(accum[:, 3] == spp)
    ours, ref = accum[:, :3] / accum[:, 3:4], g["rgb"].astype(np.float64)
    rmse = float(np.sqrt(np.mean((ours - ref) ** 2)) / np.sqrt(np.mean(ref ** 2)))
    off = float((np.abs(ours - ref).max(axis=1) > 1e-3 * np.maximum(np.abs(ref).max(axis=1), 1e-3)).mean())
    info = json.loads(str(g["info"]))
    print(json.dumps(dict(rel_rmse=rmse, pixels_off_1e3=off, mean_ours=ours.mean(axis=0).tolist(), mean_ref=ref.mean(axis=0).tolist(),
                          rays=[stats["num_basic_rays"], stats["num_shadow_rays"]], ref_rays=[info["basic_rays"], info["shadow_rays"]])))
    # SURVEY.md §8(d) allows 2 % at 1024 spp against an independent render; with the same streams the two films are the same
    # paths, so the bound here is 0.5 % (measured on B200: 2.7e-6 — a handful of flipped paths in 67 M) and the closest-hit counts agree to 0.1 %
    assert rmse <= 5e-3
    np.testing.assert_allclose(ours.mean(axis=0), ref.mean(axis=0), rtol=1e-3)
    assert stats["num_basic_rays"] == pytest.approx(info["basic_rays"], rel=1e-3)
    assert 0.5 * info["shadow_rays"] <= stats["num_shadow_rays"] <= info["shadow_rays"]  # occlusion rays of zero contributions are not shot here


def test_cross_technique_means_agree(device_scenes, flat_scenes):
    """PT, BPT, VCM and UPG converge to the same image: CornellBoxDiffuse 256 x 256 means agree to three digits (SURVEY.md §4;
    the reference's own numbers: PT 0.13754 / BPT 0.13743 / UPG 0.13759 in red; VCM is biased by design, r = 0.025: 0.13817)."""
    sc, fl = device_scenes("CornellBoxDiffuse"), flat_scenes("CornellBoxDiffuse")
    means = {}
    for tech, spp, kw in (("PT", 64, dict(beta=1.0)), ("BPT", 32, dict(beta=2.0)), ("UPG", 8, dict(beta=2.0, radius=0.01)),
                          ("VCM", 16, dict(beta=2.0, radius=0.025))):
        accum, _, _ = gpu_render(sc, fl, tech, 256, 256, spp, 4321, per_path=False, **kw)
        means[tech] = (accum[:, :3].sum(axis=0) / accum[:, 3].sum()).tolist()
    print(json.dumps(means))
    for tech in ("BPT", "UPG"):
        np.testing.assert_allclose(means[tech], means["PT"], rtol=5e-3)
    np.testing.assert_allclose(means["VCM"], means["PT"], rtol=1.5e-2)
    np.testing.assert_allclose(means["PT"], [0.13754, 0.09063, 0.02580], rtol=5e-3)  # BASELINE.md §2, the reference's CPU run
