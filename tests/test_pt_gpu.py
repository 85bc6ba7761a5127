"""Per-path and image parity of the wavefront path tracer against the UNMODIFIED reference
(north_star levels 2 and 3), on the committed goldens made by tests/golden/make_golden.py."""
import ctypes as C
import json

import numpy as np
import pytest

from conftest import golden_path

pytestmark = pytest.mark.gpu


def gpu_render(scene, flat, technique, width, height, num_samples, seed, per_path=True, **kw):
    import torch
    from master_b200 import api
    p = api.make_params(flat, technique=technique, width=width, height=height, num_samples=num_samples, seed=seed, **kw)
    accum = torch.zeros(width * height, 4, dtype=torch.float64, device="cuda")
    paths = torch.zeros(num_samples, width * height, 3, dtype=torch.float32, device="cuda") if per_path else None
    stats = api.Stats()
    api.check(api.load_library().b2pt_render(scene.handle, C.byref(p), C.c_void_p(accum.data_ptr()),
                                             C.c_void_p(paths.data_ptr() if per_path else 0), None, C.byref(stats)))
    return accum.cpu().numpy(), (paths.cpu().numpy() if per_path else None), stats.as_dict()


def image_rel_rmse(accum_ours, accum_ref):
    """north_star level 3: relative RMSE between two films at equal spp (rgb / count per pixel)."""
    a = accum_ours[:, :3] / np.maximum(accum_ours[:, 3:4], 1)
    b = accum_ref[:, :3] / np.maximum(accum_ref[:, 3:4], 1)
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.sqrt(np.mean(b ** 2)), 1e-12))


def path_agreement(ours, ref, rel=1e-4):
    """fraction of pixel-samples whose radiance matches within `rel` (relative to the path's own magnitude)."""
    scale = np.maximum(np.abs(ref).max(axis=-1), 1e-6)
    err = np.abs(ours - ref).max(axis=-1) / scale
    return float((err <= rel).mean()), err


@pytest.mark.parametrize("name", ["CornellBoxDiffuse", "CornellBoxSpecular", "MetalRings", "TestCaseFurnace", "TestCase9", "TestCase12"])
def test_pt_per_path_radiance(name, device_scenes, flat_scenes):
    g = np.load(golden_path("render_%s_PT.npz" % name))
    w, h, spp = int(g["width"]), int(g["height"]), g["paths"].shape[0]
    info = json.loads(str(g["info"]))
    accum, paths, stats = gpu_render(device_scenes(name), flat_scenes(name), "PT", w, h, spp, int(g["seed"]), beta=1.0)
    frac, err = path_agreement(paths, g["paths"])
    mean_ref, mean_ours = g["accum"][:, :3].mean(axis=0), accum[:, :3].mean(axis=0)
    print(json.dumps(dict(scene=name, paths=int(err.size), within_1e4=frac, rays=[stats["num_basic_rays"], stats["num_shadow_rays"]],
                          ref_rays=[info["basic_rays"], info["shadow_rays"]], mean_ref=mean_ref.tolist(), mean_ours=mean_ours.tolist())))
    # hits are bit-identical to Embree's (rcp_embree, trace.cuh) and the float transcendentals are evaluated in double
    # (common.cuh: libm parity), so every path takes the reference's decisions: measured 100 % on all six scenes.  The assert
    # allows a counted handful (1 path in 5 000) for the last bit of the host's libm, so a regression cannot hide
    flips = int((err > 1e-4).sum())
    print(json.dumps(dict(scene=name, flips=flips)))
    assert flips <= max(1, err.size // 5000)
    assert np.all(accum[:, 3] == spp)
    np.testing.assert_allclose(mean_ours, mean_ref, rtol=2e-3, atol=1e-5)
    # image level, equal spp, same streams: the films agree to 1 % relative RMSE (a single flipped path would show)
    assert image_rel_rmse(accum, g["accum"]) <= 1e-2
