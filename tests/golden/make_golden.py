#!/usr/bin/env python
"""Regenerates the committed fixtures under tests/golden/ from the UNMODIFIED reference.

Needs /root/reference (scenes + sources) and oracle/_ref/ref_harness (oracle/build_ref.sh).
Everything written here is an OUTPUT OF THE REFERENCE'S OWN CODE (Assimp-fork loader, Embree 2.12
intersection, PT/BPT/UPG techniques) driven with the per-(pixel, sample) RNG streams of
oracle/rng_shim.h; nothing from the product library is involved.

  scenes/<name>.b2s              flattened scene (ref_harness dump)
  rays_<name>.npz                rays + Embree hits for the three ray classes of SURVEY.md §8(d):
                                 camera rays, cosine bounce rays (closest hit, all / mesh mask),
                                 shadow segments (occlusion)
  render_<name>_<tech>.npz       per-path eye radiance, accumulated image, light image, ray counts
"""
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
sys.path.insert(0, ROOT)
from master_b200.api import HIT_DTYPE, RAY_DTYPE  # noqa: E402
from master_b200.scene import load_b2s  # noqa: E402

REF = os.environ.get("REF", "/root/reference")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
TMP = os.environ.get("TMPDIR", "/tmp") + "/b2pt_golden"

SMALL_SCENES = ["CornellBoxDiffuse", "CornellBoxSpecular", "TestCaseFurnace", "TestCase9", "MetalRings"]
SUN_SCENES = ["TestCase12", "TestCase18", "TestCase33"]  # directional (sun) lights: _connect_directional, sun_light_bsdf


def run(*args):
    out = subprocess.run([HARNESS] + [str(a) for a in args], check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


def ref_trace(scene_path, mode, rays):
    os.makedirs(TMP, exist_ok=True)
    rays.tofile(TMP + "/rays.bin")
    info = run("rays", scene_path, mode, TMP + "/rays.bin", TMP + "/hits.bin")
    return np.fromfile(TMP + "/hits.bin", dtype=HIT_DTYPE), info


def camera_rays(flat, width, height, rng, camera=0):
    """Jittered primary rays as Technique::_for_each_ray shoots them (Technique.cpp:321-331) and
    Scene::intersect sets them up (Scene.cpp:186-196), in float32."""
    f32 = np.float32
    ctx = flat.camera_context(camera, width, height)
    ys, xs = np.mgrid[0:height, 0:width]
    px = (xs.ravel() + rng.random(width * height)).astype(f32)
    py = (ys.ravel() + rng.random(width * height)).astype(f32)
    inv = f32(1.0) / f32(height)
    x = px * inv * f32(2.0) - f32(width) * inv
    y = py * inv * f32(2.0) - f32(1.0)
    d = np.stack([x, y, np.full_like(x, -ctx["focal_length_y"])], axis=1)
    d = d / np.linalg.norm(d, axis=1, keepdims=True).astype(f32)
    v2w = ctx["view_to_world"].reshape(3, 3)  # rows = columns of the glm matrix
    d = (d[:, 0:1] * v2w[0] + d[:, 1:2] * v2w[1] + d[:, 2:3] * v2w[2]).astype(f32)
    gn = -v2w[2]
    sign = np.where(d @ gn > 0, f32(1), f32(-1)).astype(f32)
    org = (ctx["position"][None, :] + sign[:, None] * gn[None, :] * f32(0.0001)).astype(f32)
    rays = np.zeros(len(d), dtype=RAY_DTYPE)
    rays["org"], rays["dir"], rays["tnear"], rays["tfar"] = org, d, 0.0, np.inf
    return rays


def geometric_normals(flat, hits):
    tv = flat.triangle_vertices()
    tri = (flat.mesh_index_offset[hits["geom"]] // 3 + hits["prim"]).astype(np.int64)
    v = tv[tri]
    n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0])
    return (n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)).astype(np.float32)


def bounce_rays(flat, rays, hits, rng):
    """Cosine-distributed continuation rays from the hit points, offset 1e-4 along the normal."""
    ok = hits["geom"] != 0xFFFFFFFF
    r, h = rays[ok], hits[ok]
    p = r["org"] + r["dir"] * h["t"][:, None]
    n = geometric_normals(flat, h)
    n = np.where((np.sum(n * r["dir"], axis=1) > 0)[:, None], -n, n)
    u1, u2 = rng.random(len(p)), rng.random(len(p))
    rad, phi = np.sqrt(u1), 2 * np.pi * u2
    a = np.where(np.abs(n[:, 0:1]) > 0.9, [[0, 1, 0]], [[1, 0, 0]])
    t = np.cross(n, a); t /= np.linalg.norm(t, axis=1, keepdims=True)
    b = np.cross(n, t)
    d = t * (rad * np.cos(phi))[:, None] + b * (rad * np.sin(phi))[:, None] + n * np.sqrt(1 - u1)[:, None]
    out = np.zeros(len(p), dtype=RAY_DTYPE)
    out["org"] = (p + n * 1e-4).astype(np.float32)
    out["dir"] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    out["tnear"], out["tfar"] = 0.0, np.inf
    return out


def shadow_rays(flat, rays, hits, rng):
    """Segments hit point -> random point on a light, dir unnormalised, t in (0, 1) (Scene.cpp:151-171)."""
    ok = hits["geom"] != 0xFFFFFFFF
    r, h = rays[ok], hits[ok]
    p = r["org"] + r["dir"] * h["t"][:, None]
    n = geometric_normals(flat, h)
    n = np.where((np.sum(n * r["dir"], axis=1) > 0)[:, None], -n, n)
    L = flat.lights[rng.integers(0, len(flat.lights), len(p))]
    tan = L["tangent"].reshape(-1, 3, 3)
    uv = (rng.random((len(p), 2)) - 0.5) * L["size"]
    q = L["position"] + uv[:, 0:1] * tan[:, 0] + uv[:, 1:2] * tan[:, 2]
    a = p + n * 1e-4
    b = q + np.where((np.sum(tan[:, 1] * (p - q), axis=1) > 0)[:, None], tan[:, 1], -tan[:, 1]) * 1e-4
    out = np.zeros(len(p), dtype=RAY_DTYPE)
    out["org"], out["dir"] = a.astype(np.float32), (b - a).astype(np.float32)
    out["tnear"], out["tfar"] = 0.0, 1.0
    return out


def make_ray_sets(flat, scene_path, width, height, seed):
    """The three ray classes with Embree's answers; returns a dict of arrays."""
    rng = np.random.default_rng(seed)
    cam = camera_rays(flat, width, height, rng)
    cam_hits, _ = ref_trace(scene_path, 0, cam)
    bounce = bounce_rays(flat, cam, cam_hits, rng)
    bounce_hits, _ = ref_trace(scene_path, 0, bounce)
    bounce_mesh_hits, _ = ref_trace(scene_path, 1, bounce)
    shadow = shadow_rays(flat, cam, cam_hits, rng)
    shadow_hits, _ = ref_trace(scene_path, 2, shadow)
    return dict(camera_rays=cam, camera_hits=cam_hits, bounce_rays=bounce, bounce_hits=bounce_hits,
                bounce_mesh_hits=bounce_mesh_hits, shadow_rays=shadow, shadow_hits=shadow_hits)


def ref_render(scene_path, technique, width, height, num_samples, **kw):
    os.makedirs(TMP, exist_ok=True)
    args = ["render", scene_path, "technique=" + technique, "width=%d" % width, "height=%d" % height,
            "num_samples=%d" % num_samples, "out=" + TMP + "/r"] + ["%s=%s" % kv for kv in kw.items()]
    info = run(*args)
    n = width * height
    return dict(info=json.dumps(info),
                paths=(np.fromfile(TMP + "/r.paths.f32", dtype=np.float32).reshape(num_samples, n, 3)
                       if str(kw.get("per_path", 1)) != "0" else np.zeros((0, n, 3), np.float32)),
                accum=np.fromfile(TMP + "/r.accum.f64", dtype=np.float64).reshape(n, 4),
                light=np.fromfile(TMP + "/r.light.f64", dtype=np.float64).reshape(n, 3))


def make_import_records():
    """tests/golden/import/*.b2i: what the Assimp fork hands to loadScene (ref_harness rawdump), the input of tests/test_import.py."""
    import lzma
    out = os.path.join(HERE, "import")
    os.makedirs(out, exist_ok=True)
    for name in ["CornellBoxDiffuse", "CornellBoxSpecular", "TestCaseFurnace", "TestCase12", "MetalRings"]:
        dst = os.path.join(out, name + ".b2i")
        print(name, run("rawdump", os.path.join(REF, "models", name + ".blend"), dst))
        if os.path.getsize(dst) > (1 << 18):
            with open(dst, "rb") as f, lzma.open(dst + ".xz", "wb", preset=9) as g:
                g.write(f.read())
            os.remove(dst)


def main():
    if "import" in sys.argv[1:]:
        return make_import_records()
    os.makedirs(os.path.join(HERE, "scenes"), exist_ok=True)
    for name in ([] if [a for a in sys.argv[1:] if not a.startswith("-")] else SMALL_SCENES):
        dst = os.path.join(HERE, "scenes", name + ".b2s")
        print(name, run("dump", os.path.join(REF, "models", name + ".blend"), dst))
        flat = load_b2s(dst)
        sets = make_ray_sets(flat, dst, 96, 96, seed=20261104)
        np.savez_compressed(os.path.join(HERE, "rays_%s.npz" % name), **sets)
    # per-path radiance / image goldens: small windows so the files stay small
    renders = [
        ("CornellBoxDiffuse", "PT", 48, 48, 4, dict(beta=1)),
        ("CornellBoxDiffuse", "BPT", 48, 48, 2, dict(beta=2)),
        ("CornellBoxSpecular", "PT", 48, 48, 2, dict(beta=1)),
        ("CornellBoxSpecular", "BPT", 32, 32, 2, dict(beta=2)),
        ("MetalRings", "PT", 48, 48, 2, dict(beta=1)),
        ("MetalRings", "BPT", 32, 32, 2, dict(beta=2)),
        ("TestCaseFurnace", "PT", 32, 32, 4, dict(beta=1)),
        ("TestCaseFurnace", "BPT", 32, 32, 2, dict(beta=2)),
        ("TestCase9", "PT", 32, 32, 2, dict(beta=1)),
        # VCM / UPG (UPG.cpp): photons = width * height per pass, one stream per photon path (domain 2)
        ("CornellBoxDiffuse", "VCM", 32, 32, 3, dict(beta=2, radius=0.1)),
        ("CornellBoxSpecular", "VCM", 32, 32, 2, dict(beta=2, radius=0.1)),
        ("MetalRings", "VCM", 32, 32, 2, dict(beta=2, radius=0.05)),
        ("TestCaseFurnace", "VCM", 32, 32, 2, dict(beta=2, radius=0.1)),
        ("TestCase12", "PT", 32, 32, 2, dict(beta=1)),
        ("TestCase12", "BPT", 32, 32, 2, dict(beta=2)),
        ("TestCase18", "BPT", 32, 32, 2, dict(beta=2)),
        ("TestCase33", "BPT", 32, 32, 2, dict(beta=2)),
        ("TestCase18", "VCM", 32, 32, 2, dict(beta=2, radius=0.1)),
        ("TestCase33", "VCM", 32, 32, 2, dict(beta=2, radius=0.05)),
        ("CornellBoxDiffuse", "UPG", 32, 32, 16, dict(beta=2, radius=0.1, per_path=0)),
        ("TestCaseFurnace", "UPG", 24, 24, 16, dict(beta=2, radius=0.1, per_path=0)),
        # merging from the camera side (--from-camera, Options.cpp:59): files are tagged VCMc / UPGc
        ("CornellBoxDiffuse", "VCM", 32, 32, 3, dict(beta=2, radius=0.1, from_light=0)),
        ("CornellBoxSpecular", "VCM", 32, 32, 2, dict(beta=2, radius=0.1, from_light=0)),
        ("MetalRings", "VCM", 32, 32, 2, dict(beta=2, radius=0.05, from_light=0)),
        ("TestCase18", "VCM", 32, 32, 2, dict(beta=2, radius=0.1, from_light=0)),
        ("CornellBoxDiffuse", "UPG", 32, 32, 16, dict(beta=2, radius=0.1, per_path=0, from_light=0)),
        ("TestCaseFurnace", "UPG", 24, 24, 16, dict(beta=2, radius=0.1, per_path=0, from_light=0)),
    ]
    # option branches of the reference that the ABI carries: same scenes, the oracle with other options.
    # (name, technique, width, height, spp, options, variant) -> render_<name>_<tag>_<variant>.npz; the options are stored in the file.
    SKY = "0.5,0.6,0.7,0.1,0.2,0.9"
    variants = [
        ("CornellBoxDiffuse", "BPT", 32, 32, 2, dict(beta=0), "beta0"),        # FixedBeta<0> (Beta.hpp:24-27)
        ("CornellBoxDiffuse", "BPT", 32, 32, 2, dict(beta=1.5), "beta1.5"),    # VariableBeta: pow (Beta.hpp:38-42)
        ("CornellBoxSpecular", "BPT", 32, 32, 2, dict(beta=1), "beta1"),       # FixedBeta<1>
        ("CornellBoxDiffuse", "VCM", 32, 32, 2, dict(beta=0, radius=0.1), "beta0"),
        ("CornellBoxDiffuse", "VCM", 32, 32, 2, dict(beta=1.5, radius=0.1), "beta1.5"),
        ("CornellBoxDiffuse", "VCM", 32, 32, 2, dict(beta=2, radius=0.1, enable_vc=0), "novc"),   # --no-vc (UPG.cpp:67,129,142,156)
        ("CornellBoxDiffuse", "VCM", 32, 32, 2, dict(beta=2, radius=0.1, enable_vm=0), "novm"),   # --no-vm (UPG.cpp:386-399,510)
        ("CornellBoxSpecular", "VCM", 32, 32, 2, dict(beta=2, radius=0.1, enable_vm=0), "novm"),
        ("CornellBoxDiffuse", "VCM", 32, 32, 2, dict(beta=2, radius=0.1, num_photons=2048), "photons2x"),  # num_photons = 2 W H
        ("CornellBoxDiffuse", "PT", 32, 32, 4, dict(beta=1, lights=0.5), "lights0.5"),            # PT.cpp:23-26
        ("CornellBoxDiffuse", "PT", 32, 32, 4, dict(beta=1, max_path=3), "maxpath3"),             # PT.cpp:27-40
        ("CornellBoxDiffuse", "PT", 32, 32, 4, dict(beta=1, max_path=1), "maxpath1"),
        ("CornellBoxDiffuse", "PT", 32, 32, 2, dict(beta=2), "beta2"),                             # pow(x, beta) (PT.cpp:72,112)
        ("MetalRings", "BPT", 32, 32, 2, dict(beta=2, sky=SKY), "sky"),                            # misses see the gradient (BPT.cpp:50, Technique.cpp:87-93)
        ("MetalRings", "VCM", 32, 32, 2, dict(beta=2, radius=0.05, sky=SKY), "sky"),               # UPG.cpp:76
        ("CornellBoxDiffuse", "UPG", 32, 32, 16, dict(beta=2, radius=0.1, per_path=0, enable_vc=0), "novc"),
        ("CornellBoxDiffuse", "UPG", 32, 32, 16, dict(beta=2, radius=0.1, per_path=0, enable_vm=0), "novm"),
        # larger windows: 16 k pixel-samples per case instead of 2-9 k
        ("CornellBoxDiffuse", "PT", 128, 128, 2, dict(beta=1), "128"),
        ("CornellBoxDiffuse", "BPT", 128, 128, 2, dict(beta=2), "128"),
        ("MetalRings", "BPT", 96, 96, 2, dict(beta=2), "96"),
        # equal-spp UPG images for a relative-RMSE bound (a localised error must not hide behind the image mean)
        ("CornellBoxDiffuse", "UPG", 32, 32, 256, dict(beta=2, radius=0.1, per_path=0), "256spp"),
        ("CornellBoxSpecular", "UPG", 32, 32, 256, dict(beta=2, radius=0.1, per_path=0), "256spp"),
        ("TestCaseFurnace", "UPG", 24, 24, 256, dict(beta=2, radius=0.1, per_path=0), "256spp"),
    ]
    tag = lambda r: r[1] + ("c" if r[5].get("from_light", 1) == 0 else "")
    only = [a for a in sys.argv[1:] if not a.startswith("-")]
    renders = [r + ("",) for r in renders] + variants
    if only:
        renders = [r for r in renders if tag(r) in only or r[0] in only or r[6] in only or ("variants" in only and r[6])]
    for name in SUN_SCENES:
        dst = os.path.join(HERE, "scenes", name + ".b2s")
        if not os.path.exists(dst):
            print(name, run("dump", os.path.join(REF, "models", name + ".blend"), dst))
    for name, tech, w, h, spp, kw, variant in renders:
        scene = os.path.join(HERE, "scenes", name + ".b2s")
        out = ref_render(scene, tech, w, h, spp, seed=1234, **kw)
        extra = {k: v for k, v in kw.items() if k in ("radius", "beta", "from_light")}
        fname = "render_%s_%s%s.npz" % (name, tag((name, tech, w, h, spp, kw)), "_" + variant if variant else "")
        np.savez_compressed(os.path.join(HERE, fname), width=w, height=h, seed=1234, spp=spp, options=json.dumps(kw), **extra, **out)
        print(fname, out["info"])


if __name__ == "__main__":
    main()
