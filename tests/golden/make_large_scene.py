#!/usr/bin/env python
"""Builds the large-mesh stand-in for BASELINE.json's CrytekSponza config: scenes/_large/LivingRoom.b2s.xz.

models/CrytekSponza.blend is stripped from the reference mount (.MISSING_LARGE_BLOBS:45); the largest
scene that IS there is models/LivingRoom.blend (567 302 triangles, 56 meshes, one camera, NO light —
Technique::render throws on it, AreaLights.cpp:217).  This script makes it renderable the way
SURVEY.md §8(d) config 4 says:

  1. `ref_harness dump --light ...` loads the .blend through the reference's own Assimp-fork loader and
     appends one area light exactly as loader.cpp:434-456 does (AreaLights::addLight + create_mesh +
     LightBSDF with the scene bounding sphere);
  2. the scene was authored for another renderer and its Phong albedos sum to 1.8 ("Invalid material",
     loader.cpp:391-394): every path would gain energy at each bounce.  diffuse and specular are scaled
     by 0.4 so the stand-in converges;
  3. 153 vertices (on a few very large wall triangles) get NaN tangent frames from the loader: their normal
     is parallel to the triangle's first edge, so `normalize(edge - dot(normal, edge) * normal)`
     (loader.cpp:337) divides by zero, and every path that lands there shoots a NaN ray that walks the
     whole BVH (10 ms each in Embree; 90 % of the CPU time before the repair).  Those frames are rebuilt
     around the vertex normal with an arbitrary perpendicular tangent;
  4. a second `ref_harness dump` of the edited .b2s re-creates every BSDF through the reference's
     constructors, so PhongBSDF::_diffuse_probability (BSDF.cpp:306-315) is the reference's own value.

The result (88 MB, 33 MB as .xz) is too big for the git history: scenes/_large/ is git-ignored but NOT
gpurun-ignored, so it travels to the GPU box like the built .so files.  bench.py falls back to the
largest committed scene (MetalRings) and says so when the file is absent.
"""
import lzma
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
sys.path.insert(0, ROOT)
from master_b200.scene import BSDF_DIFFUSE, BSDF_PHONG, load_b2s  # noqa: E402

REF = os.environ.get("REF", "/root/reference")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
OUT_DIR = os.path.join(ROOT, "scenes", "_large")
TMP = os.environ.get("TMPDIR", "/tmp") + "/b2pt_large"

# position, direction (normal), up, exitance, size — a 1.5 m panel under the ceiling (z up, ceiling at 3.17)
LIGHT = [0.4, -4.6, 3.1, 0, 0, -1, 0, 1, 0, 30, 30, 30, 1.5, 1.5]
ALBEDO_SCALE = 0.4


def main():
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(TMP, exist_ok=True)
    a, b = TMP + "/a.b2s", TMP + "/b.b2s"
    subprocess.run([HARNESS, "dump", os.path.join(REF, "models", "LivingRoom.blend"), a, "--light"] +
                   [str(x) for x in LIGHT], check=True, stderr=subprocess.DEVNULL)
    flat = load_b2s(a)
    m = flat.materials
    sel = (m["type"] == BSDF_PHONG) | (m["type"] == BSDF_DIFFUSE)
    m["diffuse"][sel] *= np.float32(ALBEDO_SCALE)
    m["specular"][sel] *= np.float32(ALBEDO_SCALE)
    t = flat.tangents.reshape(-1, 3, 3)  # rows: tangent, normal, bitangent (columns of the glm mat3)
    bad = ~np.isfinite(t).all(axis=(1, 2))
    n = t[bad, 1]
    # 64 of them (zero-area triangles, which no ray can hit) have a NaN normal as well
    n[~np.isfinite(n).all(axis=1)] = [0.0, 0.0, 1.0]
    t[bad, 1] = n
    helper = np.where(np.abs(n[:, 0:1]) < 0.9, [[1.0, 0.0, 0.0]], [[0.0, 1.0, 0.0]]).astype(np.float32)
    tan = helper - np.sum(helper * n, axis=1, keepdims=True) * n
    tan /= np.linalg.norm(tan, axis=1, keepdims=True)
    t[bad, 0] = tan
    t[bad, 2] = np.cross(n, tan)
    print("repaired %d NaN tangent frames" % int(bad.sum()))
    assert np.isfinite(flat.tangents).all()
    flat.save(b)
    final = os.path.join(TMP, "LivingRoom.b2s")
    print(subprocess.run([HARNESS, "dump", b, final], check=True, capture_output=True, text=True).stdout.strip())
    with open(final, "rb") as f:
        raw = f.read()
    with lzma.open(os.path.join(OUT_DIR, "LivingRoom.b2s.xz"), "wb", preset=6) as f:
        f.write(raw)
    print("wrote", os.path.join(OUT_DIR, "LivingRoom.b2s.xz"), os.path.getsize(os.path.join(OUT_DIR, "LivingRoom.b2s.xz")))


if __name__ == "__main__":
    main()
