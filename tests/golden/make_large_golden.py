#!/usr/bin/env python
"""Regenerates the per-path goldens of the LARGE scene (the bench workload) from the UNMODIFIED reference:

  render_LivingRoom_{PT,BPT,VCM}.npz          oracle/_ref/ref_harness          (glibc's float libm)
  render_LivingRoom_{PT,BPT,VCM}_crlibm.npz   oracle/_ref/ref_harness_crlibm   (the same objects on a correctly rounded float libm,
                                                                                oracle/cr_libm.cpp — DESIGN.md §5 "the host's libm")

Needs scenes/_large/LivingRoom.b2s.xz (tests/golden/make_large_scene.py) and the two harness binaries (oracle/build_ref.sh).
128 x 128 window of camera 0, 2 passes, seed 1234 — 32 768 pixel-samples per file; the same npz layout as make_golden.py writes.

  python tests/golden/make_large_golden.py [PT BPT VCM] [--check]      --check: compare with the committed files instead of writing
"""
import json
import lzma
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
LARGE = os.path.join(ROOT, "scenes", "_large", "LivingRoom.b2s.xz")
TMP = os.environ.get("TMPDIR", "/tmp") + "/b2pt_large_golden"
W = H = 128
SPP = 2
SEED = 1234
CASES = {  # technique -> options of the reference (Options.cpp): PT beta 1 (BASELINE config 1), BPT / VCM beta 2; VCM r0 = 0.025 like bench.py
    "PT": dict(beta=1),
    "BPT": dict(beta=2),
    "VCM": dict(beta=2, radius=0.025),
}


def render(harness, scene, tech, kw):
    args = [harness, "render", scene, "technique=" + tech, "width=%d" % W, "height=%d" % H, "num_samples=%d" % SPP, "out=" + TMP + "/r",
            "seed=%d" % SEED] + ["%s=%s" % kv for kv in kw.items()]
    out = subprocess.run(args, check=True, capture_output=True, text=True).stdout
    info = json.loads(out.strip().splitlines()[-1])
    n = W * H
    return dict(info=json.dumps(info),
                paths=np.fromfile(TMP + "/r.paths.f32", dtype=np.float32).reshape(SPP, n, 3),
                accum=np.fromfile(TMP + "/r.accum.f64", dtype=np.float64).reshape(n, 4),
                light=np.fromfile(TMP + "/r.light.f64", dtype=np.float64).reshape(n, 3))


def main():
    check = "--check" in sys.argv
    only = [a for a in sys.argv[1:] if not a.startswith("-")] or list(CASES)
    os.makedirs(TMP, exist_ok=True)
    scene = os.path.join(TMP, "LivingRoom.b2s")
    with lzma.open(LARGE, "rb") as f, open(scene, "wb") as g:
        g.write(f.read())
    for tech in only:
        kw = CASES[tech]
        for suffix, binary in (("", "ref_harness"), ("_crlibm", "ref_harness_crlibm")):
            out = render(os.path.join(ROOT, "oracle", "_ref", binary), scene, tech, kw)
            fname = os.path.join(HERE, "render_LivingRoom_%s%s.npz" % (tech, suffix))
            extra = {k: v for k, v in kw.items() if k in ("radius", "beta")}
            if check:
                old = np.load(fname)
                same = {k: bool(np.array_equal(old[k], out[k])) for k in ("paths", "accum", "light")}
                print(os.path.basename(fname), "identical to the committed file:", same)
                if not all(same.values()):
                    sys.exit(1)
            else:
                np.savez_compressed(fname, width=W, height=H, seed=SEED, spp=SPP, options=json.dumps(kw), **extra, **out)
                print(os.path.basename(fname), out["info"])


if __name__ == "__main__":
    main()
