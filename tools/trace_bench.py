#!/usr/bin/env python
"""Mrays/s micro-benchmark of the traversal kernels through b2pt_trace_device (needs a GPU).

  python tools/trace_bench.py [--scene scenes/_large/LivingRoom.b2s.xz] [--res 1024] [--reps 5] [--variants 0,1]

Ray classes: jittered camera rays, their cosine-sampled bounce rays (1st, 2nd, 3rd bounce:
increasingly incoherent), and shadow segments towards the lights.  Each class is timed with CUDA events around
`reps` launches on the current stream, rays and hits resident in HBM.  With several --variants (B2PT_TRACE_STREAM: 0 = one ray per thread, the default;
1 = persistent warp-synchronous traversal) the hits are also compared between variants.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=os.path.join(ROOT, "scenes", "_large", "LivingRoom.b2s.xz"))
    ap.add_argument("--res", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--variants", default="0")
    ap.add_argument("--bounces", type=int, default=3)
    a = ap.parse_args()
    import torch
    import make_golden as mg
    from master_b200 import api
    from master_b200.scene import load_b2s

    flat = load_b2s(a.scene)
    scene = api.DeviceSceneHandle(flat)
    rng = np.random.default_rng(12345)
    sets = {}
    rays = mg.camera_rays(flat, a.res, a.res, rng)
    sets["camera"] = (rays, api.TRACE_CLOSEST_ALL)
    prev = rays
    for b in range(1, a.bounces + 1):
        hits = scene.trace(prev, api.TRACE_CLOSEST_ALL)
        if b == 1:
            sets["shadow"] = (mg.shadow_rays(flat, prev, hits, rng), api.TRACE_OCCLUDED_MESH)
        nxt = mg.bounce_rays(flat, prev, hits, rng)
        sets["bounce%d" % b] = (nxt, api.TRACE_CLOSEST_ALL)
        prev = nxt
    out = {"scene": os.path.basename(a.scene), "info": scene.info(), "classes": {}}
    stream = torch.cuda.current_stream().cuda_stream
    ref_hits = {}
    dev = {}
    for name, (r, mode) in sets.items():
        n = len(r)
        dev[name] = (torch.from_numpy(r.view(np.float32).reshape(n, 8)).cuda(), torch.zeros(n, 5, dtype=torch.int32, device="cuda"))
        out["classes"][name] = {"rays": n}
    # a variant is "mode" or "mode:fetch:tri" (B2PT_TRACE_STREAM, and for mode 2 the trace_dyn8 thresholds, which a scene
    # reads when it is created)
    for v in a.variants.split(","):
        parts = v.split(":")
        os.environ["B2PT_TRACE_STREAM"] = parts[0]
        if len(parts) == 3:
            os.environ["B2PT_DYN_FETCH"], os.environ["B2PT_DYN_TRI"] = parts[1], parts[2]
        sc = api.DeviceSceneHandle(flat)
        for name, (r, mode) in sets.items():
            n = len(r)
            d_rays, d_hits = dev[name]
            row = out["classes"][name]
            for _ in range(2):
                sc.trace_device(d_rays.data_ptr(), n, mode, d_hits.data_ptr(), stream)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.reps):
                sc.trace_device(d_rays.data_ptr(), n, mode, d_hits.data_ptr(), stream)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.reps
            row["v%s_grays_per_s" % v] = round(n / ms * 1e-6, 4)
            h = d_hits.cpu().numpy().view(api.HIT_DTYPE).reshape(n)
            if name not in ref_hits:
                ref_hits[name] = h.copy()
            else:
                g = ref_hits[name]
                if mode == api.TRACE_OCCLUDED_MESH:
                    row["v%s_agree" % v] = float((h["geom"] == g["geom"]).mean())
                else:
                    same = (h["geom"] == g["geom"]) & (h["prim"] == g["prim"])
                    row["v%s_agree" % v] = float(same.mean())
                    row["v%s_t_bitexact" % v] = float((h["t"][same].view(np.uint32) == g["t"][same].view(np.uint32)).mean())
        sc.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
