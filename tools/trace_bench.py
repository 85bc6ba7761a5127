#!/usr/bin/env python
"""Mrays/s micro-benchmark of the traversal kernels through b2pt_trace_device, with the tree's quality beside it (needs a GPU).

  python tools/trace_bench.py [--scene scenes/_large/LivingRoom.b2s.xz] [--res 1024] [--reps 5]
                              [--builds "default:;nosplit:B2PT_SPLIT_BUDGET=0;r32:B2PT_PLOC_RADIUS=32"] [--variants 0,2]

Ray classes: jittered camera rays, their cosine-sampled bounce rays (1st, 2nd, 3rd bounce: increasingly
incoherent), and shadow segments towards the lights.  For every BUILD (a set of environment variables the builder reads when a
scene is created: B2PT_SPLIT_BUDGET, B2PT_PLOC_RADIUS, B2PT_LEAF_MAX, B2PT_BUILDER) the tool reports
  * the tree: nodes, triangle references, SAH cost and the expected node visits / triangle tests of a long random ray
    (b2pt_debug_tree_stats, from the quantised child boxes the traversal really tests);
  * per ray class: node visits and triangle tests per ray as the one-ray-per-thread traversal counts them
    (b2pt_debug_trace_counts), Grays/s of each traversal variant (B2PT_TRACE_STREAM: 0 = one ray per thread, 2 = persistent warps
    with dynamic fetch, the integrators' default) timed with CUDA events around `reps` launches on the current stream with rays
    and hits resident in HBM, and the agreement of the hits with the first build's (ids, and t bit for bit).
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

BUILD_ENV = ("B2PT_SPLIT_BUDGET", "B2PT_PLOC_RADIUS", "B2PT_LEAF_MAX", "B2PT_BUILDER", "B2PT_L2_PERSIST", "B2PT_DYN_FETCH", "B2PT_DYN_TRI")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=os.path.join(ROOT, "scenes", "_large", "LivingRoom.b2s.xz"))
    ap.add_argument("--res", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--variants", default="0,2")
    ap.add_argument("--builds", default="default:")
    ap.add_argument("--bounces", type=int, default=3)
    a = ap.parse_args()
    import torch
    import make_golden as mg
    from master_b200 import api
    from master_b200.scene import load_b2s

    flat = load_b2s(a.scene)
    out = {"scene": os.path.basename(a.scene), "triangles": int(flat.num_triangles), "builds": {}}
    stream = torch.cuda.current_stream().cuda_stream
    sets, dev, ref_hits = None, {}, {}
    for spec in a.builds.split(";"):
        label, _, envs = spec.partition(":")
        for k in BUILD_ENV:
            os.environ.pop(k, None)
        for kv in filter(None, envs.split(",")):
            k, v = kv.split("=")
            os.environ[k] = v
        sc = api.DeviceSceneHandle(flat)
        if sets is None:  # the ray sets are made once, with the first build (hits decide where bounce rays start)
            rng = np.random.default_rng(12345)
            sets = {}
            rays = mg.camera_rays(flat, a.res, a.res, rng)
            sets["camera"] = (rays, api.TRACE_CLOSEST_ALL)
            prev = rays
            for b in range(1, a.bounces + 1):
                hits = sc.trace(prev, api.TRACE_CLOSEST_ALL)
                if b == 1:
                    sets["shadow"] = (mg.shadow_rays(flat, prev, hits, rng), api.TRACE_OCCLUDED_MESH)
                nxt = mg.bounce_rays(flat, prev, hits, rng)
                sets["bounce%d" % b] = (nxt, api.TRACE_CLOSEST_ALL)
                prev = nxt
            for name, (r, mode) in sets.items():
                n = len(r)
                dev[name] = (torch.from_numpy(r.view(np.float32).reshape(n, 8)).cuda(), torch.zeros(n, 5, dtype=torch.int32, device="cuda"))
        row = {"env": envs, "info": sc.info(), "tree": sc.tree_stats(), "classes": {}}
        for name, (r, mode) in sets.items():
            n = len(r)
            counts = sc.trace_counts(r[: min(n, 1 << 18)], mode).astype(np.float64)
            c = {"rays": n, "node_visits_per_ray": round(float(counts[:, 0].mean()), 3), "triangle_tests_per_ray": round(float(counts[:, 1].mean()), 3)}
            d_rays, d_hits = dev[name]
            for v in a.variants.split(","):
                os.environ["B2PT_TRACE_STREAM"] = v
                for _ in range(2):
                    sc.trace_device(d_rays.data_ptr(), n, mode, d_hits.data_ptr(), stream)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(a.reps):
                    sc.trace_device(d_rays.data_ptr(), n, mode, d_hits.data_ptr(), stream)
                e1.record()
                torch.cuda.synchronize()
                c["v%s_grays_per_s" % v] = round(n / (e0.elapsed_time(e1) / a.reps) * 1e-6, 4)
            os.environ.pop("B2PT_TRACE_STREAM", None)
            h = d_hits.cpu().numpy().view(api.HIT_DTYPE).reshape(n)
            if name not in ref_hits:
                ref_hits[name] = h.copy()
            else:
                g = ref_hits[name]
                if mode == api.TRACE_OCCLUDED_MESH:
                    c["agree"] = float((h["geom"] == g["geom"]).mean())
                else:
                    same = (h["geom"] == g["geom"]) & (h["prim"] == g["prim"])
                    c["agree"] = float(same.mean())
                    c["t_bitexact"] = float((h["t"][same].view(np.uint32) == g["t"][same].view(np.uint32)).mean())
            row["classes"][name] = c
        out["builds"][label] = row
        sc.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
