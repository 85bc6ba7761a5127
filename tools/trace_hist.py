#!/usr/bin/env python
"""Histogram of node / triangle tests per ray (diagnostic kernel, B2PT_TRACE_VARIANT=2). Needs a GPU."""
import json, os, sys
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_golden as mg
from master_b200 import api
from master_b200.scene import load_b2s
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "scenes", "_large", "LivingRoom.b2s.xz")
res = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
flat = load_b2s(path); scene = api.DeviceSceneHandle(flat)
rng = np.random.default_rng(12345)
rays = mg.camera_rays(flat, res, res, rng)
out = {"info": scene.info()}
for name in ("camera", "bounce1", "bounce2"):
    os.environ["B2PT_TRACE_VARIANT"] = "1"
    hits = scene.trace(rays, api.TRACE_CLOSEST_ALL)
    os.environ["B2PT_TRACE_VARIANT"] = "2"
    c = scene.trace(rays, api.TRACE_CLOSEST_ALL)
    nn, nt = c["u"], c["v"]
    q = [50, 90, 99, 99.9, 99.99, 100]
    out[name] = {"rays": len(rays), "nodes_mean": float(nn.mean()), "tris_mean": float(nt.mean()),
                 "nodes_pct": dict(zip(map(str, q), np.percentile(nn, q).tolist())),
                 "tris_pct": dict(zip(map(str, q), np.percentile(nt, q).tolist()))}
    rays = mg.bounce_rays(flat, rays, hits, rng)
print(json.dumps(out, indent=1))
