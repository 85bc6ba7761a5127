#!/usr/bin/env python
"""Photon-index micro-benchmark (the reference's own is benchmark.cpp:323-657).

  python tools/grid_bench.py [--scene tests/golden/scenes/CornellBoxDiffuse.b2s] [--points 1000000,7000000] [--queries 2000,1000000]

N points uniform on the scene's triangles by area (seed 31415), Q query points drawn the same way (seed 51413) — the
generators of benchmark.cpp:330-331 — radius 0.01.  Reports, per (N, Q):
  * build: wall seconds of b2pt_grid_build (host points in, H2D inside) and the device milliseconds of the build proper;
  * query: ms per query of b2pt_grid_query_device with queries and results resident in HBM (CUDA events, `reps` launches),
    the algorithmic bytes of the cell scan — sum over the 27 cells around a query of n_cell x 16 B — and
    the achieved GB/s against MEASURED_PEAKS.json hbm_gbs;
  * the reference's v3::HashGrid3D on this host (oracle/_ref/ref_harness grid, one thread — it is single-threaded by design,
    HashGrid3D.hpp:237-325), and the author's own figures for 7 M points / 2000 queries (benchmark.cpp:654-657);
  * result check: counts equal to the reference's on the first 2000 queries.
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def points_on_triangles(flat, n, seed):
    tv = flat.triangle_vertices().astype(np.float64)
    area = 0.5 * np.linalg.norm(np.cross(tv[:, 1] - tv[:, 0], tv[:, 2] - tv[:, 0]), axis=1)
    rng = np.random.default_rng(seed)
    tri = rng.choice(len(tv), size=n, p=area / area.sum())
    u, v = rng.random(n), rng.random(n)
    flip = u + v > 1
    u[flip], v[flip] = 1 - u[flip], 1 - v[flip]
    p = tv[tri, 0] + u[:, None] * (tv[tri, 1] - tv[tri, 0]) + v[:, None] * (tv[tri, 2] - tv[tri, 0])
    return p.astype(np.float32)


def scan_bytes(points, queries, radius):
    """algorithmic bytes of the cell scans: sum over queries of the points in the 27 cells around each, x 16 B."""
    inv = np.float32(1.0) / np.float32(radius)
    cell = np.floor(points * inv).astype(np.int64)
    key = (cell[:, 0] & 0x1FFFFF) | ((cell[:, 1] & 0x1FFFFF) << 21) | ((cell[:, 2] & 0x1FFFFF) << 42)
    uk, cnt = np.unique(key, return_counts=True)
    qc = np.floor(queries * inv).astype(np.int64)
    total = 0
    for dx in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dz in (-1, 0, 1):
                k = ((qc[:, 0] + dx) & 0x1FFFFF) | (((qc[:, 1] + dy) & 0x1FFFFF) << 21) | (((qc[:, 2] + dz) & 0x1FFFFF) << 42)
                i = np.searchsorted(uk, k)
                i[i >= len(uk)] = 0
                total += int(cnt[i][uk[i] == k].sum())
    return total * 16, len(uk)


def reference_grid(points, queries, radius, tmp="/tmp/b2pt_gridbench"):
    os.makedirs(tmp, exist_ok=True)
    with open(tmp + "/in.bin", "wb") as f:
        f.write(np.uint32(len(points)).tobytes()); f.write(points.tobytes())
        f.write(np.uint32(len(queries)).tobytes()); f.write(queries.tobytes())
        f.write(np.float32(radius).tobytes())
    out = subprocess.run([HARNESS, "grid", tmp + "/in.bin", tmp + "/out.bin"], check=True, capture_output=True, text=True).stdout
    info = json.loads(out.strip().splitlines()[-1])
    raw = np.fromfile(tmp + "/out.bin", dtype=np.uint32)
    counts, at = [], 0
    for _ in range(len(queries)):
        counts.append(int(raw[at])); at += 1 + int(raw[at])
    return info, np.array(counts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=os.path.join(ROOT, "tests", "golden", "scenes", "CornellBoxDiffuse.b2s"))
    ap.add_argument("--points", default="1000000,7000000")
    ap.add_argument("--queries", default="2000,1000000")
    ap.add_argument("--radius", type=float, default=0.01)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    import torch
    from master_b200 import api
    from master_b200.scene import load_b2s
    flat = load_b2s(a.scene)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    out = {"scene": os.path.basename(a.scene), "radius": a.radius, "hbm_peak_gbs": peak,
           "reference_published": {"source": "benchmark.cpp:654-657 (author's machine, 7 000 000 points, 2000 queries, r = 0.01)",
                                   "v3_HashGrid3D": {"build_s": 5.4795, "ms_per_query": 0.0776}, "v2_HashGrid3D": {"build_s": 4.2575, "ms_per_query": 0.0772},
                                   "v2_KDTree3D": {"build_s": 15.2202, "ms_per_query": 0.0979}}, "runs": []}
    stream = torch.cuda.current_stream().cuda_stream
    for n in [int(x) for x in a.points.split(",")]:
        pts = points_on_triangles(flat, n, 31415)
        t0 = time.perf_counter()
        grid = api.PhotonGrid(pts, a.radius)
        build_wall = time.perf_counter() - t0
        info = grid.info()
        ref_info = None
        for q in [int(x) for x in a.queries.split(",")]:
            qs = points_on_triangles(flat, q, 51413)
            d_q = torch.from_numpy(qs).cuda()
            d_counts = torch.zeros(q, dtype=torch.int32, device="cuda")
            d_idx = torch.zeros(q * 8, dtype=torch.int32, device="cuda")
            for _ in range(2):
                grid.query_device(d_q.data_ptr(), q, 8, d_counts.data_ptr(), d_idx.data_ptr(), stream)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.reps):
                grid.query_device(d_q.data_ptr(), q, 8, d_counts.data_ptr(), d_idx.data_ptr(), stream)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.reps
            nbytes, cells = scan_bytes(pts, qs, a.radius)
            counts = d_counts.cpu().numpy().astype(np.int64)
            row = {"points": n, "queries": q, "non_empty_cells": cells, "build_wall_s": round(build_wall, 4), "build_device_ms": round(info["build_ms"], 3),
                   "query_launch_ms": round(ms, 4), "ms_per_query": ms / q, "found_per_query": float(counts.mean()),
                   "scan_bytes": nbytes, "scan_gbs": nbytes / (ms * 1e-3) * 1e-9, "scan_frac_of_hbm_peak": nbytes / (ms * 1e-3) * 1e-9 / peak}
            if os.path.exists(HARNESS) and q <= 2000:
                ref_info, ref_counts = reference_grid(pts, qs, a.radius)
                row["reference_same_host"] = {"build_s": ref_info["build_s"], "ms_per_query": ref_info["query_s"] / q * 1e3, "threads": 1}
                row["counts_equal_reference"] = bool(np.array_equal(counts, ref_counts))
            out["runs"].append(row)
        grid.close()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
