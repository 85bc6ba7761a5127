#!/usr/bin/env python
"""Photon-index fixtures: the reference's own rQuery golden vectors + what its v3::HashGrid3D returns on them.

/root/reference/test_data/test_case_{2,A,B,C,D,E}.case (format: benchmark.cpp:243-304 — points, queries, per-query
result POSITIONS produced by the reference's KD-tree solver, radius) are repacked into tests/golden/grid_cases.npz
together with the index lists `oracle/_ref/ref_harness grid` (= v3::HashGrid3D<T>::build + rQuery, HashGrid3D.hpp:181-325,
the reference's own header compiled unmodified) returns for the same points and queries, in callback order.
The files are 186 KB in total; nothing here is reference SOURCE.
"""
import json
import os
import struct
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
REF = os.environ.get("REF", "/root/reference")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
TMP = os.environ.get("TMPDIR", "/tmp") + "/b2pt_grid"


def read_case(path):
    buf = open(path, "rb").read()
    off = 0

    def vec3s():
        nonlocal off
        n, = struct.unpack_from("<I", buf, off); off += 4
        a = np.frombuffer(buf, "<f4", n * 3, off).reshape(n, 3).copy(); off += n * 12
        return a
    points, queries = vec3s(), vec3s()
    nres, = struct.unpack_from("<I", buf, off); off += 4
    results = [vec3s() for _ in range(nres)]
    radius, = struct.unpack_from("<f", buf, off); off += 4
    assert off == len(buf), path
    return points, queries, results, radius


def harness_grid(points, queries, radius):
    os.makedirs(TMP, exist_ok=True)
    with open(TMP + "/in.bin", "wb") as f:
        f.write(struct.pack("<I", len(points))); f.write(np.ascontiguousarray(points, "<f4").tobytes())
        f.write(struct.pack("<I", len(queries))); f.write(np.ascontiguousarray(queries, "<f4").tobytes())
        f.write(struct.pack("<f", radius))
    out = subprocess.run([HARNESS, "grid", TMP + "/in.bin", TMP + "/out.bin"], check=True, capture_output=True, text=True).stdout
    info = json.loads(out.strip().splitlines()[-1])
    raw = np.fromfile(TMP + "/out.bin", dtype="<u4")
    lists, at = [], 0
    for _ in range(len(queries)):
        c = int(raw[at]); lists.append(raw[at + 1:at + 1 + c].copy()); at += 1 + c
    return lists, info


def main():
    out = {}
    for name in ["2", "A", "B", "C", "D", "E"]:
        points, queries, results, radius = read_case(os.path.join(REF, "test_data", "test_case_%s.case" % name))
        lists, info = harness_grid(points, queries, radius)
        out["%s_points" % name] = points
        out["%s_queries" % name] = queries
        out["%s_radius" % name] = np.float32(radius)
        out["%s_kd_counts" % name] = np.array([len(r) for r in results], dtype=np.uint32)
        out["%s_kd_positions" % name] = np.concatenate(results) if results and sum(map(len, results)) else np.zeros((0, 3), "<f4")
        out["%s_grid_counts" % name] = np.array([len(l) for l in lists], dtype=np.uint32)
        out["%s_grid_indices" % name] = np.concatenate(lists).astype(np.uint32) if sum(map(len, lists)) else np.zeros(0, "<u4")
        print(name, info, "kd found", int(out["%s_kd_counts" % name].sum()))
    np.savez_compressed(os.path.join(HERE, "grid_cases.npz"), **out)


if __name__ == "__main__":
    main()
