"""The photon-index oracle is pinned: on the reference's own golden vectors (test_data/test_case_*.case, repacked by
tests/golden/make_grid_golden.py) the reference's v3::HashGrid3D::rQuery — compiled unmodified into the oracle harness —
returns exactly the result sets stored in the .case files (made by the reference's KD-tree solver, benchmark.cpp:323-337)
and exactly the brute-force sets {i : distance2(q, p_i) < r^2} in float32."""
import numpy as np
import pytest

from conftest import golden_path

CASES = ["2", "A", "B", "C", "D", "E"]


def split(counts, flat):
    out, at = [], 0
    for c in counts:
        out.append(flat[at:at + int(c)]); at += int(c)
    return out


def brute(points, q, radius):
    d = points - q[None, :]                       # glm::distance2(query, point) = length2(point - query), float32
    d2 = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]
    return np.nonzero(d2 < np.float32(radius) * np.float32(radius))[0]


@pytest.mark.parametrize("name", CASES)
def test_reference_grid_matches_its_golden_vectors(name):
    g = np.load(golden_path("grid_cases.npz"))
    pts, qs, r = g[name + "_points"], g[name + "_queries"], float(g[name + "_radius"])
    grid_lists = split(g[name + "_grid_counts"], g[name + "_grid_indices"])
    kd_lists = split(g[name + "_kd_counts"], g[name + "_kd_positions"])
    assert len(grid_lists) == len(qs) == len(kd_lists)
    for q, idx, kd in zip(qs, grid_lists, kd_lists):
        assert len(idx) == len(kd)
        assert sorted(map(tuple, pts[idx].tolist())) == sorted(map(tuple, kd.tolist()))
        assert sorted(idx.tolist()) == brute(pts, q, r).tolist()
