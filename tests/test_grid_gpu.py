"""b2pt_grid_build / b2pt_grid_query against the reference's v3::HashGrid3D results (tests/golden/grid_cases.npz),
plus size-independent properties at photon-map scale (1 M points): exact agreement with brute force on sampled
queries, empty grid, points exactly on cell borders, query radius strictness."""
import numpy as np
import pytest

from conftest import golden_path
from test_grid_oracle import CASES, brute, split

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", CASES)
def test_grid_matches_reference(name):
    from master_b200 import api
    g = np.load(golden_path("grid_cases.npz"))
    pts, qs, r = g[name + "_points"], g[name + "_queries"], float(g[name + "_radius"])
    ref = split(g[name + "_grid_counts"], g[name + "_grid_indices"])
    grid = api.PhotonGrid(pts, r)
    counts, idx = grid.query(qs, max_per_query=64)
    assert counts.tolist() == [len(x) for x in ref]
    for c, row, want in zip(counts, idx, ref):
        assert sorted(row[:c].tolist()) == sorted(want.tolist())
    grid.close()


def test_grid_one_million_points_vs_brute_force():
    from master_b200 import api
    rng = np.random.default_rng(31415)
    pts = rng.random((1_000_000, 3), dtype=np.float32) * np.float32(2.0) - np.float32(1.0)
    pts[:, 1] = np.float32(0.25) * pts[:, 1]          # flattened cloud: denser cells
    r = 0.01
    grid = api.PhotonGrid(pts, r)
    qs = pts[rng.integers(0, len(pts), 4000)] + rng.normal(0, r / 2, (4000, 3)).astype(np.float32)
    counts, idx = grid.query(qs, max_per_query=64)
    assert counts.max() <= 64
    for q, c, row in zip(qs[:400], counts[:400], idx[:400]):
        assert sorted(row[:c].tolist()) == brute(pts, q, r).tolist()
    # every query that IS a stored point finds at least itself
    c2, _ = grid.query(pts[:1000], max_per_query=0)
    assert (c2 >= 1).all()
    grid.close()


def test_grid_edge_cases():
    from master_b200 import api
    with pytest.raises(api.B2ptError):
        api.PhotonGrid(np.zeros((3, 3), np.float32), 0.0)
    empty = api.PhotonGrid(np.zeros((0, 3), np.float32), 0.1)
    c, _ = empty.query(np.zeros((5, 3), np.float32))
    assert c.tolist() == [0] * 5
    # strictness: a point at distance exactly r is NOT reported (HashGrid3D.hpp:196 uses <), cell borders and negatives work
    pts = np.array([[0.5, 0, 0], [-0.5, 0, 0], [0, 0.25, 0], [-1e-7, -1e-7, -1e-7], [0.5, 0.5, 0.5]], np.float32)
    grid = api.PhotonGrid(pts, 0.5)
    c, idx = grid.query(np.array([[0, 0, 0], [0.5, 0.5, 0.5], [10, 10, 10]], np.float32), max_per_query=8)
    assert sorted(idx[0, :c[0]].tolist()) == [2, 3]
    assert sorted(idx[1, :c[1]].tolist()) == [4]
    assert c[2] == 0
    # more hits than max_per_query: the count is still the full count
    many = api.PhotonGrid(np.zeros((100, 3), np.float32), 0.1)
    c, idx = many.query(np.zeros((1, 3), np.float32), max_per_query=8)
    assert c[0] == 100
