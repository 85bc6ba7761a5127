"""World-size-2 (and 4) gloo tests of the multi-GPU host logic: the sample split is a partition, and the single
film reduce reproduces `master merge` (sum of the dvec4 accumulators, count channel included)."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import golden_path


def _worker(rank, world, port, passes, steps, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from master_b200.dist import reduce_film, samples_of
    g = np.load(golden_path("render_CornellBoxDiffuse_PT.npz"))
    per_pass = g["paths"].astype(np.float64)  # [spp][pixels][3]: the reference's per-pass images, our stand-in for a render
    spp, pixels = per_pass.shape[:2]
    film = torch.zeros(pixels, 4, dtype=torch.float64)
    mine = []
    for step in range(steps):
        for s in samples_of(rank, world, step, passes):
            mine.append(s)
            film[:, :3] += torch.from_numpy(per_pass[s % spp])
            film[:, 3] += 1
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    reduce_film(film)
    if rank == 0:
        torch.save({"samples": gathered, "film": film}, out)
    dist.destroy_process_group()


@pytest.mark.parametrize("world,passes,steps", [(2, 2, 2), (4, 1, 2)])
def test_sample_split_and_film_reduce(world, passes, steps, tmp_path):
    out = str(tmp_path / "r.pt")
    port = 29600 + world * 7 + passes
    mp.spawn(_worker, args=(world, port, passes, steps, out), nprocs=world, join=True)
    r = torch.load(out)
    flat = sorted(s for rank in r["samples"] for s in rank)
    assert flat == list(range(world * passes * steps))  # a partition: every sample exactly once
    g = np.load(golden_path("render_CornellBoxDiffuse_PT.npz"))
    per_pass = g["paths"].astype(np.float64)
    spp = per_pass.shape[0]
    want = sum(per_pass[s % spp] for s in flat)
    np.testing.assert_allclose(r["film"][:, :3].numpy(), want, rtol=1e-12)
    assert torch.all(r["film"][:, 3] == world * passes * steps)


def test_sample_split_rejects_bad_rank():
    from master_b200.dist import sample_split
    with pytest.raises(ValueError):
        sample_split(2, 2, 0, 1)
    assert sample_split(1, 8, 3, 4) == (3 * 4 * 8 + 1, 8)


def test_technique_sample_indices_are_disjoint_across_ranks_and_calls():
    """GpuTechnique(first_sample=g, sample_stride=G): consecutive calls on G ranks cover every global sample exactly once."""
    from master_b200.technique import GpuTechnique
    world, calls, passes = 4, 5, 3
    seen = []
    for rank in range(world):
        t = GpuTechnique(None, "PT", first_sample=rank, sample_stride=world)
        for _ in range(calls):
            first, stride = t.sample_indices(passes)
            assert stride == world
            seen += [first + j * stride for j in range(passes)]
            t.statistics["num_samples"] += passes  # what _account does after a render
    assert sorted(seen) == list(range(world * calls * passes))
    t = GpuTechnique(None, "BPT")  # the single-GPU default: pass k of the run is sample k
    assert t.sample_indices(2) == (0, 1)
    t.statistics["num_samples"] = 7
    assert t.sample_indices(1) == (7, 1)
