"""The collective behind the boundary: b2pt_film_* keeps the dvec4 film in HBM, b2pt_comm_* +
b2pt_reduce_accum add the films of the ranks — merge_exr's arithmetic (Options.cpp:1340-1409) as one NCCL reduce.
One GPU is enough for the single-rank path and for a real NCCL communicator of one rank per process; the two-rank test
needs two devices (`gpurun --gpus 2`) and is skipped otherwise."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, golden_path

pytestmark = pytest.mark.gpu


def test_film_lives_on_the_device_and_reads_back(device_scenes, flat_scenes):
    from master_b200 import api
    lib = api.load_library()
    film = C.c_void_p()
    api.check(lib.b2pt_film_create(32, 24, C.byref(film)))
    p = api.make_params(flat_scenes("CornellBoxDiffuse"), technique="PT", width=32, height=24, num_samples=3, seed=9)
    stats = api.Stats()
    api.check(lib.b2pt_render(device_scenes("CornellBoxDiffuse").handle, C.byref(p), lib.b2pt_film_data(film), None, None, C.byref(stats)))
    view = np.ones((24, 32, 4))
    api.check(lib.b2pt_film_read(film, view.ctypes.data_as(C.c_void_p), 1, None))  # view += film
    assert np.all(view[:, :, 3] == 4.0) and view[:, :, :3].mean() > 1.01
    comm = api.Comm(1, 0)  # one rank: the reduce is the identity and needs no NCCL
    comm.reduce_accum(lib.b2pt_film_data(film), 32 * 24 * 4)
    again = np.zeros((24, 32, 4))
    api.check(lib.b2pt_film_read(film, again.ctypes.data_as(C.c_void_p), 0, None))
    np.testing.assert_array_equal(again, view - 1.0)
    api.check(lib.b2pt_film_clear(film, None))
    api.check(lib.b2pt_film_read(film, again.ctypes.data_as(C.c_void_p), 0, None))
    assert not again.any()
    comm.close()
    lib.b2pt_film_destroy(film)


WORKER = r"""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, %(root)r)
import torch
from master_b200 import api
rank, world = int(sys.argv[1]), int(sys.argv[2])
lib = api.load_library()
api.check(lib.b2pt_set_device(rank))
torch.cuda.set_device(rank)
ident = open(sys.argv[3], "rb").read()
comm = api.Comm(world, rank, ident)
film = torch.full((64, 4), float(rank + 1), dtype=torch.float64, device="cuda")
comm.reduce_accum(film.data_ptr(), film.numel(), root=-1)
torch.cuda.synchronize()
assert torch.all(film == sum(range(1, world + 1))), film[0]
film2 = torch.full((64, 4), float(rank + 1), dtype=torch.float64, device="cuda")
comm.reduce_accum(film2.data_ptr(), film2.numel(), root=0)
torch.cuda.synchronize()
if rank == 0:
    assert torch.all(film2 == sum(range(1, world + 1)))
comm.close()
print("ok", rank)
"""


def test_two_rank_reduce_over_nccl(tmp_path):
    import torch
    from master_b200 import api
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices (gpurun --gpus 2)")
    ident = tmp_path / "id.bin"
    ident.write_bytes(api.Comm.unique_id())
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "2", str(ident)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(2)]
    for p in procs:
        out, err = p.communicate(timeout=300)
        assert p.returncode == 0 and "ok" in out, err[-2000:]


def test_b2pt_master_two_gpus_equals_one(tmp_path):
    """`b2pt_master --gpus 2` (one host thread per device, samples split 0, 2, 4 / 1, 3, 5, one b2pt_reduce_accum) equals the
    single-GPU film of the same six passes up to the order of the atomics."""
    import torch
    host = os.path.join(ROOT, "master_b200", "host", "_build", "b2pt_master")
    if torch.cuda.device_count() < 2 or not os.path.exists(host):
        pytest.skip("needs two CUDA devices (gpurun --gpus 2) and the host binary")
    scene = golden_path("scenes", "CornellBoxDiffuse.b2s")
    a, b = str(tmp_path / "a.f64"), str(tmp_path / "b.f64")
    base = [host, scene, "--BPT", "--beta=2", "--resolution=64x48", "--num-samples=6", "--seed=5"]
    for args in (base + ["out=" + a], base + ["--gpus=2", "out=" + b]):
        r = subprocess.run(args, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    fa, fb = np.fromfile(a, dtype=np.float64).reshape(-1, 4), np.fromfile(b, dtype=np.float64).reshape(-1, 4)
    assert np.all(fa[:, 3] == 6) and np.all(fb[:, 3] == 6)
    np.testing.assert_allclose(fb, fa, rtol=1e-6, atol=1e-9)
