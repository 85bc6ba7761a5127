"""b2pt_film_errors_static (include/b2pt.h): the allocation-free variant of b2pt_film_errors, through the C ABI, against the
existing entry point.  The file is named to be collected last.

The scratch of the variant is ONE __device__ array shared by every call, which keeps the last call's contents: the second call
of each test is smaller than the first and must not see the first one's partial sums."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL = 1e-6  # the bound tests/test_errors_gpu.py holds the device's double sums to


def make(width, height, seed):
    rng = np.random.default_rng(seed)
    film = np.empty((height, width, 4))
    n = rng.integers(1, 64, size=(height, width)).astype(np.float64)
    film[:, :, :3] = rng.random((height, width, 3)) * 4.0 * n[:, :, None]
    film[:, :, 3] = n
    ref = (rng.random((height, width, 3)) * 4.0).astype(np.float32)
    return film, ref


def both(width, height, seed, window=None):
    """(rms, abs) of the existing entry point and of the static one on the same device buffers."""
    import torch
    from master_b200 import api
    lib = api.load_library()
    film, ref = make(width, height, seed)
    f, r = torch.from_numpy(film).cuda(), torch.from_numpy(ref).cuda()
    out_dev = torch.full((2,), -1.0, dtype=torch.float32, device="cuda")  # the caller's output buffer
    win = (C.c_int32 * 3)(*window) if window is not None else None
    want = (C.c_float * 2)()
    api.check(lib.b2pt_film_errors(C.c_void_p(f.data_ptr()), C.c_void_p(r.data_ptr()), width, height, win, want, None))
    stream = torch.cuda.current_stream().cuda_stream
    api.check(lib.b2pt_film_errors_static(C.c_void_p(f.data_ptr()), C.c_void_p(r.data_ptr()), width, height, win,
                                          C.c_void_p(out_dev.data_ptr()), C.c_void_p(stream)))
    torch.cuda.synchronize()
    got = out_dev.cpu().numpy()
    return (float(want[0]), float(want[1])), (float(got[0]), float(got[1]))


def test_two_calls_in_a_row_match_the_existing_entry_point():
    # 1024^2 fills 256 tiles of the scratch array, 37 x 21 one: stale partial sums of the first call would show in the second
    for width, height, seed in [(1024, 1024, 5), (37, 21, 6)]:
        want, got = both(width, height, seed)
        assert want[0] > 0.0 and want[1] > 0.0
        assert got[0] == pytest.approx(want[0], rel=REL) and got[1] == pytest.approx(want[1], rel=REL)


def test_windows_and_repeat_is_deterministic():
    first = None
    for window in [(512, 100, 16), (0, 0, 3), (300, 300, 200), (512, 100, 16)]:
        want, got = both(640, 480, 9, window)
        assert got[0] == pytest.approx(want[0], rel=REL) and got[1] == pytest.approx(want[1], rel=REL)
        if window == (512, 100, 16):
            if first is None:
                first = got
            else:
                assert got == first  # fixed summation order: bit-identical on the same input


def test_no_allocation_inside_the_call():
    """the driver's memory accounting does not move across calls (cudaMalloc would show as less free memory)."""
    import torch
    from master_b200 import api
    lib = api.load_library()
    film, ref = make(512, 512, 3)
    f, r = torch.from_numpy(film).cuda(), torch.from_numpy(ref).cuda()
    out_dev = torch.zeros(2, dtype=torch.float32, device="cuda")
    args = (C.c_void_p(f.data_ptr()), C.c_void_p(r.data_ptr()), 512, 512, None, C.c_void_p(out_dev.data_ptr()), None)
    api.check(lib.b2pt_film_errors_static(*args))  # first call on the device resolves the symbol (module already loaded)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(8):
        api.check(lib.b2pt_film_errors_static(*args))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0


def test_beyond_the_static_maximum_is_rejected():
    import torch
    from master_b200 import api
    lib = api.load_library()
    dummy = torch.zeros(16, dtype=torch.float64, device="cuda")
    out_dev = torch.full((2,), -1.0, dtype=torch.float32, device="cuda")
    # 4097 x 4096 > 2^24 pixels: refused before anything is launched or read
    rc = lib.b2pt_film_errors_static(C.c_void_p(dummy.data_ptr()), C.c_void_p(dummy.data_ptr()), 4097, 4096, None,
                                     C.c_void_p(out_dev.data_ptr()), None)
    assert rc == 4  # B2PT_ERR_CAPACITY
    assert b"at most 16777216" in lib.b2pt_last_error()
    torch.cuda.synchronize()
    assert out_dev.cpu().tolist() == [-1.0, -1.0]
    assert lib.b2pt_film_errors_static(None, None, 64, 48, None, None, None) == 1  # null arguments: B2PT_ERR_INVALID
