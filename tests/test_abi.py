"""The C-ABI library loads and exports every symbol include/b2pt.h declares; struct layouts on the Python side
match the header; without a CUDA device every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_path

HEADER = os.path.join(ROOT, "include", "b2pt.h")
DEBUG_HEADER = os.path.join(ROOT, "include", "b2pt_debug.h")


@pytest.fixture(scope="module")
def lib():
    from master_b200 import api, build
    build.build_library()
    return api.load_library()


def declared_functions(header=HEADER):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2pt_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    from master_b200 import api
    names = declared_functions()
    assert sorted(api.EXPORTS) == names
    for n in names:
        assert getattr(lib, n) is not None
    assert lib.b2pt_abi_version() == 3
    # the parity probes are a separate header: same rule, every declared symbol is exported
    debug = declared_functions(DEBUG_HEADER)
    assert sorted(api.DEBUG_EXPORTS) == debug
    for n in debug:
        assert getattr(lib, n) is not None
    assert sorted(os.listdir(os.path.join(ROOT, "include"))) == ["b2pt.h", "b2pt_debug.h"]


def test_struct_sizes_match_the_header(tmp_path):
    """sizeof() as a C compiler sees the header vs the ctypes / numpy mirrors."""
    from master_b200 import api, scene
    prog = tmp_path / "sizes.c"
    prog.write_text('#include <stdio.h>\n#include "b2pt.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                    'sizeof(B2ptMaterial),sizeof(B2ptLight),sizeof(B2ptSceneDesc),sizeof(B2ptCamera),sizeof(B2ptRenderParams),'
                    'sizeof(B2ptStats),sizeof(B2ptRay),sizeof(B2ptHit));return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [scene.MATERIAL_DTYPE.itemsize, scene.LIGHT_DTYPE.itemsize, C.sizeof(api.SceneDesc), C.sizeof(api.Camera),
            C.sizeof(api.RenderParams), C.sizeof(api.Stats), api.RAY_DTYPE.itemsize, api.HIT_DTYPE.itemsize]
    assert got == want


def test_no_cpu_fallback(lib, flat_scenes):
    """On a box without a GPU the product path must fail, not compute on the CPU."""
    from master_b200 import api
    if lib.b2pt_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(api.B2ptError) as e:
        api.DeviceSceneHandle(flat_scenes("CornellBoxDiffuse"))
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)
    pts = np.zeros((4, 3), dtype=np.float32)
    h = C.c_void_p()
    assert lib.b2pt_grid_build(pts.ctypes.data_as(C.c_void_p), 4, C.c_float(0.1), C.byref(h)) != 0


def test_argument_checks(lib):
    assert lib.b2pt_scene_create(None, None) == 1
    assert b"null" in lib.b2pt_last_error()
    assert lib.b2pt_render(None, None, None, None, None, None) == 1
    assert lib.b2pt_trace(None, None, 0, 0, None) == 1


def test_product_does_not_touch_the_oracle():
    """Nothing under master_b200/ may import, link or execute oracle/ (the oracle is test infrastructure)."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "master_b200")):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                for line in text.splitlines():
                    s = line.strip()
                    if s.startswith(("#include", "import ", "from ")) or "subprocess" in s or "dlopen" in s or "CDLL" in s:
                        assert "oracle" not in s, (f, s)
            if f.endswith(".sh"):  # build recipes: no include path, library path or source under oracle/
                for line in open(os.path.join(dirpath, f), errors="replace").read().splitlines():
                    if not line.strip().startswith("#"):
                        assert "oracle" not in line, (f, line)


def test_b2s_round_trip(tmp_path, flat_scenes):
    from master_b200.scene import load_b2s
    a = flat_scenes("CornellBoxDiffuse")
    p = str(tmp_path / "x.b2s")
    a.save(p)
    assert open(p, "rb").read() == open(golden_path("scenes", "CornellBoxDiffuse.b2s"), "rb").read()
    b = load_b2s(p)
    assert b.num_triangles == a.num_triangles == 32  # 30 mesh triangles + the 2-triangle light quad
