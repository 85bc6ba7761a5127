"""Scene ingestion without Assimp: the loader's own arithmetic — per-vertex tangent frames
(aiMeshToMesh, loader.cpp:309-369), material -> BSDF mapping (:371-406), bounding sphere (:408-432), light quads and light
BSDFs (:434-456) — restated on the host (master_b200/host/raw_import.hpp) against the reference's own classes.

tests/golden/import/*.b2i hold what the Assimp fork hands to loadScene for five of the reference's scenes (`ref_harness rawdump`,
the same importer run and flags as loader.cpp:458-487); `b2pt_master import` turns each into a .b2s; the result must be BYTE
IDENTICAL to the .b2s the Assimp-linked oracle dumps after the reference's own loader ran (tests/golden/scenes/).  CPU only."""
import lzma
import os
import subprocess

import pytest

from conftest import ROOT, golden_path

HOST = os.path.join(ROOT, "master_b200", "host", "_build", "b2pt_master")
pytestmark = pytest.mark.skipif(not os.path.exists(HOST), reason="master_b200/host/_build/b2pt_master not built")


@pytest.mark.parametrize("name", ["CornellBoxDiffuse", "CornellBoxSpecular", "TestCaseFurnace", "TestCase12", "MetalRings"])
def test_imported_scene_is_byte_identical_to_the_reference_loaders(name, tmp_path):
    src = golden_path("import", name + ".b2i")
    if not os.path.exists(src):  # the larger record is stored compressed
        src = str(tmp_path / (name + ".b2i"))
        with lzma.open(golden_path("import", name + ".b2i.xz")) as f, open(src, "wb") as g:
            g.write(f.read())
    out = str(tmp_path / (name + ".b2s"))
    r = subprocess.run([HOST, "import", src, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert open(out, "rb").read() == open(golden_path("scenes", name + ".b2s"), "rb").read()


def test_truncated_record_is_an_exception(tmp_path):
    src = golden_path("import", "CornellBoxDiffuse.b2i")
    cut = str(tmp_path / "cut.b2i")
    open(cut, "wb").write(open(src, "rb").read()[:-37])
    r = subprocess.run([HOST, "import", cut, str(tmp_path / "x.b2s")], capture_output=True, text=True)
    assert r.returncode == 1 and "truncated" in r.stderr
