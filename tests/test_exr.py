"""The reference's EXR checkpoint format: master_b200/exr.py and master_b200/host/exr_io.hpp against
each other and against the real OpenEXR library (the one OpenCV links) — CPU only.

What the reference does (exr.cpp:183-290): FLOAT channels R, G, B, denom; un-normalised sums; file rows top-down for a
y-up buffer; every metadata entry a string attribute."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

from master_b200 import exr


def image(w, h, seed=3):
    rng = np.random.default_rng(seed)
    d = rng.random((h, w, 4), dtype=np.float32) * 8.0
    d[:, :, 3] = rng.integers(1, 40, size=(h, w))  # sample counts
    return d


def metadata():
    m = exr.options_to_dict(technique="BPT", width=37, height=21, num_samples=16, beta=2.0, seed=1234, enable_seed=True)
    m.update(exr.statistics_to_dict(num_samples=16, num_basic_rays=123456789012, num_shadow_rays=42, total_time=1.5,
                                    records=[dict(sample_index=15, rms_error=0.25, abs_error=0.125, clock_time=1.5,
                                                  frame_duration=0.09, numeric_errors=0)]))
    m["measurements[100x512x512].rms_error"] = "0.250000"  # > 31 characters: long-names flag
    return m


def cv2_read(path):
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    cv2 = pytest.importorskip("cv2")
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        pytest.skip("this OpenCV build has no OpenEXR codec")
    return img


def test_python_roundtrip_and_metadata(tmp_path):
    d, m = image(37, 21), metadata()
    p = str(tmp_path / "a.exr")
    exr.save_exr(p, m, d)
    m2, d2, has_denom = exr.load_exr(p)
    assert has_denom and np.array_equal(d, d2) and m2 == {k: str(v) for k, v in m.items()}
    # the keys and number formats the reference's Options(metadata) / statistics_t(metadata) constructors parse
    assert m2["options.technique"] == "BPT" and m2["options.beta"] == "2.000000" and m2["options.width"] == "37"
    assert m2["statistics.num_basic_rays"] == "123456789012" and m2["records[15].rms_error"] == "0.250000"


def test_openexr_library_reads_what_we_write(tmp_path):
    """OpenEXR (through OpenCV) sees the same R, G, B, top-down, and is not confused by `denom` or the attributes."""
    d = image(40, 23)
    p = str(tmp_path / "b.exr")
    exr.save_exr(p, metadata(), d)
    img = cv2_read(p)  # rows top-down, channels B, G, R
    assert img.shape[:2] == (23, 40)
    np.testing.assert_array_equal(img[:, :, :3], d[::-1, :, 2::-1])


def test_we_read_what_the_openexr_library_writes(tmp_path):
    """ZIP-compressed scanline blocks of 16 lines (OpenEXR's default, what the reference's OutputFile produces)."""
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    bgr = (rng.random((50, 33, 3), dtype=np.float32) * 4.0).astype(np.float32)
    p = str(tmp_path / "c.exr")
    if not cv2.imwrite(p, bgr, [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_FLOAT]):
        pytest.skip("this OpenCV build cannot write EXR")
    _, d, has_denom = exr.load_exr(p)
    assert not has_denom and np.all(d[:, :, 3] == 1.0)
    np.testing.assert_array_equal(d[::-1, :, 2::-1], bgr)


@pytest.fixture(scope="module")
def cpp_tool(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("exr") / "exr_roundtrip")
    r = subprocess.run(["g++", "-O1", "-std=c++14", os.path.join(ROOT, "tests", "cpp", "exr_roundtrip.cpp"), "-o", out, "-lz"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("g++ / zlib not available: " + r.stderr[-300:])
    return out


def test_cpp_writer_python_reader_and_back(cpp_tool, tmp_path):
    p = str(tmp_path / "d.exr")
    subprocess.run([cpp_tool, "write", p, "37", "21"], check=True)
    m, d, has_denom = exr.load_exr(p)
    y, x = np.mgrid[0:21, 0:37]
    for c in range(4):
        np.testing.assert_array_equal(d[:, :, c], ((x * 31 + y * 17 + c * 7) % 1013) * 0.125 + (1.0 if c == 3 else 0.0))
    assert has_denom and m["options.technique"] == "BPT" and m["measurements[100x512x512].rms_error"] == "0.250000"
    img = cv2_read(p)
    np.testing.assert_array_equal(img[:, :, :3], d[::-1, :, 2::-1])
    # Python writer -> C++ reader
    q = str(tmp_path / "e.exr")
    d2 = image(19, 40, seed=9)
    exr.save_exr(q, metadata(), d2)
    info = json.loads(subprocess.run([cpp_tool, "read", q], check=True, capture_output=True, text=True).stdout)
    w = 1 + (np.arange(19 * 40) % 7)
    assert info["width"] == 19 and info["height"] == 40 and info["has_denom"] == 1
    np.testing.assert_allclose(info["sum"], (d2.reshape(-1, 4).astype(np.float64) * w[:, None]).sum(axis=0), rtol=1e-9)
    assert info["metadata"]["records[15].frame_duration"] == "0.090000"


def test_cpp_reader_takes_zip_blocks(cpp_tool, tmp_path):
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(6)
    bgr = rng.random((35, 18, 3), dtype=np.float32)
    p = str(tmp_path / "f.exr")
    if not cv2.imwrite(p, bgr, [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_FLOAT]):
        pytest.skip("this OpenCV build cannot write EXR")
    info = json.loads(subprocess.run([cpp_tool, "read", p], check=True, capture_output=True, text=True).stdout)
    rgb = bgr[::-1, :, ::-1].reshape(-1, 3).astype(np.float64)  # y-up, R G B
    w = 1 + (np.arange(35 * 18) % 7)
    np.testing.assert_allclose(info["sum"][:3], (rgb * w[:, None]).sum(axis=0), rtol=1e-9)
    assert info["has_denom"] == 0


def test_merge_is_the_multi_gpu_reduce(tmp_path):
    """`master merge` (Options.cpp:1340-1409): sums and counts add, statistics add, one record with the total duration."""
    a, b = image(16, 8, 1), image(16, 8, 2)
    ma = dict(exr.options_to_dict(technique="PT", width=16, height=8))
    mb = dict(ma)
    ma.update(exr.statistics_to_dict(num_samples=4, num_basic_rays=100, total_time=1.0,
                                     records=[dict(sample_index=3, frame_duration=0.5, clock_time=1.0)]))
    mb.update(exr.statistics_to_dict(num_samples=6, num_basic_rays=50, total_time=2.0,
                                     records=[dict(sample_index=4, frame_duration=0.25), dict(sample_index=5, frame_duration=0.25)]))
    pa, pb, pc = (str(tmp_path / n) for n in ("a.exr", "b.exr", "c.exr"))
    exr.save_exr(pa, ma, a)
    exr.save_exr(pb, mb, b)
    exr.merge_exr(pc, pa, pb)
    m, d, _ = exr.load_exr(pc)
    np.testing.assert_array_equal(d, a + b)
    assert m["statistics.num_samples"] == "10" and m["statistics.num_basic_rays"] == "150" and m["statistics.total_time"] == "3.000000"
    assert [k for k in m if k.startswith("records[")] and m["records[3].frame_duration"] == "1.000000"
    assert not any(k.startswith("records[4]") for k in m)
    mb["options.technique"] = "BPT"
    exr.save_exr(pb, mb, b)
    with pytest.raises(ValueError):
        exr.merge_exr(pc, pa, pb)
