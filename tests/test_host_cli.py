"""The reference's command line on the headless GPU driver (master_b200/host/cli.hpp): `merge` — merge_exr,
the CPU analogue of the multi-GPU film reduce (Options.cpp:1340-1409) — and the parser's error behaviour
(Options.cpp:600-612, 672-680).  CPU only: no kernel is launched."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

from master_b200 import exr

HOST = os.path.join(ROOT, "master_b200", "host", "_build", "b2pt_master")
pytestmark = pytest.mark.skipif(not os.path.exists(HOST), reason="master_b200/host/_build/b2pt_master not built")


def run(*args):
    return subprocess.run([HOST] + list(args), capture_output=True, text=True)


@pytest.fixture
def pair(tmp_path):
    def img(seed, count):
        d = np.random.default_rng(seed).random((8, 16, 4), dtype=np.float32)
        d[:, :, 3] = count
        return d
    m = exr.options_to_dict(technique="PT", width=16, height=8)
    ma, mb = dict(m), dict(m)
    ma.update(exr.statistics_to_dict(num_samples=4, num_basic_rays=100, total_time=1.0,
                                     records=[dict(sample_index=3, frame_duration=0.5, clock_time=1.0)]))
    mb.update(exr.statistics_to_dict(num_samples=2, num_basic_rays=50, total_time=2.0,
                                     records=[dict(sample_index=0, frame_duration=0.25), dict(sample_index=1, frame_duration=0.25)]))
    a, b = img(1, 4), img(2, 2)
    pa, pb = str(tmp_path / "a.exr"), str(tmp_path / "b.exr")
    exr.save_exr(pa, ma, a)
    exr.save_exr(pb, mb, b)
    return pa, pb, a, b


def test_merge_matches_the_python_mirror_and_master_merge(pair, tmp_path):
    pa, pb, a, b = pair
    pc, pd = str(tmp_path / "c.exr"), str(tmp_path / "d.exr")
    assert run("merge", pc, pa, pb).returncode == 0
    exr.merge_exr(pd, pa, pb)
    mc, dc, _ = exr.load_exr(pc)
    md, dd, _ = exr.load_exr(pd)
    np.testing.assert_array_equal(dc, a + b)
    np.testing.assert_array_equal(dc, dd)
    assert mc["statistics.num_samples"] == "6" and mc["statistics.num_basic_rays"] == "150" and mc["statistics.total_time"] == "3.000000"
    assert mc["records[3].frame_duration"] == "1.000000" and mc["options.output"] == pc
    for k in ("statistics.num_samples", "statistics.num_basic_rays", "statistics.total_time", "records[3].frame_duration", "options.technique"):
        assert mc[k] == md[k]


def test_parser_errors_are_the_references():
    assert "More than one technique specified." in run("x.b2s", "--PT", "--BPT").stderr
    assert "--from-light and --from-camera cannot be specified at the same time." in run("x.b2s", "--from-light", "--from-camera").stderr
    assert "Input file is required." in run("--BPT").stderr
    assert "Invalid value for --resolution." in run("x.b2s", "--resolution=12").stderr
    assert "Unsupported option --bogus." in run("x.b2s", "--bogus").stderr
    assert "Invalid value for --gpus." in run("x.b2s", "--gpus=0").stderr
    assert "cannot open" in run("average").stderr  # the post-processing tools of main.cpp:80-107 are not part of this driver: a scene path
    r = run("/nonexistent.b2s", "--BPT", "--beta=2")
    assert r.returncode == 1 and "cannot open" in r.stderr
