"""master_b200/csrc/rcp_table.inc is the RCPPS table of the x86 host (tools/make_rcp_table.c): when this host is an Intel
CPU with a C compiler, regenerate it and compare; the generator itself verifies that RCPPS is an 11-bit table with exact
exponent scaling before it prints anything."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _intel():
    try:
        return "GenuineIntel" in open("/proc/cpuinfo").read()
    except OSError:
        return False


def test_rcp_table_has_2048_entries_in_range():
    text = open(os.path.join(ROOT, "master_b200", "csrc", "rcp_table.inc")).read()
    vals = [int(t.rstrip("u,"), 16) for t in text.split() if t.startswith("0x")]
    assert len(vals) == 2048
    # 1 / [1, 2) lies in (0.5, 1]: exponent field 126, and the table is monotonically non-increasing with 12 fraction bits
    assert all((v >> 23) == 126 for v in vals) and all(v & 0x7FF == 0 for v in vals)
    assert all(a >= b for a, b in zip(vals, vals[1:]))
    # accuracy RCPPS documents: relative error <= 1.5 * 2^-12
    import struct
    for k in (0, 1, 777, 2047):
        r = struct.unpack("<f", struct.pack("<I", vals[k]))[0]
        assert abs(r * (1 + k / 2048.0) - 1) <= 1.5 * 2 ** -12 + 2 ** -11  # + the width of one table cell


@pytest.mark.skipif(not (_intel() and shutil.which("gcc")), reason="needs an Intel host with gcc")
def test_rcp_table_regenerates_identically(tmp_path):
    exe = str(tmp_path / "make_rcp_table")
    subprocess.check_call(["gcc", "-O1", "-msse2", os.path.join(ROOT, "tools", "make_rcp_table.c"), "-o", exe])
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    assert out == open(os.path.join(ROOT, "master_b200", "csrc", "rcp_table.inc")).read()
