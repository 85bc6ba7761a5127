"""master_b200/csrc/pow_core.h + pow_tables.inc: the device's table-driven powf (common.cuh: cr_powf), checked on the host.

The header is plain C arithmetic on doubles with explicit fma, so the SAME source compiles for the host (tools/check_pow.c) and is
compared there with libquadmath's powq rounded once to float — the value the shading stages' parity contract wants
("Float transcendentals").  The tables are regenerated with mpmath and compared with the committed file."""
import os
import re
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
CSRC = os.path.join(ROOT, "master_b200", "csrc")


def _have_quadmath():
    if not shutil.which("gcc"):
        return False
    r = subprocess.run(["gcc", "-x", "c", "-", "-o", os.devnull, "-lquadmath"], input="int main(){return 0;}", text=True, capture_output=True)
    return r.returncode == 0


@pytest.mark.skipif(not _have_quadmath(), reason="needs gcc with libquadmath")
def test_table_pow_is_the_correctly_rounded_float(tmp_path):
    exe = str(tmp_path / "check_pow")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-I", CSRC, os.path.join(ROOT, "tools", "check_pow.c"), "-o", exe,
                           "-lquadmath", "-lm"])
    r = subprocess.run([exe, "2000000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"outside the fast domain (\d+)\s+tables != correctly rounded: (\d+)", r.stdout)
    assert m, r.stdout
    assert int(m.group(1)) == 0  # every sampled argument is a positive normal float with a finite exponent
    assert int(m.group(2)) == 0  # measured: 0 of 10^8; 2 * 10^6 here to stay within seconds


def test_pow_tables_regenerate_identically():
    pytest.importorskip("mpmath")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_pow_tables.py")], check=True, capture_output=True, text=True).stdout
    assert out == open(os.path.join(CSRC, "pow_tables.inc")).read()


def test_pow_tables_shape():
    text = open(os.path.join(CSRC, "pow_tables.inc")).read()
    log_rows = re.findall(r"\{0x[0-9a-f]{16}ull, 0x[0-9a-f]{16}ull\}", text)
    assert len(log_rows) == 128
    exp_part = text.split("B2PT_POW_EXP_TABLE")[1].split("#define")[0]
    assert len(re.findall(r"0x[0-9a-f]{16}ull", exp_part)) == 64
    # 2^(0/64) = 1.0 exactly
    assert re.findall(r"0x[0-9a-f]{16}ull", exp_part)[0] == "0x3ff0000000000000ull"
