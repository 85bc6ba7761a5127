#!/usr/bin/env python
"""Per-source-line share of executed warp instructions and stall samples of one kernel launch in an .ncu-rep.

  python tools/ncu_lines.py gpurun_out/x.ncu-rep regex:bpt_trace [launch-skip] [top]
"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    skip = sys.argv[3] if len(sys.argv) > 3 else "0"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", kernel,
                          "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname, ix, lines = None, None, collections.OrderedDict()
    cur = None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            ix = {}
            for i, h in enumerate(r):
                ix.setdefault(h, i)
            continue
        if ix is None or len(r) < 8:
            continue
        if r[0].strip():
            cur = (fname, int(r[0]))
            lines.setdefault(cur, dict(src=r[1].strip()[:100], inst=0, thr=0, samp=0, lsb=0, ssb=0))
        if cur is None:
            continue
        def num(key):
            try:
                return int(r[ix[key]])
            except (ValueError, KeyError, IndexError):
                return 0
        try:
            n = int(r[ix["Instructions Executed"]])
        except (ValueError, KeyError):
            continue
        d = lines[cur]
        d["inst"] += n
        d["thr"] += num("Thread Instructions Executed")
        d["samp"] += num("# Samples")
        d["lsb"] += num("stall_long_sb")
        d["ssb"] += num("stall_short_sb")
    tot = sum(d["inst"] for d in lines.values()) or 1
    tots = sum(d["samp"] for d in lines.values()) or 1
    print("total warp instructions %d, samples %d" % (tot, tots))
    for (f, ln), d in sorted(lines.items(), key=lambda kv: -kv[1]["samp"])[:top]:
        print("%5.2f%% inst %5.2f%% samp (long_sb %4.1f%% short_sb %4.1f%%) lanes %4.1f  %s:%d | %s" % (
            100.0 * d["inst"] / tot, 100.0 * d["samp"] / tots, 100.0 * d["lsb"] / tots, 100.0 * d["ssb"] / tots, d["thr"] / max(d["inst"], 1), f, ln, d["src"]))


if __name__ == "__main__":
    main()
