#!/usr/bin/env python
"""Summarises an .ncu-rep (ncu --set full) into the few numbers the design discussion uses.

  python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.md

Reads the report with `ncu -i ... --page raw --csv` and `--page source --csv` (no GPU needed).
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__registers_per_thread", "registers/thread"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("smsp__issue_active.avg.per_cycle_active", "issue slots used / cycle / SMSP"),
    ("smsp__warps_eligible.avg.per_cycle_active", "eligible warps / cycle / SMSP"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active lanes per instruction (of 32)"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit rate %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
]


def page(rep, name, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"] + list(extra), capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows = page(rep, "raw")
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    print("# ncu summary of `%s`\n" % rep.split("/")[-1])
    print("| launch | " + " | ".join(label for _, label in KEYS) + " |")
    print("|---|" + "---|" * len(KEYS))
    for r in data:
        name = r[ix["Kernel Name"]].split("(")[0].split("::")[-1]
        cells = []
        for k, _ in KEYS:
            if k in ix:
                v = r[ix[k]]
                try:
                    v = "%.4g" % float(v)
                except ValueError:
                    pass
                cells.append("%s %s" % (v, units[ix[k]]))
            else:
                cells.append("n/a")
        print("| %s | %s |" % (name, " | ".join(cells)))
    # stall reasons + lane utilisation per opcode class from the source page of the first launch of each kernel
    seen = set()
    for li, r in enumerate(data):
        name = r[ix["Kernel Name"]].split("(")[0].split("::")[-1]
        if name in seen:
            continue
        seen.add(name)
        src = page(rep, "source", ["--launch-skip", str(li), "--launch-count", "1"])
        if len(src) < 3:
            continue
        h = src[1]
        sx = {k: i for i, k in enumerate(h)}
        rows2 = [q for q in src[2:] if q and q[0].startswith("0x")]

        def num(q, k):
            try:
                return int(q[sx[k]] or 0)
            except (ValueError, IndexError, KeyError):
                return 0
        tot_s = sum(num(q, "# Samples") for q in rows2) or 1
        tot_i = sum(num(q, "Instructions Executed") for q in rows2) or 1
        stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
        agg = sorted(((sum(num(q, s) for q in rows2), s) for s in stalls), reverse=True)[:6]
        print("\n## %s — warp-state samples (launch %d)\n" % (name, li))
        print(", ".join("%s %.1f %%" % (s, 100.0 * v / tot_s) for v, s in agg))
        ops = collections.Counter()
        lanes = collections.Counter()
        for q in rows2:
            toks = [t for t in q[sx["Source"]].split() if not t.startswith("@")]
            if not toks:
                continue
            op = toks[0].split(".")[0]
            ops[op] += num(q, "Instructions Executed")
            lanes[op] += num(q, "Thread Instructions Executed")
        print("\n| opcode | share of warp instructions | avg active lanes |\n|---|---|---|")
        for op, n in ops.most_common(12):
            print("| %s | %.1f %% | %.1f |" % (op, 100.0 * n / tot_i, lanes[op] / max(n, 1)))


if __name__ == "__main__":
    main()
