#!/usr/bin/env python
"""profiles/roofline_inputs.json from ONE `ncu --set full` capture of the closest-hit stage.

  python tools/make_roofline_inputs.py gpurun_out/x_bpt_trace_full.ncu-rep gpurun_out/x_iter.log profiles/x_summary.md

bench.py never measures under a profiler: the figures only a profiler sees (DRAM bytes, warp instructions, lanes per instruction,
L2 hit rate of the launch) are read here from the capture's first bpt_trace_kernel launch, the launch's ray count from the
iteration log of the SAME command run without ncu (B2PT_LOG_ITER=1; sampling is deterministic), and the file is stamped with the
hash of the kernel sources (bench.kernel_source_hash) so that bench.py can refuse it once they change.
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    rep, iterlog, summary = sys.argv[1:4]
    skip = int(sys.argv[4]) if len(sys.argv) > 4 else 24  # launches of bpt_(trace|shadow)_kernel skipped before the capture (gpu_session.sh full)
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, data = rows[0], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    row = next(r for r in data if "bpt_trace_kernel" in r[ix["Kernel Name"]])
    num = lambda k: float(row[ix[k]].replace(",", ""))
    unit = lambda k: rows[1][ix[k]]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    # the captured launch in launch order: lanes enqueue bursts of 8 iterations one lane after the other, two matching kernels each
    per_burst = 16
    lane, rest = divmod(skip, per_burst)
    it = rest // 2
    rays = None
    for line in open(iterlog):
        m = re.match(r"b2pt_iter lane (\d+) pass0 (\d+) it (\d+) closest (\d+)", line)
        if m and int(m.group(1)) == lane and int(m.group(3)) == it:
            rays = int(m.group(4))
            break
    if rays is None:
        raise SystemExit("no iteration %d of lane %d in %s" % (it, lane, iterlog))
    d = {
        "_comment": "Per-launch figures of the closest-hit traversal stage from ONE `ncu --set full` capture (launch %d of bpt_(trace|shadow)_kernel in "
                    "`bench.py --steps 1 --warmup 0 --spp 8` = iteration %d of lane %d); ray count from the iteration log of the same command "
                    "without ncu (B2PT_LOG_ITER=1).  bench.py copies them into its roofline object and refuses them when kernel_source_hash "
                    "differs from the sources it runs." % (skip, it, lane),
        "capture": os.path.relpath(summary, ROOT),
        "kernel_source_hash": bench.kernel_source_hash(),
        "rays_in_launch": rays,
        "duration_ms": num("gpu__time_duration.sum") * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(unit("gpu__time_duration.sum"), 1e-6),
        "dram_bytes_read": num("dram__bytes_read.sum") * scale.get(unit("dram__bytes_read.sum"), 1),
        "dram_bytes_written": num("dram__bytes_write.sum") * scale.get(unit("dram__bytes_write.sum"), 1),
        "warp_instructions": num("smsp__inst_executed.sum"),
        "active_lanes_per_instruction": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
        "issue_slots_used_per_cycle_per_scheduler": num("smsp__issue_active.avg.per_cycle_active"),
        "l2_hit_pct": num("lts__t_sector_hit_rate.pct"),
        "l1_hit_pct": num("l1tex__t_sector_hit_rate.pct"),
    }
    json.dump(d, open(os.path.join(ROOT, "profiles", "roofline_inputs.json"), "w"), indent=1)
    print(json.dumps(d, indent=1))


if __name__ == "__main__":
    main()
