#!/usr/bin/env python
"""Condenses ncu output (run under gpurun) into the small text files kept under profiles/.

    python tools/ncu_summary.py launches gpurun_out/X_launches.csv  > profiles/rNN_launches.csv
    python tools/ncu_summary.py details  gpurun_out/X_prof.ncu-rep  > profiles/rNN_details.txt
    python tools/ncu_summary.py traffic  profiles/rNN_details.txt C2:plan2 planarg_mma_kernel<16>
        -> adds DRAM bytes per launch of that kernel to profiles/ncu_traffic.json (bench.py's roofline.traffic)
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__inst_executed_op_shared_atom.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_sectors_op_red.sum", "lts__t_bytes.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__average_warp_latency_issue_stalled_short_scoreboard.pct",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
    hdr = rows[0]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    print("# every launch inside the cudaProfilerStart/Stop range of bench.py (the timed steps), in order")
    print("id,kernel,grid,block,ms")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0].replace("void ", "")
        ms = float(r[vi].replace(",", "")) / 1e6
        if "drtvam" in r[ki]:
            print(f"{r[0]},{name},{r[gi]},{r[bi]},{ms:.4f}")
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ms
    total = sum(v[1] for v in agg.values())
    print("# per kernel: name,launches,total_ms,share")
    for name, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name},{c},{t:.3f},{t / total:.4f}")


def details(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    for r in rows[2:]:
        print(f"== {r[ki]}  grid {r[hdr.index('Grid Size')]} block {r[hdr.index('Block Size')]}")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"   {k} [{units[i]}] = {r[i]}")


def traffic(path, key, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of the FIRST launch of `kernel` in a details file."""
    import json
    import os
    import re
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    found, rd, wr, ms = False, None, None, None
    for line in open(path):
        if line.startswith("=="):
            if found and rd is not None:
                break
            found = kernel in line
            continue
        if not found:
            continue
        m = re.match(r"\s+(\S+) \[(\w+)\] = ([\d.,]+)", line)
        if not m:
            continue
        v = float(m.group(3).replace(",", ""))
        if m.group(1) == "dram__bytes_read.sum":
            rd = v * scale[m.group(2)]
        elif m.group(1) == "dram__bytes_write.sum":
            wr = v * scale[m.group(2)]
        elif m.group(1) == "gpu__time_duration.sum":
            ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(m.group(2), 1.0)
    if rd is None or wr is None:
        sys.exit(f"{kernel}: no dram__bytes_* rows in {path}")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = os.path.join(root, "profiles", "ncu_traffic.json")
    table = json.load(open(out)) if os.path.exists(out) else {}
    table[key] = {"kernel": kernel, "forward_dram_bytes": rd + wr, "dram_bytes_read": rd, "dram_bytes_write": wr,
                  "ncu_duration_ms": ms, "source": os.path.relpath(os.path.abspath(path), root) +
                  " (ncu --set full --clock-control none, one launch)"}
    json.dump(table, open(out, "w"), indent=1, sort_keys=True)
    print(json.dumps(table[key]))


if __name__ == "__main__":
    {"launches": launches, "details": details, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
