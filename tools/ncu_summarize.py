#!/usr/bin/env python
"""Compact the ncu outputs that gpurun brings back into the small tracked files under profiles/.

    python tools/ncu_summarize.py launches gpurun_out/launches.csv profiles/<tag>_launches.csv
    python tools/ncu_summarize.py full     gpurun_out/prof.ncu-rep  profiles/<tag>_ncu_full_summary.csv
"""
import csv
import re
import subprocess
import sys

FULL_METRICS = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max",
]


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = name.replace("void ", "").replace("lt::", "")
    if "at::" in name:
        m = re.search(r"(\w+Functor|normal_kernel|\w+_kernel)", name)
        name = "torch:" + (m.group(1) if m else "aten")
    return name.strip()


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    k, v, u = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    g, b = hdr.index("Grid Size"), hdr.index("Block Size")
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "kernel", "grid", "block", "gpu__time_duration_us"])
        for r in rows[1:]:
            t = float(r[v]) / (1000.0 if r[u] in ("ns", "nsecond") else 1.0)
            w.writerow([r[0], short(r[k]), r[g], r[b], "%.2f" % t])


def full(src, dst):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    want = [m for m in FULL_METRICS if m in idx]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(want)
        w.writerow([units[idx[m]] for m in want])
        for r in data:
            w.writerow([short(r[idx[m]]) if m == "Kernel Name" else r[idx[m]] for m in want])


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
