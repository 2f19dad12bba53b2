#!/usr/bin/env python
"""Summarise an `ncu --set full` capture (`ncu -i x.ncu-rep --page raw --csv > raw.csv`) as a
markdown table per kernel launch: duration, occupancy, issue/LSU utilisation, DRAM traffic and
the warp-stall breakdown.  usage: ncu_summary.py raw.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}


def val(r, name, scale=1.0, fmt="%.2f"):
    if name not in idx or r[idx[name]] in ("", "n/a"):
        return "-"
    try:
        return fmt % (float(r[idx[name]].replace(",", "")) * scale)
    except ValueError:
        return r[idx[name]]


def to_mb(r, name):
    if name not in idx:
        return "-"
    u = units[idx[name]].lower()
    f = {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3}.get(u, 1.0)
    return "%.1f" % (float(r[idx[name]].replace(",", "")) * f)


cols = [
    ("time ms", lambda r: val(r, "gpu__time_duration.sum", {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(
        units[idx["gpu__time_duration.sum"]], 1.0), "%.3f")),
    ("grid x block", lambda r: "%s x %s" % (r[idx["launch__grid_size"]], r[idx["launch__block_size"]])),
    ("regs", lambda r: val(r, "launch__registers_per_thread", 1, "%d")),
    ("smem/CTA KB", lambda r: val(r, "launch__shared_mem_per_block_dynamic", {"byte": 1 / 1024, "kbyte": 1.0}.get(
        units[idx.get("launch__shared_mem_per_block_dynamic", 0)].lower(), 1.0), "%.1f")),
    ("warps active %", lambda r: val(r, "sm__warps_active.avg.pct_of_peak_sustained_active", 1, "%.1f")),
    ("issue active %", lambda r: val(r, "smsp__issue_active.avg.pct" if "smsp__issue_active.avg.pct" in idx
                                     else "sm__inst_issued.avg.pct_of_peak_sustained_active", 1, "%.1f")),
    ("IPC/SM", lambda r: val(r, "sm__inst_executed.avg.per_cycle_elapsed")),
    ("LSU wavefronts %", lambda r: val(r, "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", 1, "%.1f")),
    ("LSU writeback %", lambda r: val(r, "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed", 1, "%.1f")),
    ("smem bank conflicts M", lambda r: val(r, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", 1e-6, "%.1f")),
    ("DRAM read MB", lambda r: to_mb(r, "dram__bytes_read.sum")),
    ("DRAM write MB", lambda r: to_mb(r, "dram__bytes_write.sum")),
    ("DRAM thr %", lambda r: val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1, "%.1f")),
    ("L2 hit %", lambda r: val(r, "lts__t_sector_hit_rate.pct", 1, "%.1f")),
]
print("| # | kernel | " + " | ".join(c for c, _ in cols) + " |")
print("|---|---|" + "---|" * len(cols))
for k, r in enumerate(rows[2:]):
    name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
    print("| %d | %s | " % (k, name) + " | ".join(f(r) for _, f in cols) + " |")

stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and
          h.endswith("_per_issue_active.ratio") and "not_issued" not in h]
print()
print("Warp stall reasons, cycles per issued instruction (top 6 per launch):")
print()
for k, r in enumerate(rows[2:]):
    vals = []
    for h in stalls:
        try:
            vals.append((float(r[idx[h]]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
        except ValueError:
            pass
    vals.sort(reverse=True)
    print("- #%d: " % k + ", ".join("%s %.2f" % (n, v) for v, n in vals[:6]))
