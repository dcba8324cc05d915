"""Summarise an `ncu --set full` report: one row per captured launch with the metrics the roofline discussion uses.
usage: python tools/ncu_full_summary.py report.ncu-rep out.md [title]"""
import csv
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
title = sys.argv[3] if len(sys.argv) > 3 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("gpu__time_duration.sum", "time us"),
        ("dram__bytes_read.sum", "dram rd MB"), ("dram__bytes_write.sum", "dram wr MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("launch__registers_per_thread", "regs"),
        ("launch__shared_mem_per_block_dynamic", "dyn smem"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %")]
idx = [(hdr.index(c), n) for c, n in cols if c in hdr]
lines = [f"# {title}", "", "`ncu --set full --clock-control none` (cold cache, serialised); units as reported by ncu "
         "(MB = 1e6 bytes).", "", "| # | " + " | ".join(n for _, n in idx) + " |", "|---|" + "---|" * len(idx)]


def fmt(v, name):
    if name == "kernel":
        v = v.replace("void ", "").replace("act::", "")
        return "`" + v[:v.index("(")] + "`" if "(" in v else "`" + v[:40] + "`"
    try:
        return f"{float(v.replace(',', '')):.2f}"
    except ValueError:
        return v


for k, r in enumerate(rows[2:]):
    lines.append(f"| {k} | " + " | ".join(fmt(r[i], n) for i, n in idx) + " |")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:12]))
