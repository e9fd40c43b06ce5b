"""Markdown summary of an `ncu --set full` report: per captured kernel the duration, DRAM bytes, occupancy, issue-slot use,
lanes per instruction and the top warp-stall reasons.  usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [title]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
title = sys.argv[2] if len(sys.argv) > 2 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units = rows[0], rows[1]
col = {n: i for i, n in enumerate(h)}


def g(r, name, scale=1.0, nd=1):
    if name not in col or r[col[name]] == "":
        return "-"
    try:
        return f"{float(r[col[name]].replace(',', '')) * scale:.{nd}f}"
    except ValueError:
        return r[col[name]]


def bytes_of(r, name):
    v = float(r[col[name]].replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[col[name]], 1)


def us_of(r):
    v = float(r[col["gpu__time_duration.sum"]].replace(",", ""))
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(units[col["gpu__time_duration.sum"]], 1)


stall = [n for n in h if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")]
print(f"# {title}\n")
print("ncu --set full --clock-control none (cold caches, kernels serialised: shares, not absolute times, carry over)\n")
print("| kernel | us | DRAM read MB | DRAM write MB | DRAM GB/s | grid | regs | warps active % | issue active % | lanes / inst | L2 hit % | top stalls (warps per issued inst) |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|")
for r in rows[2:]:
    if len(r) < len(h):
        continue
    name = r[col["Kernel Name"]].split("(")[0].replace("qb::", "").replace("void ", "")
    rd, wr, us = bytes_of(r, "dram__bytes_read.sum"), bytes_of(r, "dram__bytes_write.sum"), us_of(r)
    st = sorted(((float(r[col[n]] or 0), n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]) for n in stall), reverse=True)[:3]
    print(f"| {name} | {us:.1f} | {rd / 1e6:.1f} | {wr / 1e6:.1f} | {(rd + wr) / us / 1e3:.0f} | {g(r, 'launch__grid_size', 1, 0)} | "
          f"{g(r, 'launch__registers_per_thread', 1, 0)} | {g(r, 'sm__warps_active.avg.pct_of_peak_sustained_active')} | "
          f"{g(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active')} | {g(r, 'smsp__thread_inst_executed_per_inst_executed.ratio')} | "
          f"{g(r, 'lts__t_sector_hit_rate.pct')} | " + ", ".join(f"{n} {v:.1f}" for v, n in st) + " |")
