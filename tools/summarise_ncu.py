"""profiles/traffic.json + profiles/<round>_kernels.md from the ncu launch lists written by tools/round_profiles.sh."""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rnd = sys.argv[1] if len(sys.argv) > 1 else "r2"
import subprocess
try:
    commit = subprocess.check_output(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], text=True).strip()
except Exception:
    commit = None
traffic, md = {}, ["| workload | kernel | launches/call | avg us (ncu, cold) | share | DRAM read MB | DRAM write MB |", "|---|---|---|---|---|---|---|"]
for wl in ("cfg2", "cfg3", "cfg4", "cfg5"):
    path = os.path.join(ROOT, "profiles", f"{rnd}_ncu_launches_{wl}.csv")
    if not os.path.isfile(path):
        continue
    rows = [r for r in csv.reader(open(path)) if len(r) > 8]
    h = rows[0]
    ki, mi, ui, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Unit"), h.index("Metric Value"), h.index("ID")
    per = collections.OrderedDict()
    ids = set()
    for r in rows[1:]:
        name = r[ki].split("(")[0].replace("qb::", "").replace("void ", "")
        if not any(t in name for t in ("tokenize", "stream_", "strip_", "scan", "request_", "sse_", "assemble", "set_i64")):
            continue
        v = float(r[vi].replace(",", ""))
        if r[mi] == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1.0)
        else:
            v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[ui], 1.0)
        per.setdefault(name, collections.defaultdict(list))[r[mi]].append(v)
    # bench.py --steps 2 --warmup 3 = 5 timed-style calls + 2 profile-pass calls: launches per call = count / calls
    calls = max(len(per.get("tokenize_kernel", {}).get("gpu__time_duration.sum", [])), 1)
    tot_us = sum(sum(m["gpu__time_duration.sum"]) for m in per.values()) / calls
    entry = {"per_kernel": {}, "commit": commit, "source": f"ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none; profiles/{rnd}_ncu_launches_{wl}.csv"}
    dram_total = 0.0
    for name, m in per.items():
        n = len(m["gpu__time_duration.sum"])
        us = sum(m["gpu__time_duration.sum"]) / calls
        rd, wr = sum(m["dram__bytes_read.sum"]) / calls, sum(m["dram__bytes_write.sum"]) / calls
        dram_total += rd + wr
        entry["per_kernel"][name] = {"launches_per_call": n / calls, "us_per_call": us, "dram_bytes": int(rd + wr)}
        md.append(f"| {wl} | {name} | {n / calls:.0f} | {us:.1f} | {us / tot_us * 100:.1f} % | {rd / 1e6:.1f} | {wr / 1e6:.1f} |")
    entry["dram_bytes_per_launch"] = int(dram_total)
    entry["us_per_call_ncu"] = tot_us
    traffic[wl] = entry
with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
    json.dump(traffic, f, indent=1)
with open(os.path.join(ROOT, "profiles", f"{rnd}_kernels.md"), "w") as f:
    f.write(f"# {rnd}: per-kernel durations and DRAM traffic of one call (ncu launch lists, kernels serialised, cold caches)\n\n" + "\n".join(md) + "\n")
print("\n".join(md))
