"""Hottest CUDA source lines of one kernel in an ncu report (needs -lineinfo and --import-source on).
usage: python tools/ncu_lines.py rep.ncu-rep kernel-regex [n]"""
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", f"regex:{kern}"],
                     capture_output=True, text=True).stdout
fn, h, out = "", None, []
for r in csv.reader(io.StringIO(raw)):
    if not r:
        continue
    if r[0] == "File Path":
        fn = r[1].split("/")[-1]
    elif r[0] == "Line No":
        h = r
    elif h and r[0].isdigit():
        ci, ct, cs = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
        try:
            out.append((int(r[ci]), int(r[ct]), int(r[cs]), fn, r[0], r[1].strip()[:100]))
        except ValueError:
            pass
tot, tots = sum(o[0] for o in out) or 1, sum(o[2] for o in out) or 1
print(f"total warp instructions {tot}, samples {tots}")
for o in sorted(out, reverse=True)[:n]:
    print(f"{o[0] / tot * 100:5.1f}% inst {o[2] / tots * 100:5.1f}% samples lanes {o[1] / max(o[0], 1):5.1f}  {o[3]}:{o[4]}  {o[5]}")
