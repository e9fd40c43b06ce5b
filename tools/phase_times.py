"""Per-phase time split of the fused kernels' tile chains (debug build with -DQB_PHASE_TIMING, see tools/build_dbg.sh).
usage: QB200_LIB=quorum_b200/libquorum_b200_dbg.so python tools/phase_times.py cfg3 [units]"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
units = sys.argv[2] if len(sys.argv) > 2 else "0"
import torch  # noqa: E402
import bench  # noqa: E402
from quorum_b200 import _lib  # noqa: E402

ctx = bench.Ctx()
ctx.torch, ctx.dist, ctx.rank, ctx.world, ctx.dev = torch, None, 0, 1, torch.device("cuda", 0)
W = bench.make_workload(ctx, wl, dict(bench.WORKLOADS[wl]), int(units))
L = _lib.lib()
buf = (ctypes.c_ulonglong * 16)()
for _ in range(3):
    W.step()
L.qb200_debug_phases(buf)
n = 5
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ev0.record()
for _ in range(n):
    W.step()
ev1.record()
torch.cuda.synchronize()
L.qb200_debug_phases(buf)
tot = sum(buf[i] for i in range(8)) or 1
names = ["tile record", "tokenise", "walk", "lengths+scan", "look-back", "offsets", "copy", "-"]
print(f"{wl}: {ev0.elapsed_time(ev1) / n:.3f} ms per call; warp-time split over the phases:")
for i in range(7):
    print(f"  {names[i]:14s} {100.0 * buf[i] / tot:5.1f} %   {buf[i] / n / 1e6:9.2f} warp-ms per call")
