"""e2e (host buffers) time of a workload against the number of pipeline slices (scratch measurement helper).
usage: python tools/e2e_slices.py cfg3|cfg2 [slices ...]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
counts = [int(x) for x in sys.argv[2:]] or [12, 24, 48, 96]
ctx = bench.Ctx()
ctx.torch, ctx.dist, ctx.rank, ctx.world, ctx.dev = torch, None, 0, 1, torch.device("cuda", 0)
W = bench.make_workload(ctx, wl, dict(bench.WORKLOADS[wl]))
h_data = torch.empty(W.total_in, dtype=torch.uint8).pin_memory()
h_data.copy_(W.data)
sess = W.session()
compact = os.environ.get("E2E_COMPACT") == "1" and W.wl["kind"] == "stream"
if W.wl["kind"] == "stream":                 # compact: int32 chunk lengths + per-stream byte offsets; else int64 chunk offsets
    h_len, h_b, h_sbo = W.host()
    h_a = torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(h_len.to(torch.int64), 0)]).pin_memory()
else:
    h_a, h_b = W.host()
for slices in counts:
    if compact:
        call = lambda: sess.run_compact(h_data, h_len, h_b, h_sbo, flush=True, slices=slices)  # noqa: E731
    elif W.wl["kind"] == "stream":
        call = lambda: sess.run(h_data, h_a, h_b, flush=True, slices=slices)  # noqa: E731
    else:
        call = lambda: sess.run(h_data, h_a, h_b, hide=True, streaming=False, slices=slices)  # noqa: E731
    for _ in range(2):
        call()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        call()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    print(wl, "compact" if compact else "int64", "slices", slices, round(dt * 1e3, 2), "ms", round(W.total_in / dt / 1e9, 1), "GB/s")
