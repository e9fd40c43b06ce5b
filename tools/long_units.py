"""Long responses (the units the lane walker does not take: more than QB200_LANE_STEPS KiB): the scan walker
(qb_scanwalk.cuh) against the older run-list warp walker.  usage: python tools/long_units.py [unit_kib] [dense]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from quorum_b200 import batch, synth  # noqa: E402

kib = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dense = len(sys.argv) > 2
rng = np.random.default_rng(5)
n_base = 64
base = [bytes(synth.make_unit(rng, kib * 1024, (kib // 2) if dense else 6, 0.5)) for _ in range(n_base)]
data, off = synth.pack(base)
reps = max(1, (1 << 30) // data.size)
dev = torch.device("cuda", 0)
d = torch.from_numpy(data.copy()).to(dev).repeat(reps).contiguous()
lens = torch.from_numpy(np.diff(off)).to(dev).repeat(reps)
in_off = torch.zeros(lens.numel() + 1, dtype=torch.int64, device=dev)
torch.cumsum(lens, 0, out=in_off[1:])
table = batch.tag_table(synth.TAGS5)
out = torch.empty(d.numel() + 64, dtype=torch.uint8, device=dev)
n = lens.numel()
oo = torch.empty(n + 1, dtype=torch.int64, device=dev)
scratch = torch.empty(batch.lib().qb200_strip_concat_scratch_bytes(n, n, out.numel()), dtype=torch.uint8, device=dev)
res = {}
for mode in ("1", "0"):
    os.environ["QB200_SCANWALK"] = mode
    for _ in range(3):
        batch.strip_concat_batch(table, d, in_off, None, None, hide=True, streaming=False, out_bytes=out, out_off=oo, scratch=scratch)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        batch.strip_concat_batch(table, d, in_off, None, None, hide=True, streaming=False, out_bytes=out, out_off=oo, scratch=scratch)
    e1.record()
    torch.cuda.synchronize()
    from quorum_b200 import _lib
    _lib.profile_begin(torch.cuda.current_stream().cuda_stream)
    batch.strip_concat_batch(table, d, in_off, None, None, hide=True, streaming=False, out_bytes=out, out_off=oo, scratch=scratch)
    split = ", ".join(f"{k.replace('_kernel', '')} {v:.3f}" for k, v in _lib.profile_end())
    res[mode] = (e0.elapsed_time(e1) / 10, int(oo[-1].item()), out[: int(oo[-1].item())].clone(), split)
print(f"{n} responses x {kib} KiB ({'dense' if dense else 'sparse'} tags), {d.numel() / 1e9:.2f} GB: scan walker {res['1'][0]:.3f} ms, "
      f"run-list warp walker {res['0'][0]:.3f} ms per call [{res['1'][3]} | {res['0'][3]}]; outputs equal: {res['1'][1] == res['0'][1] and bool(torch.equal(res['1'][2], res['0'][2]))}")
