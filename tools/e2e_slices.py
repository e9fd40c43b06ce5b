"""e2e (host buffers) time of the cfg2 stream feed against the number of pipeline slices (scratch measurement helper)."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import quorum_b200 as qb
from quorum_b200 import synth

dev = torch.device("cuda", 0)
n = 65536
base = synth.cfg2_base()
idx = synth.gather_units(base, n, seed=1002)
data = torch.from_numpy(np.ascontiguousarray(base[idx]).reshape(-1)).pin_memory()
co = torch.arange(0, data.numel() + 1, 64, dtype=torch.int64).pin_memory()
so = torch.arange(0, n * 256 + 1, 256, dtype=torch.int64).pin_memory()
sess = qb.batch.HostStreamFeeder(qb.batch.tag_table(synth.TAGS5), dev, data.numel(), co.numel() - 1, n)
for slices in (4, 8, 12, 16, 24, 32):
    for _ in range(2):
        sess.run(data, co, so, flush=True, slices=slices)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        sess.run(data, co, so, flush=True, slices=slices)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    print(slices, round(dt * 1e3, 2), "ms", round(data.numel() / dt / 1e9, 1), "GB/s")
