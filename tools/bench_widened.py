"""Throughput of the rows section 8 marks "next" (f1 SSE framing, f2 SSE parsing, f3 prompt assembly) on one GPU, with the
reference's own Python for the same step timed on one host core beside it.  One JSON line per row, same shape as
bench.py's (metric, value, roofline on algorithmic bytes against MEASURED_PEAKS.json).  Usage: python tools/bench_widened.py"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import quorum_b200 as qb  # noqa: E402
from quorum_b200 import batch, synth  # noqa: E402
from quorum_b200.streaming import parse_sse_chunk  # noqa: E402

dev = torch.device("cuda", 0)
peak = 6532.5
try:
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
        peak = float(json.load(f)["hbm_gbs"])
except Exception:
    pass


def timed(fn, steps=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def kernel_split(fn, steps=5):
    from quorum_b200 import _lib as qlib
    qlib.profile_begin(torch.cuda.current_stream().cuda_stream)
    for _ in range(steps):
        fn()
    agg = {}
    for name, ms in qlib.profile_end():
        agg[name] = agg.get(name, 0.0) + ms / steps
    return {k: round(v, 4) for k, v in agg.items()}


def line(row, workload, in_bytes, alg_bytes, ms, units, unit_name, cpu_gbs, cpu_sample, kernels=None):
    ach = alg_bytes / (ms * 1e-3) / 1e9
    print(json.dumps({
        "row": row, "metric": "GB/s (input bytes)", "value": in_bytes / (ms * 1e-3) / 1e9, "unit": "GB/s", "n_gpus": 1,
        "ms_per_step": ms, "config": {"workload": workload}, unit_name + "_per_s": units / (ms * 1e-3),
        "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "algorithmic_bytes_per_launch": int(alg_bytes), "kernel": "whole call, CUDA events", "kernels_ms": kernels},
        "cpu_baseline": {"value": cpu_gbs, "unit": "GB/s", "cores": 1, "kind": "reference", "sample": cpu_sample}}), flush=True)


# deltas: the 64-byte chunks of cfg2 streams (multi-byte text, tags), 2 Mi of them
base = synth.cfg2_base(256)
deltas_np = np.ascontiguousarray(base).reshape(-1, 64)
rng = np.random.default_rng(3)
n_delta = 1 << 21
pick = rng.integers(0, deltas_np.shape[0], n_delta)
flat = np.ascontiguousarray(deltas_np[pick]).reshape(-1)
d_bytes = torch.from_numpy(flat).to(dev)
d_off = torch.arange(0, n_delta * 64 + 1, 64, dtype=torch.int64, device=dev)
texts = [bytes(deltas_np[i]).decode("utf-8", "replace") for i in pick[:20000]]

# ---- f1: frame every delta as the reference's SSE event -------------------------------------------------------------
env = [qb.sse.event_envelope(qb.sse.backend_event_id(i), 1700000000) for i in range(8)]
ip = torch.from_numpy((np.arange(n_delta) % 8).astype(np.int32)).to(dev)
out = qb.sse.frame_events_device(d_bytes, d_off, ip, [e[0] for e in env], env[0][1])
total_out = int(out[1][-1].item())
ms = timed(lambda: qb.sse.frame_events_device(d_bytes, d_off, ip, [e[0] for e in env], env[0][1]))
t0 = time.perf_counter()
for i, t in enumerate(texts):
    ev = {"id": f"chatcmpl-parallel-{i % 8}", "object": "chat.completion.chunk", "created": 1700000000, "model": "parallel-proxy",
          "choices": [{"index": 0, "delta": {"content": t}, "finish_reason": None}]}
    f"data: {json.dumps(ev)}\n\n".encode()
cpu = len(texts) * 64 / (time.perf_counter() - t0) / 1e9
line("f1 SSE output framing", f"{n_delta} deltas x 64 B -> data: json.dumps(event) (oai_proxy.py:629-652)", n_delta * 64,
     n_delta * 64 + total_out + 16 * n_delta, ms, n_delta, "events", cpu, "20000 deltas, json.dumps + encode, 1 core",
     kernel_split(lambda: qb.sse.frame_events_device(d_bytes, d_off, ip, [e[0] for e in env], env[0][1])))

# ---- f2: parse HTTP chunks of 1-4 such events ------------------------------------------------------------------------
ev_bytes = [f"data: {json.dumps({'id': 'chatcmpl-1', 'object': 'chat.completion.chunk', 'created': 1700000000, 'model': 'm', 'choices': [{'index': 0, 'delta': {'content': t}, 'finish_reason': None}]})}\n\n".encode() for t in texts]
chunks = []
k = 0
while k < len(ev_bytes):
    m = 1 + (k % 4)
    chunks.append(b"".join(ev_bytes[k:k + m]))
    k += m
tile, tile_off = synth.pack(chunks)
reps = max(1, (256 << 20) // tile.size)
http = np.tile(tile, reps)
http_off = np.concatenate([[0]] + [tile_off[1:] + r * tile.size for r in range(reps)]).astype(np.int64)
h_bytes, h_off = torch.from_numpy(http).to(dev), torch.from_numpy(http_off).to(dev)
res = qb.sse.parse_http_chunks_device(h_bytes, h_off)
n_ev, out_total = res[1].numel() - 1, int(res[1][-1].item())
assert int(res[4].item()) == 0
ms = timed(lambda: qb.sse.parse_http_chunks_device(h_bytes, h_off))
t0 = time.perf_counter()
for c in chunks:
    parse_sse_chunk(c)
cpu = tile.size / (time.perf_counter() - t0) / 1e9
line("f2 SSE input parsing", f"{len(http_off) - 1} HTTP chunks of 1-4 events, {n_ev} events (oai_proxy.py:578-617)", http.size,
     http.size + out_total + 8 * (len(http_off) - 1) + 9 * n_ev, ms, n_ev, "events", cpu, f"{len(chunks)} chunks, decode/split/json.loads, 1 core",
     kernel_split(lambda: qb.sse.parse_http_chunks_device(h_bytes, h_off)))

# ---- f3: aggregator prompts: 4 answers of 4 KiB per request, labels, default template --------------------------------
b3 = synth.cfg3_base(1024)
n_resp = 1 << 17
idx = synth.gather_units(b3, n_resp, seed=5)
r_bytes = torch.from_numpy(np.ascontiguousarray(b3[idx]).reshape(-1)).to(dev)
r_off = torch.arange(0, n_resp * 4096 + 1, 4096, dtype=torch.int64, device=dev)
req_off = torch.arange(0, n_resp + 1, 4, dtype=torch.int64, device=dev)
n_req = n_resp // 4
tmpl = qb.aggregate.DEFAULT_TEMPLATE.split("{responses}")
heads = [("Original query: what is the answer to question %d?\n\n" % q + tmpl[0]).encode() for q in range(n_req)]
labels = [f"Response from LLM{i + 1}:\n".encode() for i in range(4)]
packed = qb.aggregate.pack_heads(heads, dev)
o = qb.aggregate.assemble_prompts_device(r_bytes, r_off, req_off, packed, labels, b"\n\n---\n\n", tmpl[1].encode())
out_total = int(o[1][-1].item())
ms = timed(lambda: qb.aggregate.assemble_prompts_device(r_bytes, r_off, req_off, packed, labels, b"\n\n---\n\n", tmpl[1].encode()), steps=10)
sample = [bytes(b3[i]).decode("utf-8", "replace") for i in idx[:4000]]
t0 = time.perf_counter()
for q in range(1000):
    parts = [f"Response from LLM{i + 1}:\n" + sample[4 * q + i] for i in range(4)]
    ("Original query: q\n\n" + qb.aggregate.DEFAULT_TEMPLATE.replace("{responses}", "\n\n---\n\n".join(parts))).encode()
cpu = 4000 * 4096 / (time.perf_counter() - t0) / 1e9
line("f3 aggregate prompt assembly", f"{n_req} requests x 4 answers x 4 KiB (oai_proxy.py:406-423; heads, labels and answers resident on the device)",
     n_resp * 4096, n_resp * 4096 + out_total + 16 * n_resp, ms, n_req, "requests", cpu, "1000 requests, str join/replace/encode, 1 core",
     kernel_split(lambda: qb.aggregate.assemble_prompts_device(r_bytes, r_off, req_off, packed, labels, b"\n\n---\n\n", tmpl[1].encode())))
