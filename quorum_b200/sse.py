"""SSE output framing on the GPU: the step right after the filter.

The reference turns every non-empty filtered delta into one server-sent event
    f"data: {json.dumps(stream_event)}\\n\\n".encode()          (src/quorum/oai_proxy.py:629-652)
and the joined text into the final event                         (src/quorum/oai_proxy.py:847-862).
Only `delta.content` varies, so the envelope is rendered once per backend with the same json.dumps call and the
per-delta work - JSON string escaping (CPython's ensure_ascii encoder) and assembly - runs as one batched kernel pair.
"""
from __future__ import annotations

import ctypes
import json
from typing import List, Optional, Sequence, Tuple

import torch

from . import batch as _b
from ._lib import QB200_SKIP_EMPTY, check, lib
from .filter import _device, _enc, _to_device

_MARK = '"content": ""'


def event_envelope(event_id: str, created: int, finish_reason: Optional[str] = None) -> Tuple[bytes, bytes]:
    """(prefix, suffix) such that prefix + json_escape(content) + suffix == f"data: {json.dumps(event)}\\n\\n".encode()
    for the reference's event dict (same keys, same order, oai_proxy.py:629-645 / 847-859)."""
    event = {
        "id": event_id,
        "object": "chat.completion.chunk",
        "created": created,
        "model": "parallel-proxy",
        "choices": [{"index": 0, "delta": {"content": ""}, "finish_reason": finish_reason}],
    }
    s = f"data: {json.dumps(event)}\n\n"
    k = s.index(_MARK) + len(_MARK) - 1          # position of the closing quote of the empty content string
    return s[:k].encode(), s[k:].encode()


def backend_event_id(i: int) -> str:
    return f"chatcmpl-parallel-{i}"             # oai_proxy.py:630


FINAL_EVENT_ID = "chatcmpl-parallel-final"       # oai_proxy.py:848


def frame_events_device(in_bytes: torch.Tensor, in_off: torch.Tensor, item_prefix: Optional[torch.Tensor],
                        prefixes: Sequence[bytes], suffix: bytes, *, skip_empty: bool = True):
    """Device tensors in, device tensors out: event k is out_bytes[out_off[k]:out_off[k+1]] (empty when skipped)."""
    dev = in_off.device
    n = in_off.numel() - 1
    pb = torch.frombuffer(bytearray(b"".join(prefixes) + b"\0"), dtype=torch.uint8).to(dev)
    po = [0]
    for p in prefixes:
        po.append(po[-1] + len(p))
    po_t = torch.tensor(po, dtype=torch.int64).to(dev)
    sf = torch.frombuffer(bytearray(suffix + b"\0"), dtype=torch.uint8).to(dev)
    cap = 6 * in_bytes.numel() + n * (max(len(p) for p in prefixes) + len(suffix)) + 16
    with torch.cuda.device(dev):
        out = torch.empty(cap, dtype=torch.uint8, device=dev)
        out_off = torch.empty(n + 1, dtype=torch.int64, device=dev)
        scratch = torch.empty(lib().qb200_sse_frame_scratch_bytes(n), dtype=torch.uint8, device=dev)
        check(lib().qb200_sse_frame_batch(
            _b._ptr(in_bytes), _b._ptr(in_off), n, _b._ptr(item_prefix), _b._ptr(pb), _b._ptr(po_t), len(prefixes), _b._ptr(sf),
            len(suffix), QB200_SKIP_EMPTY if skip_empty else 0, _b._ptr(out), cap, _b._ptr(out_off), _b._ptr(scratch),
            scratch.numel(), _b._stream_ptr(dev)))
    return out, out_off


def frame_events(texts: Sequence[str], backend_index: Sequence[int], created: int, *, skip_empty: bool = True,
                 device=None) -> List[bytes]:
    """The reference's per-delta SSE events for a batch of filtered deltas: texts[k] came from backend backend_index[k].
    Returns one bytes object per text (b"" where the reference emits nothing because the delta is empty)."""
    if not texts:
        return []
    dev = _device(device)
    ids = sorted(set(int(i) for i in backend_index))
    slot = {i: k for k, i in enumerate(ids)}
    env = [event_envelope(backend_event_id(i), created) for i in ids]
    data, off = _b.pack_host([_enc(t) for t in texts])
    ip = torch.tensor([slot[int(i)] for i in backend_index], dtype=torch.int32)
    out, out_off = frame_events_device(_to_device(data, dev), _to_device(off, dev), ip.to(dev), [e[0] for e in env], env[0][1],
                                       skip_empty=skip_empty)
    return _b.split_host(out, out_off)


def frame_final_event(text: str, created: int, device=None) -> bytes:
    """The final aggregated event (finish_reason "stop", oai_proxy.py:847-860)."""
    dev = _device(device)
    prefix, suffix = event_envelope(FINAL_EVENT_ID, created, "stop")
    data, off = _b.pack_host([_enc(text)])
    out, out_off = frame_events_device(_to_device(data, dev), _to_device(off, dev), None, [prefix], suffix, skip_empty=False)
    return _b.split_host(out, out_off)[0]
