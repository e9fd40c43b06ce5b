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
from typing import List, Optional, Sequence, Tuple  # noqa: F401

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


# ---- SSE input parsing (row f2): HTTP chunks of the backends' streams -> deltas ------------------------------------------

def parse_http_chunks_device(http_bytes: torch.Tensor, http_off: torch.Tensor):
    """Device buffers in, device buffers out (qb200_sse_split_batch + qb200_sse_parse_batch).

    Returns (out_bytes, out_off, ev_kind, ev_first, n_host): event k of chunk c (k in [ev_first[c], ev_first[c+1])) has
    kind ev_kind[k] (0 no delta, 1 delta, 2 host decides) and content out_bytes[out_off[k]:out_off[k+1]].  One host
    synchronisation (the number of events sizes the second call)."""
    dev = http_off.device
    n_http = http_off.numel() - 1
    with torch.cuda.device(dev):
        ev_first = torch.empty(n_http + 1, dtype=torch.int64, device=dev)
        scratch = torch.empty(lib().qb200_sse_split_scratch_bytes(n_http), dtype=torch.uint8, device=dev)
        check(lib().qb200_sse_split_batch(_b._ptr(http_bytes), _b._ptr(http_off), n_http, _b._ptr(ev_first), _b._ptr(scratch),
                                          scratch.numel(), _b._stream_ptr(dev)))
        n_events = int(ev_first[-1].item())
        out = torch.empty(max(int(http_bytes.numel()), 16), dtype=torch.uint8, device=dev)
        out_off = torch.empty(n_events + 1, dtype=torch.int64, device=dev)
        kind = torch.empty(max(n_events, 1), dtype=torch.uint8, device=dev)
        n_host = torch.empty(1, dtype=torch.int64, device=dev)
        scratch = torch.empty(lib().qb200_sse_parse_scratch_bytes(n_events), dtype=torch.uint8, device=dev)
        check(lib().qb200_sse_parse_batch(_b._ptr(http_bytes), _b._ptr(http_off), n_http, _b._ptr(ev_first), n_events, _b._ptr(out),
                                          out.numel(), _b._ptr(out_off), _b._ptr(kind), _b._ptr(n_host), _b._ptr(scratch),
                                          scratch.numel(), _b._stream_ptr(dev)))
    return out, out_off, kind[:n_events], ev_first, n_host


def parse_http_chunks(chunks: Sequence[bytes], device=None) -> List[Optional[List[str]]]:
    """parse_sse_chunk (quorum_b200/streaming.py, i.e. oai_proxy.py:578-617) over a batch of HTTP chunks on the device:
    per chunk the list of deltas in order, or None when the chunk does not decode.  A chunk with an event the device
    parser is not certain about (kind 2) is parsed by the host function instead, exceptions included."""
    from .streaming import parse_sse_chunk
    from .filter import _dec
    if not chunks:
        return []
    dev = _device(device)
    data, off = _b.pack_host([bytes(c) for c in chunks])
    d_data = _to_device(data if data.numel() else torch.zeros(16, dtype=torch.uint8), dev)
    out, out_off, kind, ev_first, n_host = parse_http_chunks_device(d_data, _to_device(off, dev))
    first = ev_first.cpu().tolist()
    kinds = kind.cpu().numpy()
    pieces = _b.split_host(out, out_off)
    res: List[Optional[List[str]]] = []
    for c, chunk in enumerate(chunks):
        a, b = first[c], first[c + 1]
        if a == b:
            res.append(None)
        elif (kinds[a:b] == 2).any():
            res.append(parse_sse_chunk(bytes(chunk)))
        else:
            res.append([_dec(pieces[k]) for k in range(a, b) if kinds[k] == 1])
    return res


def parse_http_chunks_ex(chunks: Sequence[bytes], device=None) -> List[Tuple[Optional[List[str]], bool]]:
    """parse_sse_events (quorum_b200/streaming.py) over a batch of HTTP chunks on the device: per chunk (deltas, ended).
    A chunk with an event the device parser is not certain about (kind 2) is parsed by the host function instead."""
    from .streaming import parse_sse_events
    from .filter import _dec
    if not chunks:
        return []
    dev = _device(device)
    data, off = _b.pack_host([bytes(c) for c in chunks])
    d_data = _to_device(data if data.numel() else torch.zeros(16, dtype=torch.uint8), dev)
    out, out_off, kind, ev_first, n_host = parse_http_chunks_device(d_data, _to_device(off, dev))
    first = ev_first.cpu().tolist()
    kinds = kind.cpu().numpy()
    pieces = _b.split_host(out, out_off)
    res: List[Tuple[Optional[List[str]], bool]] = []
    for c, chunk in enumerate(chunks):
        a, b = first[c], first[c + 1]
        if a == b:
            res.append((None, False))
        elif (kinds[a:b] == 2).any():
            res.append(parse_sse_events(bytes(chunk)))
        else:
            res.append(([_dec(pieces[k]) for k in range(a, b) if kinds[k] == 1], False))
    return res


# ---- the whole streaming step on the device: HTTP chunks in, SSE events out ---------------------------------------------

class DeviceStreamPipeline:
    """What progress_streaming_aggregator does per HTTP chunk of every backend (oai_proxy.py:578-652), batched over
    `n_streams` streams with the data staying on the device between the three steps: parse the chunks (row f2), feed the
    deltas through the ThinkingTagFilter of their stream with carried state (a3), frame every non-empty filtered delta
    as the reference's SSE event (f1).  `backend_of_stream[s]` is the i of `chatcmpl-parallel-{i}` for stream s.
    `created` is the event time stamp; the reference stamps every event with int(loop.time()) (oai_proxy.py:634), so pass
    the current value to step() when it matters (the constructor's value is the default)."""

    def __init__(self, thinking_tags: Sequence[str], backend_of_stream: Sequence[int], created: int, device=None,
                 hide_intermediate_think: bool = True, suppress_individual_responses: bool = False):
        self.dev = _device(device)
        self.n = len(backend_of_stream)
        self.table = _b.tag_table(list(thinking_tags))
        self.hide = hide_intermediate_think
        self.suppress = suppress_individual_responses
        self.created = created
        self.state = _b.new_stream_state(self.n, self.dev)
        self.ids = sorted(set(int(i) for i in backend_of_stream))
        slot = {i: k for k, i in enumerate(self.ids)}
        self.slot_of_stream = torch.tensor([slot[int(i)] for i in backend_of_stream], dtype=torch.int32).to(self.dev)
        self.content = [[] for _ in range(self.n)]  # filtered bytes per stream, for the final join (oai_proxy.py:624)
        self.ended = [False] * self.n               # stream ended by an exception (oai_proxy.py:660-673): contributes nothing

    def abandon(self, s: int) -> None:
        """The stream's chunk iterator raised: nothing more is read from it, its content never reaches the final join."""
        self.ended[s] = True

    def step(self, chunks: Sequence[Optional[bytes]], created: Optional[int] = None) -> List[List[bytes]]:
        """chunks[s] = the HTTP chunk that arrived for stream s (None: nothing).  Returns, per stream, the SSE events to
        send on, in order."""
        from .streaming import parse_sse_events
        if len(chunks) != self.n:
            raise ValueError("one entry per stream")
        dev = self.dev
        created = self.created if created is None else created
        env = [event_envelope(backend_event_id(i), created) for i in self.ids]
        prefixes, suffix = [e[0] for e in env], env[0][1]
        live = [bytes(c) if (c is not None and not self.ended[s]) else b"" for s, c in enumerate(chunks)]
        data, off = _b.pack_host(live)
        d_data = _to_device(data if data.numel() else torch.zeros(16, dtype=torch.uint8), dev)
        out, out_off, kind, ev_first, n_host = parse_http_chunks_device(d_data, _to_device(off, dev))
        if int(n_host.item()):  # an event the device is not certain about: those chunks go through the reference's code path
            first, kinds = ev_first.cpu().tolist(), kind.cpu().numpy()
            pieces = _b.split_host(out, out_off)
            deltas = []
            for s, c in enumerate(live):
                a, b = first[s], first[s + 1]
                if (kinds[a:b] == 2).any():
                    ds, ended = parse_sse_events(c)     # the events before a raising one still count; the stream is over
                    deltas.append([_enc(t) for t in (ds or [])])
                    if ended:
                        self.ended[s] = True
                else:
                    deltas.append([pieces[k] for k in range(a, b) if kinds[k] == 1])
            flat = [d for ds in deltas for d in ds]
            hb, ho = _b.pack_host(flat)
            out, out_off = _to_device(hb if hb.numel() else torch.zeros(16, dtype=torch.uint8), dev), _to_device(ho, dev)
            so = [0]
            for ds in deltas:
                so.append(so[-1] + len(ds))
            ev_first = torch.tensor(so, dtype=torch.int64).to(dev)
        n_events = out_off.numel() - 1
        if self.hide:
            f_bytes, f_off, self.state = _b.stream_feed_batch(self.table, out, out_off, ev_first, self.state, state_out=self.state)
        else:
            f_bytes, f_off = out, out_off
        first = ev_first.cpu().tolist()
        texts = _b.split_host(f_bytes, f_off)
        if self.suppress:   # oai_proxy.py:625-628: nothing is sent per delta, the content still accumulates
            events = [b""] * n_events
        else:
            # the stream of every event: events are grouped by stream in order
            counts = ev_first[1:] - ev_first[:-1]
            ev_slot = torch.repeat_interleave(self.slot_of_stream, counts, output_size=n_events)
            framed, framed_off = frame_events_device(f_bytes, f_off, ev_slot, prefixes, suffix, skip_empty=True)
            events = _b.split_host(framed, framed_off)
        res: List[List[bytes]] = []
        for s in range(self.n):
            a, b = first[s], first[s + 1]
            self.content[s].extend(texts[a:b])
            res.append([e for e in events[a:b] if e])
        return res

    def finish(self) -> List[str]:
        """flush() every filter (oai_proxy.py:739-741: returns "", drops a held partial) and return each stream's content
        ("" for a stream that ended in an exception: the reference never stores it)."""
        from .filter import _dec
        _b.stream_flush_batch(self.state)
        return ["" if self.ended[s] else _dec(b"".join(c)) for s, c in enumerate(self.content)]


def tail_events_device(strategy, contents: Sequence[str], created: int, device=None) -> List[bytes]:
    """What follows the last backend (oai_proxy.py:759-885) with the join and the framing on the device: the final event
    (strip_thinking_tags + join by qb200_strip_concat_batch, event by qb200_sse_frame_batch), or the error event when no
    backend has content, nothing when skip_final_aggregation; then `[DONE]`."""
    from . import streaming as st
    out: List[bytes] = []
    if not strategy.skip_final_aggregation:
        if any(contents):
            out.append(frame_final_event(strategy.combine_stream_final(contents), created, device))
        else:
            out.append(st.error_event(created))
    out.append(st.DONE_EVENT)
    return out
