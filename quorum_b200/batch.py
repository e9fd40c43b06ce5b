"""Batched device entry points: torch.uint8 / int64 CUDA tensors in, CUDA tensors out, through the C ABI.

These are the calls a batching front end makes (one call per batch of responses or of stream chunks).  torch is
used for device memory and the current stream only; all compute is in libquorum_b200.so.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import QB200_CHAIN, QB200_FLUSH, QB200_HIDE, QB200_STREAM_JOIN, STATE_BYTES, check, lib


class TagTable:
    """Host-side handle of the tag table (qb200_tags_create).  Mirrors how the reference turns `thinking_tags` into its
    patterns (oai_proxy.py:137-138, 270-274); unsupported names raise ValueError instead of being approximated."""

    def __init__(self, tags: Sequence[str]):
        tags = list(tags)
        for t in tags:
            if not isinstance(t, str):
                raise TypeError("thinking_tags must be a list of str")
        try:
            bs = [t.encode("ascii") for t in tags]
        except UnicodeEncodeError:
            raise ValueError("quorum_b200: tag names must be ASCII")
        off = (ctypes.c_int32 * (len(bs) + 1))()
        for i, b in enumerate(bs):
            off[i + 1] = off[i] + len(b)
        names = ctypes.create_string_buffer(b"".join(bs) + b"\0")
        h = ctypes.c_void_p()
        check(lib().qb200_tags_create(ctypes.cast(names, ctypes.c_void_p), ctypes.cast(off, ctypes.c_void_p), len(bs),
                                      ctypes.byref(h)))
        self._h = h
        self._destroy = lib().qb200_tags_destroy
        self.tags = tags

    @property
    def handle(self) -> ctypes.c_void_p:
        return self._h

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                self._destroy(h)
            except Exception:
                pass
            self._h = None


_TABLES: dict = {}


def tag_table(tags: Sequence[str]) -> TagTable:
    key = tuple(tags)
    t = _TABLES.get(key)
    if t is None:
        if len(_TABLES) >= 256:                       # bounded: drop the oldest entry, not the whole cache
            _TABLES.pop(next(iter(_TABLES)))
        t = _TABLES[key] = TagTable(key)
    return t


def _require_cuda(t: torch.Tensor, dtype: torch.dtype, name: str) -> None:
    if not t.is_cuda:
        raise _lib.QuorumB200Error(f"{name} must be a CUDA tensor: quorum_b200 has no CPU path")
    if t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous {dtype} tensor")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream_ptr(dev: torch.device):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def strip_concat_batch(tags: TagTable, in_bytes: torch.Tensor, in_off: torch.Tensor,
                       req_off: Optional[torch.Tensor], sep: Optional[torch.Tensor], *, hide: bool, streaming: bool,
                       out_bytes: Optional[torch.Tensor] = None, out_off: Optional[torch.Tensor] = None,
                       resp_out_len: Optional[torch.Tensor] = None, scratch: Optional[torch.Tensor] = None,
                       want_resp_len: bool = False, chain: bool = False):
    """strip_thinking_tags over every response + the concatenate join per request (qb200_strip_concat_batch).

    in_bytes: uint8[total]; in_off: int64[n_resp+1]; req_off: int64[n_req+1] or None (one response per request);
    sep: uint8[sep_len] device tensor (for the streaming-final join pass "\\n"+separator, oai_proxy.py:841).
    Returns (out_bytes, out_off, resp_out_len or None); request q is out_bytes[out_off[q]:out_off[q+1]].
    chain=True (QB200_CHAIN): this call is one slice of a larger batch, see include/quorum_b200.h.
    Nothing is synchronised: results are ordered on the current stream.
    """
    dev = in_off.device
    _require_cuda(in_off, torch.int64, "in_off")
    _require_cuda(in_bytes, torch.uint8, "in_bytes")
    n_resp = in_off.numel() - 1
    if req_off is not None:
        _require_cuda(req_off, torch.int64, "req_off")
        n_req = req_off.numel() - 1
    else:
        n_req = n_resp
    sep_len = 0 if sep is None else sep.numel()
    if sep is not None:
        _require_cuda(sep, torch.uint8, "sep")
    with torch.cuda.device(dev):
        if out_bytes is None:
            out_bytes = torch.empty(in_bytes.numel() + sep_len * max(n_resp, 1) + 16, dtype=torch.uint8, device=dev)
        if out_off is None:
            out_off = torch.empty(n_req + 1, dtype=torch.int64, device=dev)
        if resp_out_len is None and want_resp_len:
            resp_out_len = torch.empty(max(n_resp, 1), dtype=torch.int64, device=dev)
        need = lib().qb200_strip_concat_scratch_bytes(n_resp, n_req, out_bytes.numel())
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        flags = (QB200_HIDE if hide else 0) | (QB200_STREAM_JOIN if streaming else 0) | (QB200_CHAIN if chain else 0)
        check(lib().qb200_strip_concat_batch(
            tags.handle, _ptr(in_bytes), _ptr(in_off), n_resp, _ptr(req_off), n_req, _ptr(sep), sep_len, flags,
            _ptr(out_bytes), out_bytes.numel(), _ptr(out_off), _ptr(resp_out_len), _ptr(scratch), scratch.numel(),
            _stream_ptr(dev)))
    return out_bytes, out_off, resp_out_len


def new_stream_state(n_streams: int, device) -> torch.Tensor:
    """Fresh per-stream state (depth 0, nothing held): uint8[n_streams, 32], layout of qb200_stream_state."""
    st = torch.zeros((n_streams, STATE_BYTES), dtype=torch.uint8, device=device)
    st[:, 4:8] = 0xFF  # tail_len = -1
    return st


def stream_feed_batch(tags: TagTable, chunk_bytes: torch.Tensor, chunk_off: torch.Tensor, stream_off: torch.Tensor,
                      state: Optional[torch.Tensor] = None, *, flush: bool = False,
                      state_out: Optional[torch.Tensor] = None, out_bytes: Optional[torch.Tensor] = None,
                      out_chunk_off: Optional[torch.Tensor] = None, scratch: Optional[torch.Tensor] = None,
                      chain: bool = False):
    """ThinkingTagFilter.feed for every chunk of every stream, in order, state carried (qb200_stream_feed_batch).

    chunk_bytes: uint8[total]; chunk_off: int64[n_chunks+1]; stream_off: int64[n_streams+1] (chunk indices).
    state: uint8[n_streams, 32] from a previous call / new_stream_state, or None for fresh streams.
    Returns (out_bytes, out_chunk_off, state_out): the feed of chunk c returned
    out_bytes[out_chunk_off[c]:out_chunk_off[c+1]].
    chain=True (QB200_CHAIN): this call is one slice of a larger batch — chunk_bytes / chunk_off / out_bytes /
    out_chunk_off are the whole batch's tensors, stream_off (and the state tensors) are views of the slice's streams, and
    the slice's output continues where the previous slice's call left out_chunk_off.
    """
    dev = chunk_off.device
    _require_cuda(chunk_off, torch.int64, "chunk_off")
    _require_cuda(stream_off, torch.int64, "stream_off")
    _require_cuda(chunk_bytes, torch.uint8, "chunk_bytes")
    n_chunks = chunk_off.numel() - 1
    n_streams = stream_off.numel() - 1
    if state is not None:
        _require_cuda(state, torch.uint8, "state")
        if state.numel() != n_streams * STATE_BYTES:
            raise ValueError("state must be uint8[n_streams, 32]")
    with torch.cuda.device(dev):
        if out_bytes is None:
            out_bytes = torch.empty(chunk_bytes.numel() + _lib.HELD_MAX * n_streams + 16, dtype=torch.uint8, device=dev)
        if out_chunk_off is None:
            out_chunk_off = torch.empty(n_chunks + 1, dtype=torch.int64, device=dev)
        if state_out is None:
            state_out = torch.empty((n_streams, STATE_BYTES), dtype=torch.uint8, device=dev)
        need = lib().qb200_stream_feed_scratch_bytes(n_chunks, n_streams, out_bytes.numel())
        if scratch is None or scratch.numel() < need:
            scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        check(lib().qb200_stream_feed_batch(
            tags.handle, _ptr(chunk_bytes), _ptr(chunk_off), n_chunks, _ptr(stream_off), n_streams, _ptr(state),
            _ptr(state_out), (QB200_FLUSH if flush else 0) | (QB200_CHAIN if chain else 0), _ptr(out_bytes),
            out_bytes.numel(), _ptr(out_chunk_off), _ptr(scratch), scratch.numel(), _stream_ptr(dev)))
    return out_bytes, out_chunk_off, state_out


def stream_feed_batch_noalloc(tags: TagTable, chunk_bytes: torch.Tensor, chunk_off: torch.Tensor,
                              stream_off: torch.Tensor, state: Optional[torch.Tensor], *, flush: bool,
                              state_out: Optional[torch.Tensor], out_bytes: torch.Tensor,
                              out_chunk_off: torch.Tensor) -> None:
    """stream_feed_batch for hosts that must not allocate: every buffer, outputs included, is passed in by the caller and
    this wrapper creates no tensor.  The library keeps its scratch in a fixed device array, so a call handles at most
    `_lib.NOALLOC_MAX_STREAMS` streams, calls on one device must not overlap, and the first call on a device should be
    made outside stream capture (qb200_stream_feed_batch_noalloc, include/quorum_b200.h)."""
    dev = chunk_off.device
    for t, dt, nm in ((chunk_off, torch.int64, "chunk_off"), (stream_off, torch.int64, "stream_off"),
                      (chunk_bytes, torch.uint8, "chunk_bytes"), (out_bytes, torch.uint8, "out_bytes"),
                      (out_chunk_off, torch.int64, "out_chunk_off")):
        _require_cuda(t, dt, nm)
    n_chunks = chunk_off.numel() - 1
    n_streams = stream_off.numel() - 1
    if out_chunk_off.numel() < n_chunks + 1:
        raise ValueError("out_chunk_off needs n_chunks + 1 entries")
    with torch.cuda.device(dev):
        check(lib().qb200_stream_feed_batch_noalloc(
            tags.handle, _ptr(chunk_bytes), _ptr(chunk_off), n_chunks, _ptr(stream_off), n_streams, _ptr(state),
            _ptr(state_out), QB200_FLUSH if flush else 0, _ptr(out_bytes), out_bytes.numel(), _ptr(out_chunk_off),
            _stream_ptr(dev)))


def stream_flush_batch(state: torch.Tensor) -> torch.Tensor:
    """ThinkingTagFilter.flush for every stream (oai_proxy.py:352-371): drops the held partial, keeps depth."""
    _require_cuda(state, torch.uint8, "state")
    with torch.cuda.device(state.device):
        check(lib().qb200_stream_flush_batch(_ptr(state), state.numel() // STATE_BYTES, _stream_ptr(state.device)))
    return state


def offsets_from_lengths32(lengths: torch.Tensor, off: torch.Tensor, scratch: torch.Tensor) -> None:
    """off[i + 1] = off[0] + lengths[0] + .. + lengths[i] on the device (qb200_offsets_from_lengths32); off[0] must hold the start."""
    n = lengths.numel()
    with torch.cuda.device(off.device):
        check(lib().qb200_offsets_from_lengths32(_ptr(lengths), n, _ptr(off), _ptr(scratch), scratch.numel(), _stream_ptr(off.device)))


def lengths32_from_offsets(off: torch.Tensor, lengths: torch.Tensor) -> None:
    """lengths[i] = off[i + 1] - off[i] as int32 on the device (qb200_lengths32_from_offsets)."""
    with torch.cuda.device(off.device):
        check(lib().qb200_lengths32_from_offsets(_ptr(off), lengths.numel(), _ptr(lengths), _stream_ptr(off.device)))


# ---- host <-> device packing helpers -----------------------------------------------------------------------------

def pack_host(items: Sequence[bytes]) -> Tuple[torch.Tensor, torch.Tensor]:
    """Concatenate byte strings into pinned host tensors (uint8[total], int64[n+1])."""
    lens = torch.tensor([len(b) for b in items], dtype=torch.int64)
    off = torch.zeros(len(items) + 1, dtype=torch.int64)
    torch.cumsum(lens, 0, out=off[1:])
    total = int(off[-1])
    data = torch.frombuffer(bytearray(b"".join(items)), dtype=torch.uint8) if total else torch.empty(0, dtype=torch.uint8)
    if torch.cuda.is_available() and total >= (1 << 20):   # pinning costs a cudaHostAlloc: only worth it for real batches
        data, off = data.pin_memory(), off.pin_memory()    # (the per-call drop-ins feed a few hundred bytes at a time)
    return data, off


def split_host(out_bytes: torch.Tensor, out_off: torch.Tensor) -> List[bytes]:
    """Device (or host) flat output -> list of bytes, one per offset interval."""
    off = out_off.cpu().tolist()
    raw = out_bytes[: off[-1]].cpu().numpy().tobytes() if off[-1] else b""
    return [raw[off[i]:off[i + 1]] for i in range(len(off) - 1)]


# ---- host-buffer sessions: the call a host application makes (pinned host memory in, pinned host memory out) -------

def default_slices(n_bytes: int) -> int:
    """Pipeline slices of a host-buffer call: about 48 MB of input each, 16 to 64 of them.  Only the first upload and the
    last download are not overlapped, so more slices help until the per-slice kernel launches show (measured on one B200:
    cfg3, 4.1 GB: 24 slices 46.4, 48: 48.4, 64: 49.0 GB/s; cfg2, 1 GiB: 12: 42.8, 24: 44.4, 48: 45.3, 96: 39.3)."""
    return max(16, min(64, n_bytes // (48 << 20)))


class HostStreamFeeder:
    """Reusable host-facing session for stream_feed_batch: device staging buffers and pinned result buffers are
    allocated once for a maximum batch shape, every run() copies host -> device, runs the kernels, copies the released
    bytes and per-chunk offsets device -> host."""

    def __init__(self, tags: TagTable, device, max_bytes: int, max_chunks: int, max_streams: int):
        self.tags, self.dev = tags, torch.device(device)
        d = self.dev
        self.d_data = torch.empty(max_bytes, dtype=torch.uint8, device=d)
        self.d_co = torch.empty(max_chunks + 1, dtype=torch.int64, device=d)
        self.d_so = torch.empty(max_streams + 1, dtype=torch.int64, device=d)
        self.d_out = torch.empty(max_bytes + _lib.HELD_MAX * max_streams + 16, dtype=torch.uint8, device=d)
        self.d_oco = torch.empty(max_chunks + 1, dtype=torch.int64, device=d)
        self.d_state = new_stream_state(max_streams, d)
        self.d_scratch = torch.empty(lib().qb200_stream_feed_scratch_bytes(max_chunks, max_streams, self.d_out.numel()), dtype=torch.uint8,
                                     device=d)
        self.h_out = torch.empty(self.d_out.numel(), dtype=torch.uint8).pin_memory()
        self.h_oco = torch.empty(max_chunks + 1, dtype=torch.int64).pin_memory()

    def run(self, h_data: torch.Tensor, h_chunk_off: torch.Tensor, h_stream_off: torch.Tensor, *, flush: bool,
            carry_state: bool = False, slices: Optional[int] = None):
        """Returns (h2d_bytes, d2h_bytes); results are in self.h_out[: self.h_oco[n_chunks]] and self.h_oco.

        The batch is fed in `slices` slices of streams (QB200_CHAIN): while slice k is being filtered, slice k+1 is on
        its way up and slice k-1's output on its way down (PCIe is full duplex), so the call costs about
        max(upload, download) instead of their sum (measured on cfg2: 4 slices 36.3, 8: 39.3, 16: 41.9, 24: 42.9 GB/s)."""
        nb, nc, ns = h_data.numel(), h_chunk_off.numel() - 1, h_stream_off.numel() - 1
        if slices is None:
            slices = default_slices(nb)
        if ns < 4 * slices or nb < (1 << 22):
            slices = 1
        dev = self.dev
        cur = torch.cuda.current_stream(dev)
        if not hasattr(self, "_s_in"):
            self._s_in, self._s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        s_in, s_out = self._s_in, self._s_out
        s_in.wait_stream(cur)
        s_out.wait_stream(cur)
        d_data, d_co, d_so, d_oco = self.d_data[:nb], self.d_co[:nc + 1], self.d_so[:ns + 1], self.d_oco[:nc + 1]
        with torch.cuda.stream(s_in):
            d_so.copy_(h_stream_off, non_blocking=True)
            d_co[nc:nc + 1].copy_(h_chunk_off[nc:nc + 1], non_blocking=True)     # the kernels bound their reads by chunk_off[n_chunks]
        sb = [ns * k // slices for k in range(slices + 1)]                      # stream bounds
        cb = [int(h_stream_off[i]) for i in sb]                                # chunk bounds (host reads, pinned)
        bb = [int(h_chunk_off[c]) for c in cb]                                 # byte bounds
        ev_o, pending = [], None

        def drain(k):                                                          # slice k's bytes -> host, once its end is known
            ev_o[k].synchronize()
            o0 = 0 if k == 0 else int(self.h_oco[cb[k]])
            o1 = int(self.h_oco[cb[k + 1]])
            with torch.cuda.stream(s_out):
                self.h_out[o0:o1].copy_(self.d_out[o0:o1], non_blocking=True)

        for k in range(slices):
            with torch.cuda.stream(s_in):
                d_data[bb[k]:bb[k + 1]].copy_(h_data[bb[k]:bb[k + 1]], non_blocking=True)
                d_co[cb[k]:cb[k + 1] + 1].copy_(h_chunk_off[cb[k]:cb[k + 1] + 1], non_blocking=True)
                ev_in = torch.cuda.Event()
                ev_in.record(s_in)
            cur.wait_event(ev_in)
            st = self.d_state[sb[k]:sb[k + 1]]
            stream_feed_batch(self.tags, d_data, d_co, d_so[sb[k]:sb[k + 1] + 1], st if carry_state else None, flush=flush,
                              state_out=st, out_bytes=self.d_out, out_chunk_off=d_oco, scratch=self.d_scratch, chain=k > 0)
            ev_c = torch.cuda.Event()
            ev_c.record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_c)
                lo = cb[k] + (1 if k else 0)
                self.h_oco[lo:cb[k + 1] + 1].copy_(d_oco[lo:cb[k + 1] + 1], non_blocking=True)
                e = torch.cuda.Event()
                e.record(s_out)
                ev_o.append(e)
            if pending is not None:
                drain(pending)
            pending = k
        drain(pending)
        s_out.synchronize()
        cur.wait_stream(s_out)
        total = int(self.h_oco[nc])
        return nb + 8 * (nc + 1) + 8 * (ns + 1), total + 8 * (nc + 1)


    def run_compact(self, h_data: torch.Tensor, h_chunk_len: torch.Tensor, h_stream_off: torch.Tensor,
                    h_stream_byte_off: torch.Tensor, *, flush: bool, carry_state: bool = False, slices: Optional[int] = None):
        """run() with 32-bit per-chunk LENGTHS over PCIe instead of 64-bit offsets, both ways: for 64-byte chunks the int64
        offset arrays are 11 % of the upload and 20 % of the download.  h_chunk_len int32[n_chunks]; h_stream_off
        int64[n_streams + 1] (chunk indices); h_stream_byte_off int64[n_streams + 1] (byte offset of every stream's first
        chunk, last entry = total bytes).  The offsets the kernels need are built on the device
        (qb200_offsets_from_lengths32) and the per-chunk output lengths are narrowed there (qb200_lengths32_from_offsets).
        Returns (h2d_bytes, d2h_bytes); results: self.h_out[:total] and self.h_olen[:n_chunks] (int32 length of every
        feed's return value), total = self.h_ends[-1] of the call."""
        nb, nc, ns = h_data.numel(), h_chunk_len.numel(), h_stream_off.numel() - 1
        if slices is None:
            slices = default_slices(nb)
        if ns < 4 * slices or nb < (1 << 22):
            slices = 1
        dev = self.dev
        cur = torch.cuda.current_stream(dev)
        if not hasattr(self, "_s_in"):
            self._s_in, self._s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        if not hasattr(self, "d_len"):
            mc = self.d_co.numel() - 1
            self.d_len = torch.empty(mc, dtype=torch.int32, device=dev)
            self.d_olen = torch.empty(mc, dtype=torch.int32, device=dev)
            self.h_olen = torch.empty(mc, dtype=torch.int32).pin_memory()
            self.d_cscratch = torch.empty(lib().qb200_offsets_scratch_bytes(mc), dtype=torch.uint8, device=dev)
            self.h_ends_buf = torch.zeros(257, dtype=torch.int64).pin_memory()
        s_in, s_out = self._s_in, self._s_out
        s_in.wait_stream(cur)
        s_out.wait_stream(cur)
        d_data, d_co, d_so, d_oco = self.d_data[:nb], self.d_co[:nc + 1], self.d_so[:ns + 1], self.d_oco[:nc + 1]
        slices = min(slices, 256)
        h_ends = self.h_ends_buf[:slices + 1]
        h_ends[0] = 0
        with torch.cuda.stream(s_in):
            d_so.copy_(h_stream_off, non_blocking=True)
            d_co[nc:nc + 1].copy_(h_stream_byte_off[ns:ns + 1], non_blocking=True)   # the kernels bound their reads by chunk_off[n_chunks]
            d_co[0:1].zero_()
        sb = [ns * k // slices for k in range(slices + 1)]                      # stream bounds
        cb = [int(h_stream_off[i]) for i in sb]                                # chunk bounds (host reads, pinned)
        bb = [int(h_stream_byte_off[i]) for i in sb]                           # byte bounds
        ev_o, pending = [], None

        def drain(k):                                                          # slice k's bytes -> host, once its end is known
            ev_o[k].synchronize()
            o0, o1 = int(h_ends[k]), int(h_ends[k + 1])
            with torch.cuda.stream(s_out):
                self.h_out[o0:o1].copy_(self.d_out[o0:o1], non_blocking=True)

        for k in range(slices):
            with torch.cuda.stream(s_in):
                d_data[bb[k]:bb[k + 1]].copy_(h_data[bb[k]:bb[k + 1]], non_blocking=True)
                self.d_len[cb[k]:cb[k + 1]].copy_(h_chunk_len[cb[k]:cb[k + 1]], non_blocking=True)
                ev_in = torch.cuda.Event()
                ev_in.record(s_in)
            cur.wait_event(ev_in)
            offsets_from_lengths32(self.d_len[cb[k]:cb[k + 1]], d_co[cb[k]:], self.d_cscratch)   # continues from d_co[cb[k]]
            st = self.d_state[sb[k]:sb[k + 1]]
            stream_feed_batch(self.tags, d_data, d_co, d_so[sb[k]:sb[k + 1] + 1], st if carry_state else None, flush=flush,
                              state_out=st, out_bytes=self.d_out, out_chunk_off=d_oco, scratch=self.d_scratch, chain=k > 0)
            lengths32_from_offsets(d_oco[cb[k]:], self.d_olen[cb[k]:cb[k + 1]])
            ev_c = torch.cuda.Event()
            ev_c.record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_c)
                self.h_olen[cb[k]:cb[k + 1]].copy_(self.d_olen[cb[k]:cb[k + 1]], non_blocking=True)
                h_ends[k + 1:k + 2].copy_(d_oco[cb[k + 1]:cb[k + 1] + 1], non_blocking=True)
                e = torch.cuda.Event()
                e.record(s_out)
                ev_o.append(e)
            if pending is not None:
                drain(pending)
            pending = k
        drain(pending)
        s_out.synchronize()
        cur.wait_stream(s_out)
        self.h_ends = h_ends
        total = int(h_ends[slices])
        return nb + 4 * nc + 16 * (ns + 1), total + 4 * nc + 8 * slices


class HostStripConcat:
    """Reusable host-facing session for strip_concat_batch (see HostStreamFeeder)."""

    def __init__(self, tags: TagTable, device, max_bytes: int, max_resp: int, max_req: int, sep: bytes):
        self.tags, self.dev = tags, torch.device(device)
        d = self.dev
        self.sep = torch.frombuffer(bytearray(sep), dtype=torch.uint8).to(d) if sep else None
        self.d_data = torch.empty(max_bytes, dtype=torch.uint8, device=d)
        self.d_io = torch.empty(max_resp + 1, dtype=torch.int64, device=d)
        self.d_ro = torch.empty(max_req + 256, dtype=torch.int64, device=d)  # every slice of run() has its own rebased copy (<= 255 slices)
        self.d_out = torch.empty(max_bytes + len(sep) * max_resp + 16, dtype=torch.uint8, device=d)
        self.d_oo = torch.empty(max_req + 1, dtype=torch.int64, device=d)
        self.d_scratch = torch.empty(lib().qb200_strip_concat_scratch_bytes(max_resp, max_req, self.d_out.numel()),
                                     dtype=torch.uint8, device=d)
        self.h_out = torch.empty(self.d_out.numel(), dtype=torch.uint8).pin_memory()
        self.h_oo = torch.empty(max_req + 1, dtype=torch.int64).pin_memory()

    def run(self, h_data: torch.Tensor, h_in_off: torch.Tensor, h_req_off: torch.Tensor, *, hide: bool, streaming: bool,
            slices: Optional[int] = None):
        """Returns (h2d_bytes, d2h_bytes); results are in self.h_out[: self.h_oo[n_req]] and self.h_oo.

        The batch is processed in `slices` slices of requests (QB200_CHAIN) on three streams: the upload of slice k+1,
        the kernels of slice k and the download of slice k-1 overlap, so the call costs about max(upload, download)."""
        nb, nr, nq = h_data.numel(), h_in_off.numel() - 1, h_req_off.numel() - 1
        if slices is None:
            slices = default_slices(nb)
        slices = min(slices, 255)
        if nq < 4 * slices or nb < (1 << 22):
            slices = 1
        dev = self.dev
        cur = torch.cuda.current_stream(dev)
        if not hasattr(self, "_s_in"):
            self._s_in, self._s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
            self.h_ro = torch.empty(self.d_ro.numel(), dtype=torch.int64).pin_memory()
        s_in, s_out = self._s_in, self._s_out
        s_in.wait_stream(cur)
        s_out.wait_stream(cur)
        data, io, ro, oo = self.d_data[:nb], self.d_io[:nr + 1], self.d_ro, self.d_oo[:nq + 1]
        qb_ = [nq * k // slices for k in range(slices + 1)]                    # request bounds
        rb = [int(h_req_off[q]) for q in qb_]                                  # response bounds
        bb = [int(h_in_off[r]) for r in rb]                                    # byte bounds
        # a slice's requests as response indices relative to the slice's first response (one entry shared by neighbours
        # would need two values, so every slice gets its own nq_k + 1 entries)
        h_ro, pos, ro_at = self.h_ro, 0, []
        for k in range(slices):
            m = qb_[k + 1] - qb_[k] + 1
            torch.sub(h_req_off[qb_[k]:qb_[k + 1] + 1], rb[k], out=h_ro[pos:pos + m])
            ro_at.append((pos, m))
            pos += m
        ev_o, pending = [], None

        def drain(k):                                                          # slice k's bytes -> host, once its end is known
            ev_o[k].synchronize()
            o0 = 0 if k == 0 else int(self.h_oo[qb_[k]])
            o1 = int(self.h_oo[qb_[k + 1]])
            with torch.cuda.stream(s_out):
                self.h_out[o0:o1].copy_(self.d_out[o0:o1], non_blocking=True)

        for k in range(slices):
            p0, m = ro_at[k]
            with torch.cuda.stream(s_in):
                data[bb[k]:bb[k + 1]].copy_(h_data[bb[k]:bb[k + 1]], non_blocking=True)
                io[rb[k]:rb[k + 1] + 1].copy_(h_in_off[rb[k]:rb[k + 1] + 1], non_blocking=True)
                ro[p0:p0 + m].copy_(h_ro[p0:p0 + m], non_blocking=True)
                ev_in = torch.cuda.Event()
                ev_in.record(s_in)
            cur.wait_event(ev_in)
            strip_concat_batch(self.tags, data, io[rb[k]:rb[k + 1] + 1], ro[p0:p0 + m], self.sep, hide=hide, streaming=streaming,
                               out_bytes=self.d_out, out_off=oo[qb_[k]:qb_[k + 1] + 1], scratch=self.d_scratch, chain=k > 0)
            ev_c = torch.cuda.Event()
            ev_c.record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_c)
                lo = qb_[k] + (1 if k else 0)
                self.h_oo[lo:qb_[k + 1] + 1].copy_(oo[lo:qb_[k + 1] + 1], non_blocking=True)
                e = torch.cuda.Event()
                e.record(s_out)
                ev_o.append(e)
            if pending is not None:
                drain(pending)
            pending = k
        drain(pending)
        s_out.synchronize()
        cur.wait_stream(s_out)
        total = int(self.h_oo[nq])
        return nb + 8 * (nr + 1) + 8 * (nq + 1), total + 8 * (nq + 1)
