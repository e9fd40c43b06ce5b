"""Drop-in replacements for the reference's two filter names, computed on the GPU.

    quorum.oai_proxy.strip_thinking_tags   (src/quorum/oai_proxy.py:120-139)
    quorum.oai_proxy.ThinkingTagFilter     (src/quorum/oai_proxy.py:262-371)

Same names, signatures, return values and exception types.  The per-call forms exist for parity and for patching
into the reference (quorum_b200.install); throughput comes from the *_many / batch forms, which put a whole batch
through one launch sequence.
"""
from __future__ import annotations

import struct
from typing import List, Optional, Sequence

import torch

from . import batch as _b


def _enc(s: str) -> bytes:
    return s.encode("utf-8", "surrogatepass")


def _dec(b: bytes) -> str:
    return b.decode("utf-8", "surrogatepass")


def _device(device=None) -> torch.device:
    if device is not None:
        return torch.device(device)
    if not torch.cuda.is_available():
        raise _b._lib.QuorumB200Error("quorum_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _to_device(data: torch.Tensor, dev: torch.device) -> torch.Tensor:
    return data.to(dev, non_blocking=True)


def strip_thinking_tags_many(contents: Sequence[str], tags: List[str], hide_intermediate: bool = True,
                             device=None) -> List[str]:
    """strip_thinking_tags over a list of contents in one batch."""
    if not hide_intermediate:
        return list(contents)
    for c in contents:
        if not isinstance(c, str):
            raise TypeError(f"expected string or bytes-like object, got '{type(c).__name__}'")
    if not contents:
        return []
    dev = _device(device)
    table = _b.tag_table(tags)
    data, off = _b.pack_host([_enc(c) for c in contents])
    out, out_off, _ = _b.strip_concat_batch(table, _to_device(data, dev), _to_device(off, dev), None, None,
                                            hide=True, streaming=False)
    return [_dec(x) for x in _b.split_host(out, out_off)]


def strip_thinking_tags(content: str, tags: List[str], hide_intermediate: bool = True) -> str:
    """Strip reasoning tags from content (reference signature, oai_proxy.py:120-122).

    hide_intermediate False returns `content` unchanged, not even stripped (oai_proxy.py:134-135)."""
    if not hide_intermediate:
        return content
    return strip_thinking_tags_many([content], tags, True)[0]


class ThinkingTagFilterBatch:
    """N independent ThinkingTagFilter objects advanced together: one launch sequence per call.

    State (depth + held partial tag per stream) lives on the device between calls, exactly what the reference keeps in
    `thinking_depth` / `buffer` (oai_proxy.py:275-276)."""

    def __init__(self, tags: Sequence[str], n_streams: int, device=None):
        self.allowed_tags = [t.lower() for t in tags]
        self._table = _b.tag_table(list(tags))
        self.n_streams = n_streams
        self._dev = _device(device)
        self._state = _b.new_stream_state(n_streams, self._dev)

    def feed_chunks(self, chunks_per_stream: Sequence[Sequence[str]]) -> List[List[str]]:
        """chunks_per_stream[s] = the texts fed to stream s, in order.  Returns the return value of every feed."""
        if len(chunks_per_stream) != self.n_streams:
            raise ValueError("one chunk list per stream")
        flat: List[bytes] = []
        soff = [0]
        for chunks in chunks_per_stream:
            for c in chunks:
                if not isinstance(c, str):
                    raise TypeError(f'can only concatenate str (not "{type(c).__name__}") to str')
                flat.append(_enc(c))
            soff.append(len(flat))
        data, coff = _b.pack_host(flat)
        stream_off = torch.tensor(soff, dtype=torch.int64)
        out, oco, self._state = _b.stream_feed_batch(
            self._table, _to_device(data, self._dev), _to_device(coff, self._dev), stream_off.to(self._dev), self._state,
            state_out=self._state)
        pieces = [_dec(x) for x in _b.split_host(out, oco)]
        return [pieces[soff[s]:soff[s + 1]] for s in range(self.n_streams)]

    def feed(self, texts: Sequence[str]) -> List[str]:
        """One feed per stream."""
        return [r[0] for r in self.feed_chunks([[t] for t in texts])]

    def flush(self) -> List[str]:
        _b.stream_flush_batch(self._state)
        return [""] * self.n_streams

    def states(self):
        """[(depth, tail_len, held_bytes)] per stream (device -> host)."""
        raw = self._state.cpu().numpy().tobytes()
        out = []
        for s in range(self.n_streams):
            depth, tail, hl = struct.unpack_from("<iiB", raw, 32 * s)
            out.append((depth, tail, raw[32 * s + 9: 32 * s + 9 + hl]))
        return out


class ThinkingTagFilter:
    """Incremental thinking-tag filter with the reference's interface (oai_proxy.py:262-371):
    `feed(text) -> str`, `flush() -> str`, attributes `allowed_tags`, `buffer`, `thinking_depth`."""

    def __init__(self, tags):
        self.allowed_tags = [tag.lower() for tag in tags]
        self._impl = ThinkingTagFilterBatch(list(tags), 1)
        self._buffer = b""
        self._depth = 0

    @property
    def thinking_depth(self) -> int:
        return self._depth

    @property
    def buffer(self) -> str:
        return _dec(self._buffer)

    def feed(self, text: str) -> str:
        if not isinstance(text, str):
            raise TypeError(f'can only concatenate str (not "{type(text).__name__}") to str')
        out = self._impl.feed([text])[0]
        depth, tail, held = self._impl.states()[0]
        raw = _enc(text)
        if depth == 0:
            self._buffer = held                      # oai_proxy.py:321-325
        elif tail >= 0:
            self._buffer = raw[len(raw) - tail:]     # everything after the last consumed tag (oai_proxy.py:338, 344)
        else:
            self._buffer = self._buffer + raw        # oai_proxy.py:332-334: wait for more text
        self._depth = depth
        return out

    def flush(self) -> str:
        self._impl.flush()
        self._buffer = b""
        return ""
