"""Streaming-chunk assembly around the filter (host side of src/quorum/oai_proxy.py:578-659, 739-741, 759-764).

`StreamAssembler` restates what the reference does with every HTTP chunk of one backend's SSE stream — strict UTF-8
decode, split into events, `data: ` prefix, `[DONE]`, JSON parse, `choices[0].delta.content` — and pushes all deltas
of the chunk through the GPU filter in one call.  `ConcatenateStreamSession` holds one assembler per backend like
`tag_filters` (oai_proxy.py:543-545) and produces the final joined text (oai_proxy.py:759-764, 841).
The same parsing runs on the GPU over a batch of chunks in quorum_b200.sse.parse_http_chunks (section 8 row f2), which
hands any event it is not certain about back to parse_sse_chunk below.
"""
from __future__ import annotations

import json
from typing import List, Optional, Sequence

from .concatenate import ConcatenateStrategy
from .filter import ThinkingTagFilterBatch


def parse_sse_chunk(chunk: bytes) -> Optional[List[str]]:
    """Deltas carried by one HTTP chunk, in order; None when the chunk does not decode (it is dropped whole,
    oai_proxy.py:580-588).  Events that are not `data: `, `[DONE]`, fail to parse as JSON, or carry no
    choices[0].delta.content contribute nothing (oai_proxy.py:595-617, 653-659)."""
    try:
        text = chunk.decode()
    except UnicodeDecodeError:
        return None
    deltas: List[str] = []
    for event in text.strip().split("\n\n"):
        if not event.strip():
            continue
        if not event.startswith("data: "):
            continue
        data = event[6:].strip()
        if data == "[DONE]":
            continue
        try:
            parsed = json.loads(data)
            if "choices" in parsed and parsed["choices"]:
                delta = parsed["choices"][0].get("delta", {})
                if "content" in delta:
                    deltas.append(delta["content"])
        except json.JSONDecodeError:
            continue
    return deltas


class ConcatenateStreamSession:
    """One streaming request over `n_backends` backends (progress_streaming_aggregator, oai_proxy.py:489-885, hot lines only)."""

    def __init__(self, strategy: ConcatenateStrategy, n_backends: int, device=None, device_parse: bool = False):
        self.strategy = strategy
        self.n = n_backends
        self.device, self.device_parse = device, device_parse  # device_parse: SSE/JSON parsing on the GPU too (quorum_b200.sse)
        self.all_content = ["" for _ in range(n_backends)]          # oai_proxy.py:552
        self._content = ["" for _ in range(n_backends)]
        self._filters = (ThinkingTagFilterBatch(strategy.thinking_tags, n_backends, device)
                         if strategy.hide_intermediate_think else None)  # oai_proxy.py:543-545

    def feed_http_chunks(self, chunks: Sequence[Optional[bytes]]) -> List[List[str]]:
        """chunks[i] = the next HTTP chunk of backend i (None = nothing arrived).  Returns, per backend, the filtered
        deltas to re-emit as SSE events (empty strings are not emitted, oai_proxy.py:625-628)."""
        if self.device_parse:
            from . import sse
            present = [i for i, c in enumerate(chunks) if c is not None]
            parsed = sse.parse_http_chunks([chunks[i] for i in present], self.device)
            deltas = [[] for _ in chunks]
            for i, d in zip(present, parsed):
                deltas[i] = d or []
        else:
            deltas = [(parse_sse_chunk(c) or []) if c is not None else [] for c in chunks]
        if self._filters is not None:
            safe = self._filters.feed_chunks(deltas)               # oai_proxy.py:618-621
        else:
            safe = deltas                                          # oai_proxy.py:622-623
        emit: List[List[str]] = []
        for i in range(self.n):
            self._content[i] += "".join(safe[i])                   # oai_proxy.py:624
            emit.append([] if self.strategy.suppress_individual_responses else [t for t in safe[i] if t])
        return emit

    def finish_backend(self, i: int) -> None:
        """oai_proxy.py:739-741 — flush() always returns "" so the content is what the feeds returned."""
        self.all_content[i] = self._content[i]

    def finish(self) -> Optional[str]:
        """Final combined text, or None when there is no final event (skip_final_aggregation, or no content at all)."""
        if self._filters is not None:
            self._filters.flush()
        for i in range(self.n):
            self.finish_backend(i)
        return self.strategy.combine_stream_final(self.all_content)
