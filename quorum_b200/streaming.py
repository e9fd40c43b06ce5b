"""Streaming-chunk assembly around the filter (host side of src/quorum/oai_proxy.py:530-541, 578-659, 739-741, 759-764,
841-885).

`parse_sse_events` restates what the reference does with every HTTP chunk of one backend's SSE stream — strict UTF-8
decode, split into events, `data: ` prefix, `[DONE]`, JSON parse, `choices[0].delta.content` — including how it ends:
a JSONDecodeError skips the event, any other exception ends that backend's stream for good (oai_proxy.py:653-673: the
`except Exception` sits outside the `async for`, and its `continue` skips the `all_content[i] = content` of :739-741,
so the backend contributes nothing to the final join although the events it already produced were sent).
`ConcatenateStreamSession` holds one filter per backend like `tag_filters` (oai_proxy.py:543-545), pushes all deltas of
a round of chunks through the GPU filter in one call, and renders the same SSE events as the reference (role event,
per-delta events, final / error event, `[DONE]`).
The same parsing runs on the GPU over a batch of chunks in quorum_b200.sse (section 8 row f2), which hands any event it
is not certain about back to parse_sse_events below.
"""
from __future__ import annotations

import json
from typing import List, Optional, Sequence, Tuple

from .concatenate import ConcatenateStrategy
from .filter import ThinkingTagFilterBatch


def parse_sse_events(chunk: bytes) -> Tuple[Optional[List[str]], bool]:
    """(deltas, ended) for one HTTP chunk.  deltas: the `delta.content` strings in order, None when the chunk does not
    decode (it is dropped whole, oai_proxy.py:580-588).  ended: an event made the reference's parsing block raise
    something other than JSONDecodeError (a non-object payload, a non-list `choices`, a non-str content, ...): the
    deltas before it still count, the backend's stream is over (oai_proxy.py:667-673)."""
    try:
        text = chunk.decode()
    except UnicodeDecodeError:
        return None, False
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
                    c = delta["content"]
                    if not isinstance(c, str):   # feed(c) / content += c raise TypeError (oai_proxy.py:618-624)
                        return deltas, True
                    deltas.append(c)
        except json.JSONDecodeError:
            continue
        except Exception:  # noqa: BLE001 - whatever it is, the reference's handler is `except Exception` (oai_proxy.py:667)
            return deltas, True
    return deltas, False


def parse_sse_chunk(chunk: bytes) -> Optional[List[str]]:
    """Deltas carried by one HTTP chunk, in order; None when the chunk does not decode.  The parsing block of
    oai_proxy.py:595-617 as it is: exceptions other than JSONDecodeError propagate to the caller (in the reference they
    end the backend's stream; parse_sse_events is the form that reports that instead of raising)."""
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


def _event(event_id: str, created: int, delta: dict, finish_reason) -> bytes:
    """The reference's event dict (same keys, same order) through the same json.dumps call (oai_proxy.py:629-646)."""
    ev = {"id": event_id, "object": "chat.completion.chunk", "created": created, "model": "parallel-proxy",
          "choices": [{"index": 0, "delta": delta, "finish_reason": finish_reason}]}
    return f"data: {json.dumps(ev)}\n\n".encode()


def role_event(created: int) -> bytes:
    """oai_proxy.py:530-541."""
    return _event("chatcmpl-parallel", created, {"role": "assistant"}, None)


def delta_event(i: int, text: str, created: int) -> bytes:
    """oai_proxy.py:629-646."""
    return _event(f"chatcmpl-parallel-{i}", created, {"content": text}, None)


def final_event(text: str, created: int) -> bytes:
    """oai_proxy.py:847-860."""
    return _event("chatcmpl-parallel-final", created, {"content": text}, "stop")


def error_event(created: int) -> bytes:
    """oai_proxy.py:863-881."""
    return _event("error", created, {"content": "Error: All backends failed to provide content"}, "error")


DONE_EVENT = b"data: [DONE]\n\n"   # oai_proxy.py:883


def tail_events(strategy: ConcatenateStrategy, all_content: Sequence[str], created: int, combined: Optional[str] = None) -> List[bytes]:
    """What follows the last backend (oai_proxy.py:759-885): the final event (or the error event when no backend has
    content; nothing when skip_final_aggregation), then `[DONE]`.  `combined` may carry the already joined text."""
    out: List[bytes] = []
    if not strategy.skip_final_aggregation:
        if any(all_content):
            if combined is None:
                combined = strategy.combine_stream_final(all_content)
            out.append(final_event(combined, created))
        else:
            out.append(error_event(created))
    out.append(DONE_EVENT)
    return out


class ConcatenateStreamSession:
    """One streaming request over `n_backends` backends (progress_streaming_aggregator, oai_proxy.py:489-885, hot lines only)."""

    def __init__(self, strategy: ConcatenateStrategy, n_backends: int, device=None, device_parse: bool = False):
        self.strategy = strategy
        self.n = n_backends
        self.device, self.device_parse = device, device_parse  # device_parse: SSE/JSON parsing on the GPU too (quorum_b200.sse)
        self.all_content = ["" for _ in range(n_backends)]          # oai_proxy.py:552
        self._content = ["" for _ in range(n_backends)]
        self._ended = [False] * n_backends                          # stream ended by an exception: contributes nothing
        self._filters = (ThinkingTagFilterBatch(strategy.thinking_tags, n_backends, device)
                         if strategy.hide_intermediate_think else None)  # oai_proxy.py:543-545

    def abandon_backend(self, i: int) -> None:
        """The backend's chunk iterator raised (oai_proxy.py:660-673): nothing more is read from it and its content never
        reaches `all_content`."""
        self._ended[i] = True

    def feed_http_chunks(self, chunks: Sequence[Optional[bytes]]) -> List[List[str]]:
        """chunks[i] = the next HTTP chunk of backend i (None = nothing arrived).  Returns, per backend, the filtered
        deltas to re-emit as SSE events (empty strings are not emitted, oai_proxy.py:625-628)."""
        live = [c if (c is not None and not self._ended[i]) else None for i, c in enumerate(chunks)]
        if self.device_parse:
            from . import sse
            present = [i for i, c in enumerate(live) if c is not None]
            parsed = sse.parse_http_chunks_ex([live[i] for i in present], self.device)
            results = [([], False) for _ in live]
            for i, r in zip(present, parsed):
                results[i] = r
        else:
            results = [parse_sse_events(c) if c is not None else ([], False) for c in live]
        deltas = [r[0] or [] for r in results]
        if self._filters is not None:
            safe = self._filters.feed_chunks(deltas)               # oai_proxy.py:618-621
        else:
            safe = deltas                                          # oai_proxy.py:622-623
        emit: List[List[str]] = []
        for i in range(self.n):
            self._content[i] += "".join(safe[i])                   # oai_proxy.py:624
            emit.append([] if self.strategy.suppress_individual_responses else [t for t in safe[i] if t])
            if results[i][1]:
                self._ended[i] = True
        return emit

    def feed_http_chunks_events(self, chunks: Sequence[Optional[bytes]], created: int) -> List[List[bytes]]:
        """feed_http_chunks with every returned delta rendered as the reference's SSE event (oai_proxy.py:629-652)."""
        return [[delta_event(i, t, created) for t in ts] for i, ts in enumerate(self.feed_http_chunks(chunks))]

    def finish_backend(self, i: int) -> None:
        """oai_proxy.py:739-741 — flush() always returns "" so the content is what the feeds returned; a stream that
        ended in an exception never gets here in the reference."""
        if not self._ended[i]:
            self.all_content[i] = self._content[i]

    def finish(self) -> Optional[str]:
        """Final combined text, or None when there is no final event (skip_final_aggregation, or no content at all)."""
        if self._filters is not None:
            self._filters.flush()
        for i in range(self.n):
            self.finish_backend(i)
        return self.strategy.combine_stream_final(self.all_content)

    def finish_events(self, created: int) -> List[bytes]:
        """finish() rendered: final event or error event (none with skip_final_aggregation), then `[DONE]`."""
        combined = self.finish()
        return tail_events(self.strategy, self.all_content, created, combined)
