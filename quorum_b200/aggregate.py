"""Aggregate-strategy prompt assembly and usage sums on the GPU.

Mirrors what the reference does between the backends' answers and the aggregator call: filter every answer
(`strip_thinking_tags(content, thinking_tags, hide_intermediate=hide_final_think)`, src/quorum/oai_proxy.py:1193-1199), then build
    query_format.format(query=user_query) + prompt_template.replace("{responses}", separator.join(label_i + answer_i))
(:406-423, configuration keys and defaults :1204-1268).  The per-request constants (formatted query, template text around
the placeholder, labels) are rendered on the host with the same str methods; the filtering, the lengths, the prefix sum
and every byte of the answers move on the device in one batched pass.  The aggregator call itself stays where it is.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence

import torch

from . import batch as _b
from ._lib import check, lib
from .filter import _dec, _device, _enc, _to_device

DEFAULT_TEMPLATE = ("You have received the following responses regarding the user's query:\n\n{responses}\n\n"
                    "Provide a concise synthesis of these responses.")


def pack_heads(heads: Sequence[bytes], dev):
    """The per-request heads as device buffers (bytes, offsets, total bytes), for assemble_prompts_device."""
    hb, ho = _b.pack_host(list(heads))
    return _to_device(hb if hb.numel() else torch.zeros(1, dtype=torch.uint8), dev), _to_device(ho, dev), int(ho[-1])


def assemble_prompts_device(resp_bytes: torch.Tensor, resp_off: torch.Tensor, req_off: torch.Tensor, heads,
                            labels: Sequence[bytes], sep: bytes, tail: bytes):
    """Device buffers in, device buffers out: prompt q is out[out_off[q]:out_off[q+1]]."""
    dev = resp_off.device
    n_resp, n_req = resp_off.numel() - 1, req_off.numel() - 1

    def dbytes(b: bytes):
        return torch.frombuffer(bytearray(b + b"\0"), dtype=torch.uint8).to(dev)

    head_b = head_o = None
    head_total = 0
    if isinstance(heads, tuple):  # already on the device: (bytes, offsets, total) from pack_heads
        head_b, head_o, head_total = heads
    elif heads is not None:
        head_b, head_o, head_total = pack_heads(heads, dev)
    lo = [0]
    for lab in labels:
        lo.append(lo[-1] + len(lab))
    label_b, label_o = dbytes(b"".join(labels)), torch.tensor(lo, dtype=torch.int32).to(dev)
    sep_t, tail_t = dbytes(sep), dbytes(tail)
    cap = int(resp_bytes.numel()) + head_total + n_req * len(tail) + n_resp * (len(sep) + max((len(x) for x in labels), default=0)) + 16
    with torch.cuda.device(dev):
        out = torch.empty(cap, dtype=torch.uint8, device=dev)
        out_off = torch.empty(n_req + 1, dtype=torch.int64, device=dev)
        scratch = torch.empty(lib().qb200_assemble_scratch_bytes(n_resp, n_req), dtype=torch.uint8, device=dev)
        check(lib().qb200_assemble_prompts_batch(
            _b._ptr(resp_bytes), _b._ptr(resp_off), n_resp, _b._ptr(req_off), n_req, _b._ptr(head_b), _b._ptr(head_o), _b._ptr(label_b),
            _b._ptr(label_o), len(labels), _b._ptr(sep_t), len(sep), _b._ptr(tail_t), len(tail), _b._ptr(out), cap, _b._ptr(out_off),
            _b._ptr(scratch), scratch.numel(), _b._stream_ptr(dev)))
    return out, out_off


def usage_sums_device(usage: torch.Tensor, req_off: torch.Tensor) -> torch.Tensor:
    """usage: int64 [n_resp, 3] on the device; returns int64 [n_req, 3] (oai_proxy.py:1298-1311 per request)."""
    dev = usage.device
    n_req = req_off.numel() - 1
    with torch.cuda.device(dev):
        out = torch.empty((n_req, 3), dtype=torch.int64, device=dev)
        check(lib().qb200_usage_sums_batch(_b._ptr(usage.contiguous()), _b._ptr(req_off), n_req, _b._ptr(out), _b._stream_ptr(dev)))
    return out


@dataclass
class AggregatePromptBuilder:
    """Same configuration keys and defaults as the reference's aggregate strategy (oai_proxy.py:1204-1268).  The answers
    are filtered before they reach the prompt exactly as the reference does it (oai_proxy.py:1166-1199): with
    strip_thinking_tags(content, thinking_tags, hide_intermediate=hide_final_think), both keys read from the block of the
    strategy named by iterations.aggregation.strategy, hide_final_think defaulting to False."""
    separator: str = "\n\n---\n\n"
    include_original_query: bool = True
    query_format: str = "Original query: {query}\n\n"
    include_source_names: bool = False
    source_label_format: str = "Response from {backend_name}:\n"
    prompt_template: str = DEFAULT_TEMPLATE
    thinking_tags: List[str] = field(default_factory=lambda: ["think", "reason", "reasoning", "thought"])
    hide_final_think: bool = False
    device: Any = None

    @classmethod
    def from_config(cls, config: Dict[str, Any], device=None) -> "AggregatePromptBuilder":
        agg = config.get("strategy", {}).get("aggregate", {})
        tmpl = agg.get("prompt_template", DEFAULT_TEMPLATE)
        if "{intermediate_results}" in tmpl:  # :1244-1247
            tmpl = tmpl.replace("{intermediate_results}", "{responses}")
        name = config.get("iterations", {}).get("aggregation", {}).get("strategy", "concatenate")   # :1166-1173
        it = config.get("strategy", {}).get(name, {})
        return cls(separator=agg.get("intermediate_separator", "\n\n---\n\n"),
                   include_original_query=agg.get("include_original_query", True),
                   query_format=agg.get("query_format", "Original query: {query}\n\n"),
                   include_source_names=agg.get("include_source_names", False),
                   source_label_format=agg.get("source_label_format", "Response from {backend_name}:\n"),
                   prompt_template=tmpl,
                   thinking_tags=it.get("thinking_tags", ["think", "reason", "reasoning", "thought"]),
                   hide_final_think=it.get("hide_final_think", False), device=device)

    def build_prompts(self, contents_per_request: Sequence[Sequence[str]], user_queries: Sequence[str]) -> List[str]:
        """One aggregator prompt per request.  contents_per_request[q] are the backends' raw answers of request q."""
        n_req = len(contents_per_request)
        if n_req == 0:
            return []
        dev = _device(self.device)
        flat = [c for cs in contents_per_request for c in cs]
        for c in flat:
            if not isinstance(c, str):
                raise TypeError("expected string or bytes-like object, got %r" % type(c).__name__)
        heads = [self.query_format.format(query=q) if self.include_original_query else "" for q in user_queries]
        parts = self.prompt_template.split("{responses}")
        max_b = max((len(cs) for cs in contents_per_request), default=0)
        labels = [self.source_label_format.format(backend_name=f"LLM{i + 1}") for i in range(max_b)] if self.include_source_names else []
        if len(parts) == 1:  # no placeholder: the answers do not appear in the prompt
            return [h + self.prompt_template for h in heads]
        # filter every answer (one answer per "request": no join yet), stays on the device
        data, off = _b.pack_host([_enc(c) for c in flat])
        d_data, d_off = _to_device(data if data.numel() else torch.zeros(16, dtype=torch.uint8), dev), _to_device(off, dev)
        table = _b.tag_table(self.thinking_tags)
        f_bytes, f_off, _ = _b.strip_concat_batch(table, d_data, d_off, None, None, hide=self.hide_final_think, streaming=False)
        ro = [0]
        for cs in contents_per_request:
            ro.append(ro[-1] + len(cs))
        d_ro = torch.tensor(ro, dtype=torch.int64).to(dev)
        if len(parts) == 2:
            out, out_off = assemble_prompts_device(f_bytes, f_off, d_ro, [(_enc(h + parts[0])) for h in heads], [_enc(x) for x in labels],
                                                   _enc(self.separator), _enc(parts[1]))
            return [_dec(x) for x in _b.split_host(out, out_off)]
        # the placeholder occurs more than once: the join is built once on the device and substituted by str.replace
        out, out_off = assemble_prompts_device(f_bytes, f_off, d_ro, None, [_enc(x) for x in labels], _enc(self.separator), b"")
        joined = [_dec(x) for x in _b.split_host(out, out_off)]
        return [h + self.prompt_template.replace("{responses}", j) for h, j in zip(heads, joined)]
