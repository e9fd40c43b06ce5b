"""The concatenate aggregation strategy on the GPU.

The reference has no function for it: the join is written inline twice,
    non-streaming   processed = [strip_thinking_tags(c, tags, hide_final_think) ...]; separator.join(processed)
                    (src/quorum/oai_proxy.py:1194-1199, 1286; fallbacks 1279, 1284)
    streaming final [strip_thinking_tags(t, ...) for t in all_content if t]; f"\\n{separator}".join(...)
                    (src/quorum/oai_proxy.py:759-764, 841; fallbacks 834, 839)
Here both are one fused pass: filter + strip + prefix sum over kept lengths and separators + copy to the final offsets.
Config keys keep the reference's names (oai_proxy.py:1058-1071, 1175-1185).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import torch

from . import batch as _b
from .filter import _dec, _device, _enc, _to_device

DEFAULT_THINKING_TAGS = ["think", "reason", "reasoning", "thought"]   # oai_proxy.py:1063-1065, 1180-1182


def concatenate_many(requests: Sequence[Sequence[str]], separator: str, *, tags: Sequence[str],
                     hide_final_think: bool, streaming: bool, device=None) -> List[str]:
    """Join every request's contents.  streaming=False: empties kept, `separator` between items;
    streaming=True: contents that are empty before filtering are dropped and the joiner is "\\n"+separator."""
    for r in requests:
        for c in r:
            if not isinstance(c, str):   # re.sub / str.join both raise TypeError in the reference
                raise TypeError(f"expected str instance, {type(c).__name__} found")
    if not requests:
        return []
    dev = _device(device)
    flat = [_enc(c) for r in requests for c in r]
    roff = [0]
    for r in requests:
        roff.append(roff[-1] + len(r))
    data, off = _b.pack_host(flat)
    sep_b = _enc(("\n" + separator) if streaming else separator)
    sep_t = torch.frombuffer(bytearray(sep_b), dtype=torch.uint8).to(dev) if sep_b else None
    out, out_off, _ = _b.strip_concat_batch(
        _b.tag_table(list(tags)), _to_device(data, dev), _to_device(off, dev),
        torch.tensor(roff, dtype=torch.int64).to(dev), sep_t, hide=hide_final_think, streaming=streaming)
    return [_dec(x) for x in _b.split_host(out, out_off)]


def concatenate(contents: Sequence[str], separator: str, *, tags: Sequence[str] = tuple(DEFAULT_THINKING_TAGS),
                hide_final_think: bool = False, streaming: bool = False) -> str:
    """One request: the reference's inline join (see module docstring)."""
    return concatenate_many([list(contents)], separator, tags=tags, hide_final_think=hide_final_think,
                            streaming=streaming)[0]


@dataclass
class ConcatenateStrategy:
    """The `strategy.concatenate` config block with the reference's keys and defaults
    (oai_proxy.py:1058-1071 streaming, 1175-1185 non-streaming; config.yaml:29-40)."""
    separator: str = "\n"
    hide_intermediate_think: bool = True
    hide_final_think: bool = False
    thinking_tags: List[str] = field(default_factory=lambda: list(DEFAULT_THINKING_TAGS))
    skip_final_aggregation: bool = False
    suppress_individual_responses: bool = False

    @classmethod
    def from_config(cls, config: dict, strategy_name: Optional[str] = None,
                    request_body: Optional[dict] = None) -> "ConcatenateStrategy":
        """Reads config["strategy"][name] like oai_proxy.py:1049-1071 (name defaults to
        config["iterations"]["aggregation"]["strategy"], itself defaulting to "concatenate").  `request_body` is the
        parsed request: its "suppress_individual_responses" key, when present, overrides the config (oai_proxy.py:1072-1075)."""
        if strategy_name is None:
            strategy_name = config.get("iterations", {}).get("aggregation", {}).get("strategy", "concatenate")
        c = config.get("strategy", {}).get(strategy_name, {})
        if request_body is not None and "suppress_individual_responses" in request_body:
            c = dict(c, suppress_individual_responses=request_body.get("suppress_individual_responses"))
        return cls(
            separator=c.get("separator", "\n"),
            hide_intermediate_think=c.get("hide_intermediate_think", True),
            hide_final_think=c.get("hide_final_think", False),
            thinking_tags=c.get("thinking_tags", list(DEFAULT_THINKING_TAGS)),
            skip_final_aggregation=c.get("skip_final_aggregation", False),
            suppress_individual_responses=c.get("suppress_individual_responses", False),
        )

    def combine(self, contents: Sequence[str]) -> str:
        """Non-streaming path (oai_proxy.py:1194-1199 + 1286).  skip_final_aggregation is not consulted there."""
        return concatenate(contents, self.separator, tags=self.thinking_tags, hide_final_think=self.hide_final_think,
                           streaming=False)

    def combine_many(self, requests: Sequence[Sequence[str]]) -> List[str]:
        return concatenate_many(requests, self.separator, tags=self.thinking_tags,
                                hide_final_think=self.hide_final_think, streaming=False)

    def combine_stream_final(self, all_content: Sequence[str]) -> Optional[str]:
        """Streaming final event (oai_proxy.py:759-764 + 841).  None when skip_final_aggregation is set (no final event)
        or when no backend produced content (the reference then emits its error event, oai_proxy.py:863-881)."""
        if self.skip_final_aggregation or not any(all_content):
            return None
        return concatenate(all_content, self.separator, tags=self.thinking_tags,
                           hide_final_think=self.hide_final_think, streaming=True)
