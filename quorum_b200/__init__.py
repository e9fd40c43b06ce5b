"""quorum_b200 — the thinking-tag filter and concatenate strategy of andrewginns/quorum as hand-written sm_100a CUDA.

Reference-shaped surface (drop-in for src/quorum/oai_proxy.py):
    strip_thinking_tags, ThinkingTagFilter, ConcatenateStrategy / concatenate, install()
Batched surface (what a batching front end calls; torch CUDA tensors through the C ABI of include/quorum_b200.h):
    batch.strip_concat_batch, batch.stream_feed_batch, batch.stream_flush_batch
There is no CPU implementation in this package: without libquorum_b200.so and a CUDA device every call raises.
"""
from ._lib import QuorumB200Error, launch_count  # noqa: F401
from .concatenate import ConcatenateStrategy, concatenate, concatenate_many  # noqa: F401
from .filter import (ThinkingTagFilter, ThinkingTagFilterBatch, strip_thinking_tags,  # noqa: F401
                     strip_thinking_tags_many)
from .install import install, uninstall  # noqa: F401
from .streaming import ConcatenateStreamSession, parse_sse_chunk, parse_sse_events  # noqa: F401
from . import aggregate, batch, sharding, sse  # noqa: F401
from .aggregate import AggregatePromptBuilder  # noqa: F401

__all__ = [
    "strip_thinking_tags", "strip_thinking_tags_many", "ThinkingTagFilter", "ThinkingTagFilterBatch",
    "ConcatenateStrategy", "concatenate", "concatenate_many", "ConcatenateStreamSession", "parse_sse_chunk", "parse_sse_events",
    "install", "uninstall", "batch", "sharding", "sse", "aggregate", "AggregatePromptBuilder", "QuorumB200Error", "launch_count",
]
