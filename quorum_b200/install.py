"""Patch point: rebinding two module attributes is the reference-native way to replace the hot path
(its own tests rebind `quorum.oai_proxy.config` the same way, tests/test_parallel_backends.py:177)."""
from __future__ import annotations

from .filter import ThinkingTagFilter, strip_thinking_tags


def install(module=None):
    """Rebind `strip_thinking_tags` and `ThinkingTagFilter` in `quorum.oai_proxy` (or the module given) to the GPU
    implementations.  Returns the previous pair so a caller can restore them."""
    if module is None:
        import quorum.oai_proxy as module  # noqa: PLC0415
    previous = (module.strip_thinking_tags, module.ThinkingTagFilter)
    module.strip_thinking_tags = strip_thinking_tags
    module.ThinkingTagFilter = ThinkingTagFilter
    return previous


def uninstall(previous, module=None):
    if module is None:
        import quorum.oai_proxy as module  # noqa: PLC0415
    module.strip_thinking_tags, module.ThinkingTagFilter = previous
