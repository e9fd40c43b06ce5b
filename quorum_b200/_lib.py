"""ctypes binding of the C ABI declared in include/quorum_b200.h.  Fails loudly when the CUDA library is missing:
there is no CPU path in this package."""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libquorum_b200.so")

QB200_HIDE, QB200_STREAM_JOIN, QB200_FLUSH, QB200_CHAIN, QB200_SKIP_EMPTY = 1, 2, 4, 8, 16
HELD_MAX = 23
STATE_BYTES = 32

EXPORTS = [
    "qb200_last_error", "qb200_tags_create", "qb200_tags_destroy", "qb200_strip_concat_scratch_bytes",
    "qb200_strip_concat_batch", "qb200_stream_feed_scratch_bytes", "qb200_stream_feed_batch",
    "qb200_stream_flush_batch", "qb200_launch_count", "qb200_stream_feed_batch_noalloc",
    "qb200_sse_frame_scratch_bytes", "qb200_sse_frame_batch", "qb200_profile_begin", "qb200_profile_end",
    "qb200_assemble_scratch_bytes", "qb200_assemble_prompts_batch", "qb200_usage_sums_batch",
    "qb200_sse_split_scratch_bytes", "qb200_sse_split_batch", "qb200_sse_parse_scratch_bytes", "qb200_sse_parse_batch",
    "qb200_offsets_scratch_bytes", "qb200_offsets_from_lengths32", "qb200_lengths32_from_offsets",
]
NOALLOC_MAX_STREAMS = 65536

_lib = None


class QuorumB200Error(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise QuorumB200Error(
            f"{LIB_PATH} is missing: build it with `python -m quorum_b200.build` (nvcc, sm_100a). "
            "quorum_b200 has no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i64, i32, u32, sz = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_uint32, ctypes.c_size_t
    L.qb200_last_error.restype = ctypes.c_char_p
    L.qb200_last_error.argtypes = []
    L.qb200_tags_create.restype = ctypes.c_int
    L.qb200_tags_create.argtypes = [vp, vp, i32, ctypes.POINTER(vp)]
    L.qb200_tags_destroy.restype = None
    L.qb200_tags_destroy.argtypes = [vp]
    L.qb200_strip_concat_scratch_bytes.restype = sz
    L.qb200_strip_concat_scratch_bytes.argtypes = [i64, i64, i64]
    L.qb200_strip_concat_batch.restype = ctypes.c_int
    L.qb200_strip_concat_batch.argtypes = [vp, vp, vp, i64, vp, i64, vp, i32, u32, vp, i64, vp, vp, vp, sz, vp]
    L.qb200_stream_feed_scratch_bytes.restype = sz
    L.qb200_stream_feed_scratch_bytes.argtypes = [i64, i64, i64]
    L.qb200_stream_feed_batch.restype = ctypes.c_int
    L.qb200_stream_feed_batch.argtypes = [vp, vp, vp, i64, vp, i64, vp, vp, u32, vp, i64, vp, vp, sz, vp]
    L.qb200_stream_feed_batch_noalloc.restype = ctypes.c_int
    L.qb200_stream_feed_batch_noalloc.argtypes = [vp, vp, vp, i64, vp, i64, vp, vp, u32, vp, i64, vp, vp]
    L.qb200_sse_frame_scratch_bytes.restype = sz
    L.qb200_sse_frame_scratch_bytes.argtypes = [i64]
    L.qb200_sse_frame_batch.restype = ctypes.c_int
    L.qb200_sse_frame_batch.argtypes = [vp, vp, i64, vp, vp, vp, i32, vp, i32, u32, vp, i64, vp, vp, sz, vp]
    L.qb200_assemble_scratch_bytes.restype = sz
    L.qb200_assemble_scratch_bytes.argtypes = [i64, i64]
    L.qb200_assemble_prompts_batch.restype = ctypes.c_int
    L.qb200_assemble_prompts_batch.argtypes = [vp, vp, i64, vp, i64, vp, vp, vp, vp, i32, vp, i32, vp, i32, vp, i64, vp, vp, sz, vp]
    L.qb200_offsets_scratch_bytes.restype = sz
    L.qb200_offsets_scratch_bytes.argtypes = [i64]
    L.qb200_offsets_from_lengths32.restype = ctypes.c_int
    L.qb200_offsets_from_lengths32.argtypes = [vp, i64, vp, vp, sz, vp]
    L.qb200_lengths32_from_offsets.restype = ctypes.c_int
    L.qb200_lengths32_from_offsets.argtypes = [vp, i64, vp, vp]
    L.qb200_usage_sums_batch.restype = ctypes.c_int
    L.qb200_usage_sums_batch.argtypes = [vp, vp, i64, vp, vp]
    L.qb200_sse_split_scratch_bytes.restype = sz
    L.qb200_sse_split_scratch_bytes.argtypes = [i64]
    L.qb200_sse_split_batch.restype = ctypes.c_int
    L.qb200_sse_split_batch.argtypes = [vp, vp, i64, vp, vp, sz, vp]
    L.qb200_sse_parse_scratch_bytes.restype = sz
    L.qb200_sse_parse_scratch_bytes.argtypes = [i64]
    L.qb200_sse_parse_batch.restype = ctypes.c_int
    L.qb200_sse_parse_batch.argtypes = [vp, vp, i64, vp, i64, vp, i64, vp, vp, vp, vp, sz, vp]
    L.qb200_profile_begin.restype = ctypes.c_int
    L.qb200_profile_begin.argtypes = [vp]
    L.qb200_profile_end.restype = i32
    L.qb200_profile_end.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(ctypes.c_float), i32]
    L.qb200_stream_flush_batch.restype = ctypes.c_int
    L.qb200_stream_flush_batch.argtypes = [vp, i64, vp]
    L.qb200_launch_count.restype = i64
    L.qb200_launch_count.argtypes = []
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().qb200_last_error().decode("utf-8", "replace")
        if rc == -2:
            raise ValueError(f"quorum_b200: {msg}")
        raise QuorumB200Error(f"quorum_b200 error {rc}: {msg}")


def launch_count() -> int:
    return int(lib().qb200_launch_count())


def profile_begin(stream_ptr: int = 0) -> None:
    """Start recording one CUDA event per kernel launch of this library (see include/quorum_b200.h)."""
    check(lib().qb200_profile_begin(ctypes.c_void_p(stream_ptr)))


def profile_end(max_entries: int = 4096):
    """[(kernel name, milliseconds)] for the launches since profile_begin, in launch order."""
    names = (ctypes.c_char_p * max_entries)()
    ms = (ctypes.c_float * max_entries)()
    n = lib().qb200_profile_end(names, ms, max_entries)
    if n < 0:
        check(n)
    return [(names[i].decode(), float(ms[i])) for i in range(n)]
