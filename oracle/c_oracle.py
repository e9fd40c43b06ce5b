"""ctypes binding of oracle/quorum_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Importable only from tests/, __graft_entry__.smoke() and bench.py's CPU legs.  Builds the shared
library with `make -C oracle` on first use when it is missing (gcc only).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "libquorum_oracle.so")

FLAG_HIDE, FLAG_STREAM_JOIN, FLAG_FLUSH = 1, 2, 4
HELD_MAX = 23

STATE_DTYPE = np.dtype([("depth", "<i4"), ("tail_len", "<i4"), ("held_len", "u1"), ("held", "u1", (HELD_MAX,))])
assert STATE_DTYPE.itemsize == 32


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "quorum_oracle.c")
    if force or not os.path.isfile(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "-s"] + (["-B"] if force else []))
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build())
        vp, i64, i32, u32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_uint32
        L.qo_tags_init.argtypes = [vp, vp, vp, i32]
        L.qo_tags_init.restype = ctypes.c_int
        L.qo_strip.argtypes = [vp, vp, i64, ctypes.c_int, vp]
        L.qo_strip.restype = i64
        L.qo_filter_feed.argtypes = [vp, vp, vp, i64, vp]
        L.qo_filter_feed.restype = i64
        L.qo_filter_flush.argtypes = [vp]
        L.qo_filter_flush.restype = None
        L.qo_strip_concat_batch.argtypes = [vp, vp, vp, i64, vp, i64, vp, i32, u32, vp, i64, vp, vp]
        L.qo_strip_concat_batch.restype = ctypes.c_int
        L.qo_stream_feed_batch.argtypes = [vp, vp, vp, i64, vp, i64, vp, vp, u32, vp, i64, vp]
        L.qo_stream_feed_batch.restype = ctypes.c_int
        assert L.qo_sizeof_state() == STATE_DTYPE.itemsize
        _lib = L
    return _lib


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data


def pack_tags(tags: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    bs = [t.encode("ascii") for t in tags]
    off = np.zeros(len(bs) + 1, dtype=np.int32)
    off[1:] = np.cumsum([len(b) for b in bs])
    names = np.frombuffer(b"".join(bs) + b"\0", dtype=np.uint8).copy()
    return names, off


class Tags:
    def __init__(self, tags: Sequence[str]):
        L = lib()
        self._buf = ctypes.create_string_buffer(L.qo_sizeof_tags())
        names, off = pack_tags(tags)
        rc = L.qo_tags_init(ctypes.addressof(self._buf), _ptr(names), _ptr(off), len(tags))
        if rc != 0:
            raise ValueError(f"qo_tags_init failed: {rc}")

    @property
    def ptr(self):
        return ctypes.addressof(self._buf)


def strip_bytes(content: bytes, tags: Tags, hide: bool = True) -> bytes:
    out = ctypes.create_string_buffer(max(len(content), 1))
    n = lib().qo_strip(tags.ptr, content, len(content), int(hide), out)
    return out.raw[:n]


class FilterBytes:
    def __init__(self, tags: Tags):
        self.tags = tags
        self.state = np.zeros(1, dtype=STATE_DTYPE)
        self.state["tail_len"] = -1

    def feed(self, text: bytes) -> bytes:
        out = ctypes.create_string_buffer(len(text) + HELD_MAX + 1)
        n = lib().qo_filter_feed(self.tags.ptr, _ptr(self.state), text, len(text), out)
        return out.raw[:n]

    def flush(self) -> bytes:
        lib().qo_filter_flush(_ptr(self.state))
        return b""

    @property
    def depth(self) -> int:
        return int(self.state["depth"][0])

    @property
    def held(self) -> bytes:
        return bytes(self.state["held"][0][: int(self.state["held_len"][0])])


def strip_concat_batch(tags: Tags, in_bytes: np.ndarray, in_off: np.ndarray, req_off: Optional[np.ndarray],
                       sep: bytes, flags: int):
    """Returns (out_bytes[:total], out_off[n_req+1], resp_out_len[n_resp])."""
    n_resp = len(in_off) - 1
    n_req = n_resp if req_off is None else len(req_off) - 1
    cap = int(in_off[-1]) + len(sep) * max(n_resp, 1) + 16
    out = np.empty(cap, dtype=np.uint8)
    out_off = np.empty(n_req + 1, dtype=np.int64)
    rlen = np.empty(max(n_resp, 1), dtype=np.int64)
    sep_a = np.frombuffer(sep + b"\0", dtype=np.uint8)
    rc = lib().qo_strip_concat_batch(tags.ptr, _ptr(in_bytes), _ptr(in_off), n_resp, _ptr(req_off), n_req,
                                     _ptr(sep_a), len(sep), flags, _ptr(out), cap, _ptr(out_off), _ptr(rlen))
    if rc != 0:
        raise RuntimeError(f"qo_strip_concat_batch failed: {rc}")
    return out[: int(out_off[-1])], out_off, rlen[:n_resp]


def stream_feed_batch(tags: Tags, chunk_bytes: np.ndarray, chunk_off: np.ndarray, stream_off: np.ndarray,
                      state_in: Optional[np.ndarray], flags: int):
    """Returns (out_bytes[:total], out_chunk_off[n_chunks+1], state_out[n_streams])."""
    n_chunks = len(chunk_off) - 1
    n_streams = len(stream_off) - 1
    cap = int(chunk_off[-1]) + HELD_MAX * n_streams + 16
    out = np.empty(cap, dtype=np.uint8)
    oco = np.empty(n_chunks + 1, dtype=np.int64)
    state_out = np.zeros(max(n_streams, 1), dtype=STATE_DTYPE)
    rc = lib().qo_stream_feed_batch(tags.ptr, _ptr(chunk_bytes), _ptr(chunk_off), n_chunks, _ptr(stream_off), n_streams,
                                    _ptr(state_in), _ptr(state_out), flags, _ptr(out), cap, _ptr(oco))
    if rc != 0:
        raise RuntimeError(f"qo_stream_feed_batch failed: {rc}")
    return out[: int(oco[-1])], oco, state_out[:n_streams]


def json_escape(content: bytes) -> bytes:
    """Body of json.dumps(str) (ensure_ascii) for UTF-8 bytes, via the C restatement."""
    L = lib()
    if not hasattr(L, "_qo_json_ready"):
        L.qo_json_escape.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        L.qo_json_escape.restype = ctypes.c_int64
        L.qo_sse_frame_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int, ctypes.c_void_p,
                                         ctypes.c_int64, ctypes.c_void_p]
        L.qo_sse_frame_batch.restype = ctypes.c_int
        L._qo_json_ready = True
    out = ctypes.create_string_buffer(6 * len(content) + 1)
    n = L.qo_json_escape(content, len(content), out)
    return out.raw[:n]


def sse_frame_batch(in_bytes: np.ndarray, in_off: np.ndarray, item_prefix: Optional[np.ndarray], prefixes: Sequence[bytes],
                    suffix: bytes, skip_empty: bool):
    """Returns (out_bytes[:total], out_off[n+1])."""
    json_escape(b"")  # binds the functions
    n = len(in_off) - 1
    pb = np.frombuffer(b"".join(prefixes) + b"\0", dtype=np.uint8)
    po = np.zeros(len(prefixes) + 1, dtype=np.int64)
    po[1:] = np.cumsum([len(p) for p in prefixes])
    cap = 6 * int(in_off[-1]) + n * (max(len(p) for p in prefixes) + len(suffix)) + 16
    out = np.empty(cap, dtype=np.uint8)
    out_off = np.empty(n + 1, dtype=np.int64)
    sa = np.frombuffer(suffix + b"\0", dtype=np.uint8)
    rc = lib().qo_sse_frame_batch(_ptr(in_bytes), _ptr(in_off), n, _ptr(item_prefix), _ptr(pb), _ptr(po), _ptr(sa), len(suffix),
                                  int(skip_empty), _ptr(out), cap, _ptr(out_off))
    if rc != 0:
        raise RuntimeError(f"qo_sse_frame_batch failed: {rc}")
    return out[: int(out_off[-1])], out_off
