"""Multi-GPU partitioning: requests / streams are independent, so ranks take contiguous,
byte-balanced ranges of units and never exchange data; the host concatenates the per-rank outputs."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def partition_by_bytes(unit_bytes: Sequence[int], world_size: int) -> List[Tuple[int, int]]:
    """Contiguous [lo, hi) unit ranges per rank, balanced by byte count.  A unit (one request with all its backend
    responses, or one stream with all its chunks) is never split across ranks."""
    n = len(unit_bytes)
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    csum = np.concatenate([[0], np.cumsum(np.asarray(unit_bytes, dtype=np.int64))])
    total = int(csum[-1])
    cuts = [0]
    for r in range(1, world_size):
        target = total * r // world_size
        k = int(np.searchsorted(csum, target, side="left"))
        k = min(max(k, cuts[-1]), n)
        cuts.append(k)
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(world_size)]


def shard_offsets(offsets: np.ndarray, group_off: np.ndarray, lo: int, hi: int):
    """Slice a two-level ragged layout (groups of items over a flat byte buffer) to groups [lo, hi).
    Returns (byte_lo, byte_hi, item_off_local, group_off_local)."""
    i0, i1 = int(group_off[lo]), int(group_off[hi])
    b0, b1 = int(offsets[i0]), int(offsets[i1])
    return b0, b1, offsets[i0:i1 + 1] - b0, group_off[lo:hi + 1] - i0


# ---- product entry points: one host batch over several GPUs --------------------------------------------------------------

def _devices(devices):
    import torch
    if devices is None:
        n = torch.cuda.device_count()
        if n == 0:
            from ._lib import QuorumB200Error
            raise QuorumB200Error("quorum_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        devices = list(range(n))
    return [torch.device("cuda", d) if isinstance(d, int) else torch.device(d) for d in devices]


def _run_parts(jobs):
    """Run one callable per shard, each on its own host thread (the CUDA calls release the GIL), re-raise the first error."""
    import threading
    results, errors = [None] * len(jobs), []

    def work(k):
        try:
            results[k] = jobs[k]()
        except BaseException as e:  # noqa: BLE001 - re-raised below on the calling thread
            errors.append(e)
    threads = [threading.Thread(target=work, args=(k,)) for k in range(len(jobs))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return results


def strip_concat_sharded(tags, h_data, h_in_off, h_req_off, sep: bytes, *, hide: bool, streaming: bool, devices=None):
    """strip_thinking_tags + concatenate of ONE host batch over several GPUs (BASELINE.json:
    "independent requests shard across the GPUs by batch index with no NCCL collective, only the host gathers").

    h_data uint8[total], h_in_off int64[n_resp+1], h_req_off int64[n_req+1]: host tensors (pinned for speed).  Requests are
    cut into contiguous, byte-balanced ranges (partition_by_bytes), range k runs on devices[k] through the host-buffer
    session batch.HostStripConcat (upload, kernels and download pipelined), the host concatenates the outputs.
    Returns (out_bytes uint8[*], out_off int64[n_req+1]) on the host; request q is out_bytes[out_off[q]:out_off[q+1]].
    `devices` may name a device more than once (its shards then run one after the other on that device's streams)."""
    import torch
    from . import batch as _b
    devs = _devices(devices)
    io = h_in_off.numpy() if hasattr(h_in_off, "numpy") else np.asarray(h_in_off)
    ro = h_req_off.numpy() if hasattr(h_req_off, "numpy") else np.asarray(h_req_off)
    n_req = len(ro) - 1
    req_bytes = io[ro[1:]] - io[ro[:-1]]
    parts = partition_by_bytes(req_bytes, len(devs))
    table = tags if isinstance(tags, _b.TagTable) else _b.tag_table(list(tags))
    pin = torch.cuda.is_available()

    def job(k):
        lo, hi = parts[k]
        b0, b1, off_l, req_l = shard_offsets(io, ro, lo, hi)
        if hi == lo:
            return np.zeros(0, dtype=np.uint8), np.zeros(1, dtype=np.int64)
        dev = devs[k]
        with torch.cuda.device(dev):
            sess = _b.HostStripConcat(table, dev, max(b1 - b0, 16), len(off_l) - 1, hi - lo, sep)
            h_io = torch.from_numpy(np.ascontiguousarray(off_l))
            h_ro = torch.from_numpy(np.ascontiguousarray(req_l))
            if pin:
                h_io, h_ro = h_io.pin_memory(), h_ro.pin_memory()
            sess.run(h_data[b0:b1], h_io, h_ro, hide=hide, streaming=streaming)
            torch.cuda.synchronize(dev)
            oo = sess.h_oo[: hi - lo + 1].numpy().copy()
            return sess.h_out[: int(oo[-1])].numpy().copy(), oo
    res = _run_parts([lambda k=k: job(k) for k in range(len(devs))])
    out_off = np.zeros(n_req + 1, dtype=np.int64)
    base = 0
    for (lo, hi), (ob, oo) in zip(parts, res):
        out_off[lo + 1:hi + 1] = base + oo[1:]
        base += int(oo[-1])
    out = np.concatenate([ob for ob, _ in res]) if res else np.zeros(0, dtype=np.uint8)
    return torch.from_numpy(out), torch.from_numpy(out_off)


def stream_feed_sharded(tags, h_data, h_chunk_off, h_stream_off, *, flush: bool, devices=None):
    """ThinkingTagFilter.feed of ONE host batch of streams over several GPUs: streams are cut into contiguous,
    byte-balanced ranges, range k runs on devices[k] through batch.HostStreamFeeder, the host concatenates.
    Returns (out_bytes uint8[*], out_chunk_off int64[n_chunks+1], state uint8[n_streams, 32]) on the host."""
    import torch
    from . import batch as _b
    from ._lib import STATE_BYTES
    devs = _devices(devices)
    co = h_chunk_off.numpy() if hasattr(h_chunk_off, "numpy") else np.asarray(h_chunk_off)
    so = h_stream_off.numpy() if hasattr(h_stream_off, "numpy") else np.asarray(h_stream_off)
    n_streams, n_chunks = len(so) - 1, len(co) - 1
    stream_bytes = co[so[1:]] - co[so[:-1]]
    parts = partition_by_bytes(stream_bytes, len(devs))
    table = tags if isinstance(tags, _b.TagTable) else _b.tag_table(list(tags))
    pin = torch.cuda.is_available()

    def job(k):
        lo, hi = parts[k]
        b0, b1, off_l, str_l = shard_offsets(co, so, lo, hi)
        if hi == lo:
            return np.zeros(0, dtype=np.uint8), np.zeros(1, dtype=np.int64), np.zeros((0, STATE_BYTES), dtype=np.uint8)
        dev = devs[k]
        with torch.cuda.device(dev):
            sess = _b.HostStreamFeeder(table, dev, max(b1 - b0, 16), len(off_l) - 1, hi - lo)
            h_co = torch.from_numpy(np.ascontiguousarray(off_l))
            h_so = torch.from_numpy(np.ascontiguousarray(str_l))
            if pin:
                h_co, h_so = h_co.pin_memory(), h_so.pin_memory()
            sess.run(h_data[b0:b1], h_co, h_so, flush=flush)
            torch.cuda.synchronize(dev)
            oo = sess.h_oco[: len(off_l)].numpy().copy()
            return sess.h_out[: int(oo[-1])].numpy().copy(), oo, sess.d_state[: hi - lo].cpu().numpy()
    res = _run_parts([lambda k=k: job(k) for k in range(len(devs))])
    out_off = np.zeros(n_chunks + 1, dtype=np.int64)
    base = 0
    for (lo, hi), (ob, oo, _) in zip(parts, res):
        c0, c1 = int(so[lo]), int(so[hi])
        out_off[c0 + 1:c1 + 1] = base + oo[1:]
        base += int(oo[-1])
    out = np.concatenate([r[0] for r in res]) if res else np.zeros(0, dtype=np.uint8)
    state = np.concatenate([r[2] for r in res]) if res else np.zeros((0, STATE_BYTES), dtype=np.uint8)
    return torch.from_numpy(out), torch.from_numpy(out_off), torch.from_numpy(state)
