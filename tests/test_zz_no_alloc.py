"""The allocation-free entry point qb200_stream_feed_batch_noalloc against qb200_stream_feed_batch.

Some hosts cannot give the library scratch memory at call time (all device memory already reserved by their own
pools, a binding that only owns its input/output buffers, or a loop where cudaMalloc's device-wide synchronisation is
not acceptable), so the library carries a fixed scratch array for them.  That array is shared by every call and keeps
the previous call's contents, which is why this test makes TWO calls in a row with different batches and requires both
to match the ordinary entry point bit for bit (this path is integer/byte work: the tolerance is zero).

Collected last on purpose (file name), so a run that stops at the first failure has already run everything else; it
passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _batch(seed, n_streams, dev):
    from quorum_b200 import synth
    base = synth.cfg2_base(32, chunks=64, chunk_bytes=64, seed=seed)
    idx = synth.gather_units(base, n_streams, seed=seed + 1)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    chunk_off = np.arange(0, data.size + 1, 64, dtype=np.int64)
    stream_off = np.arange(0, n_streams * 64 + 1, 64, dtype=np.int64)
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    return t(data), t(chunk_off), t(stream_off)


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    import quorum_b200 as qb
    from quorum_b200 import _lib, batch

    dev = torch.device("cuda", 0)
    table = batch.tag_table(["think", "reason", "reasoning", "thought", "Thought"])
    for call, (seed, n_streams) in enumerate([(41, 700), (42, 333)]):        # second batch is smaller and different:
        data, chunk_off, stream_off = _batch(seed, n_streams, dev)          # stale scratch from call 1 must not matter
        n_chunks = chunk_off.numel() - 1
        # reference result: the ordinary entry point with caller-owned scratch
        exp_out, exp_oco, exp_state = batch.stream_feed_batch(table, data, chunk_off, stream_off, None, flush=False)
        # every buffer of the no-alloc call is created here, by the caller, before the call
        out = torch.zeros(data.numel() + _lib.HELD_MAX * n_streams + 16, dtype=torch.uint8, device=dev)
        oco = torch.zeros(n_chunks + 1, dtype=torch.int64, device=dev)
        state = torch.zeros((n_streams, _lib.STATE_BYTES), dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        batch.stream_feed_batch_noalloc(table, data, chunk_off, stream_off, None, flush=False, state_out=state,
                                        out_bytes=out, out_chunk_off=oco)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before, "the wrapper must not allocate through torch either"
        total = int(exp_oco[-1].item())
        assert total > 0
        assert torch.equal(oco, exp_oco), f"call {call}: per-chunk offsets differ"
        assert torch.equal(out[:total], exp_out[:total]), f"call {call}: output bytes differ"
        assert torch.equal(state, exp_state), f"call {call}: carried state differs"


def _batch_few_chunks(seed, n_streams, dev):
    """Many short streams (4 chunks x 64 B each), so that batch sizes near the limit stay small."""
    from quorum_b200 import synth
    base = synth.cfg2_base(48, chunks=4, chunk_bytes=64, seed=seed)
    idx = synth.gather_units(base, n_streams, seed=seed + 1)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    chunk_off = np.arange(0, data.size + 1, 64, dtype=np.int64)
    stream_off = np.arange(0, n_streams * 4 + 1, 4, dtype=np.int64)
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    return t(data), t(chunk_off), t(stream_off)


def test_noalloc_at_the_scan_split_and_at_the_limit():
    """Two more calls in a row on the shared array: 8,193 streams (the first size at which the prefix sum over the streams
    takes three launches and keeps its partial sums in the array's last KiB) and 65,536 streams (the documented limit, the
    whole array in use); a small batch in between must not see what the large one left behind."""
    from quorum_b200 import _lib, batch

    dev = torch.device("cuda", 0)
    table = batch.tag_table(["think", "reason", "reasoning", "thought", "Thought"])
    for call, (seed, n_streams) in enumerate([(51, 8193), (52, _lib.NOALLOC_MAX_STREAMS), (53, 77), (54, 8192)]):
        data, chunk_off, stream_off = _batch_few_chunks(seed, n_streams, dev)
        n_chunks = chunk_off.numel() - 1
        exp_out, exp_oco, exp_state = batch.stream_feed_batch(table, data, chunk_off, stream_off, None, flush=True)
        out = torch.zeros(data.numel() + _lib.HELD_MAX * n_streams + 16, dtype=torch.uint8, device=dev)
        oco = torch.zeros(n_chunks + 1, dtype=torch.int64, device=dev)
        state = torch.zeros((n_streams, _lib.STATE_BYTES), dtype=torch.uint8, device=dev)
        batch.stream_feed_batch_noalloc(table, data, chunk_off, stream_off, None, flush=True, state_out=state,
                                        out_bytes=out, out_chunk_off=oco)
        torch.cuda.synchronize()
        total = int(exp_oco[-1].item())
        assert total > 0
        assert torch.equal(oco, exp_oco), f"call {call} ({n_streams} streams): per-chunk offsets differ"
        assert torch.equal(out[:total], exp_out[:total]), f"call {call} ({n_streams} streams): output bytes differ"
        assert torch.equal(state, exp_state), f"call {call} ({n_streams} streams): carried state differs"


def test_noalloc_rejects_batches_over_its_limit():
    from quorum_b200 import _lib, batch

    dev = torch.device("cuda", 0)
    table = batch.tag_table(["think"])
    n_streams = _lib.NOALLOC_MAX_STREAMS + 1
    stream_off = torch.zeros(n_streams + 1, dtype=torch.int64, device=dev)       # streams without chunks
    chunk_off = torch.zeros(1, dtype=torch.int64, device=dev)
    data = torch.zeros(16, dtype=torch.uint8, device=dev)
    out = torch.zeros(16, dtype=torch.uint8, device=dev)
    oco = torch.zeros(1, dtype=torch.int64, device=dev)
    with pytest.raises(_lib.QuorumB200Error, match="QB200_NOALLOC_MAX_STREAMS"):
        batch.stream_feed_batch_noalloc(table, data, chunk_off, stream_off, None, flush=False, state_out=None,
                                        out_bytes=out, out_chunk_off=oco)
