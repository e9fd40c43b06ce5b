"""The multi-GPU product calls: one host batch, partitioned by bytes over the devices, outputs gathered
on the host - sharded == unsharded == oracle, byte for byte, on the adversarial cfg5 mix.  The same-device variant runs
the partition / gather logic on a one-GPU box; the multi-device variant needs at least two GPUs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _cfg5(total, seed):
    from quorum_b200 import synth
    units = synth.cfg5_units(total, seed=seed)
    data, in_off = synth.pack(units)
    rng = np.random.default_rng(seed)
    cuts = np.unique(np.concatenate([[0, len(units)], rng.integers(0, len(units), len(units) // 3)])).astype(np.int64)
    return units, data, in_off, cuts


def _check_strip(devices):
    from oracle import c_oracle
    from quorum_b200 import sharding, synth
    units, data, in_off, cuts = _cfg5(5 << 20, 61)
    sep = synth.SEP_MULTIBYTE.encode()
    pin = lambda a: torch.from_numpy(np.array(a, copy=True)).pin_memory()  # noqa: E731
    for streaming in (False, True):
        flags = c_oracle.FLAG_HIDE | (c_oracle.FLAG_STREAM_JOIN if streaming else 0)
        e_out, e_off, _ = c_oracle.strip_concat_batch(c_oracle.Tags(synth.TAGS5), data, in_off, cuts, sep, flags)
        out, off = sharding.strip_concat_sharded(synth.TAGS5, pin(data), pin(in_off), pin(cuts), sep, hide=True,
                                                 streaming=streaming, devices=devices)
        assert np.array_equal(off.numpy(), e_off), streaming
        assert np.array_equal(out.numpy(), e_out[: int(e_off[-1])]), streaming


def _check_stream(devices):
    from oracle import c_oracle
    from quorum_b200 import sharding, synth
    units, data, in_off, _ = _cfg5(3 << 20, 62)
    rng = np.random.default_rng(9)
    chunk_off = np.unique(np.concatenate([in_off, rng.integers(0, data.size, data.size // 90)])).astype(np.int64)
    stream_off = np.searchsorted(chunk_off, in_off).astype(np.int64)
    pin = lambda a: torch.from_numpy(np.array(a, copy=True)).pin_memory()  # noqa: E731
    e_out, e_oco, e_st = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), data, chunk_off, stream_off, None, c_oracle.FLAG_FLUSH)
    out, oco, st = sharding.stream_feed_sharded(synth.TAGS5, pin(data), pin(chunk_off), pin(stream_off), flush=True, devices=devices)
    assert np.array_equal(oco.numpy(), e_oco)
    assert np.array_equal(out.numpy(), e_out[: int(e_oco[-1])])
    got = st.numpy().reshape(-1).view(c_oracle.STATE_DTYPE)
    assert np.array_equal(got["depth"], e_st["depth"]) and np.array_equal(got["held_len"], e_st["held_len"])


def test_sharded_calls_on_one_device_used_three_times():
    _check_strip([0, 0, 0])
    _check_stream([0, 0, 0])


def test_sharded_calls_over_all_devices():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least two GPUs")
    devs = list(range(torch.cuda.device_count()))
    _check_strip(devs)
    _check_stream(devs)
