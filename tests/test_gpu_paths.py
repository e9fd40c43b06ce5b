"""The two implementations behind the C ABI must agree bit for bit: the record-driven multi-kernel pipeline (default,
QB200_PATH=token) and the byte-rescanning walkers (QB200_PATH=general, also what the other falls back to for units that do
not fit its tables).  Also: full BASELINE.json sizes with sampled oracle checks
and size-independent properties."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def qb():
    import quorum_b200
    from quorum_b200 import _lib
    _lib.lib()
    return quorum_b200


@pytest.fixture()
def path_env():
    old = os.environ.get("QB200_PATH")
    yield
    if old is None:
        os.environ.pop("QB200_PATH", None)
    else:
        os.environ["QB200_PATH"] = old


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a).copy()).to(dev)


def _stream_all_paths(qb, tags, data, chunk_off, stream_off, flush):
    dev = torch.device("cuda", 0)
    table = qb.batch.tag_table(tags)
    res = {}
    for path in ("token", "general"):
        os.environ["QB200_PATH"] = path
        out, oco, st = qb.batch.stream_feed_batch(table, _t(data, dev), _t(chunk_off, dev), _t(stream_off, dev), None,
                                                  flush=flush)
        torch.cuda.synchronize()
        n = int(oco[-1].item())
        res[path] = (out[:n].cpu().numpy(), oco.cpu().numpy(), st.cpu().numpy())
    for a, b in zip(res["token"], res["general"]):
        assert np.array_equal(a, b)
    return res["token"]


def test_stream_paths_agree_on_cfg2_and_ragged_shapes(qb, path_env):
    from oracle import c_oracle
    from quorum_b200 import synth
    base = synth.cfg2_base(96)
    idx = synth.gather_units(base, 700, seed=5)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    chunk_off = np.arange(0, data.size + 1, 64, dtype=np.int64)
    stream_off = np.arange(0, 700 * 256 + 1, 256, dtype=np.int64)
    out, oco, _ = _stream_all_paths(qb, synth.TAGS5, data, chunk_off, stream_off, True)
    exp_out, exp_oco, _ = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), data, chunk_off, stream_off, None,
                                                     c_oracle.FLAG_FLUSH)
    assert np.array_equal(oco, exp_oco) and np.array_equal(out, exp_out)
    # ragged: adversarial units of 64 B .. 64 KiB cut into uneven chunks (some empty), odd alignment everywhere
    rng = np.random.default_rng(3)
    units = synth.cfg5_units(6 << 20, seed=77)
    chunks, soff = [], [0]
    for u in units:
        p = 0
        while p < len(u):
            q = min(len(u), p + int(rng.integers(0, 200)))
            while q < len(u) and (u[q] & 0xC0) == 0x80:
                q += 1
            chunks.append(u[p:q])
            p = q
        soff.append(len(chunks))
    cdata, coff = synth.pack(chunks)
    out, oco, _ = _stream_all_paths(qb, synth.TAGS5, cdata, coff, np.asarray(soff, dtype=np.int64), False)
    exp_out, exp_oco, _ = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), cdata, coff,
                                                     np.asarray(soff, dtype=np.int64), None, 0)
    assert np.array_equal(oco, exp_oco) and np.array_equal(out, exp_out)


def test_strip_paths_agree(qb, path_env):
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    units = synth.cfg5_units(8 << 20, seed=91) + [synth.make_unit(np.random.default_rng(2), 300000, 9, 0.5)]
    data, in_off = synth.pack(units)
    req_off = np.unique(np.concatenate([[0, len(units)], np.random.default_rng(1).integers(0, len(units), 200)])).astype(np.int64)
    table = qb.batch.tag_table(synth.TAGS5)
    sep = _t(np.frombuffer(synth.SEP_MULTIBYTE.encode(), dtype=np.uint8), dev)
    from oracle import c_oracle
    for streaming in (True, False):
        res = {}
        for path in ("token", "general"):
            os.environ["QB200_PATH"] = path
            out, off, rl = qb.batch.strip_concat_batch(table, _t(data, dev), _t(in_off, dev), _t(req_off, dev), sep, hide=True,
                                                       streaming=streaming, want_resp_len=True)
            torch.cuda.synchronize()
            n = int(off[-1].item())
            res[path] = (out[:n].cpu().numpy(), off.cpu().numpy(), rl.cpu().numpy()[: len(units)])
        for a, b in zip(res["token"], res["general"]):
            assert np.array_equal(a, b), streaming
        flags = c_oracle.FLAG_HIDE | (c_oracle.FLAG_STREAM_JOIN if streaming else 0)
        e_out, e_off, _ = c_oracle.strip_concat_batch(c_oracle.Tags(synth.TAGS5), data, in_off, req_off, synth.SEP_MULTIBYTE.encode(), flags)
        assert np.array_equal(res["token"][1], e_off) and np.array_equal(res["token"][0], e_out)


def test_record_slot_overflow_falls_back(qb):
    """More '<' per KiB than the tokenizer has record slots for, and more kept runs than a descriptor holds."""
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    tags = synth.TAGS5
    dense = ("a<b <c> <<think>x</think>> </reason> <th " * 400).encode()            # '<' every few bytes
    many_runs = ("keep <think>drop</think> " * 300).encode()                        # 300 kept runs in one unit
    units = [dense, many_runs, b"plain text", dense[:1500], many_runs[:900]]
    data, in_off = synth.pack(units)
    exp = c_oracle.strip_concat_batch(c_oracle.Tags(tags), data, in_off, None, b"", c_oracle.FLAG_HIDE)
    out, off, _ = qb.batch.strip_concat_batch(qb.batch.tag_table(tags), _t(data, dev), _t(in_off, dev), None, None, hide=True,
                                              streaming=False)
    torch.cuda.synchronize()
    assert np.array_equal(off.cpu().numpy(), exp[1]) and np.array_equal(out[: int(exp[1][-1])].cpu().numpy(), exp[0])
    chunk_off = np.unique(np.concatenate([in_off, np.arange(0, data.size, 97)])).astype(np.int64)
    stream_off = np.searchsorted(chunk_off, in_off).astype(np.int64)
    exp = c_oracle.stream_feed_batch(c_oracle.Tags(tags), data, chunk_off, stream_off, None, 0)
    out, oco, st = qb.batch.stream_feed_batch(qb.batch.tag_table(tags), _t(data, dev), _t(chunk_off, dev), _t(stream_off, dev),
                                              None, flush=False)
    torch.cuda.synchronize()
    assert np.array_equal(oco.cpu().numpy(), exp[1]) and np.array_equal(out[: int(exp[1][-1])].cpu().numpy(), exp[0])
    assert np.array_equal(st.cpu().numpy().reshape(-1).view(c_oracle.STATE_DTYPE)["depth"], exp[2]["depth"])


def test_cfg2_full_size_sampled_against_oracle(qb):
    """BASELINE.json configs[1] at full size (65,536 streams x 256 chunks x 64 B = 1 GiB): every per-chunk offset is
    checked through size-independent properties, 3,000 random streams byte for byte against the oracle."""
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    n_streams, chunks, cb = 65536, 256, 64
    base = synth.cfg2_base(256)
    idx = synth.gather_units(base, n_streams, seed=1000)
    data = torch.from_numpy(base).to(dev).index_select(0, torch.from_numpy(idx).to(dev)).reshape(-1).contiguous()
    chunk_off = torch.arange(0, data.numel() + 1, cb, dtype=torch.int64, device=dev)
    stream_off = torch.arange(0, n_streams * chunks + 1, chunks, dtype=torch.int64, device=dev)
    out, oco, st = qb.batch.stream_feed_batch(qb.batch.tag_table(synth.TAGS5), data, chunk_off, stream_off, None, flush=True)
    torch.cuda.synchronize()
    lens = oco[1:] - oco[:-1]
    assert int(oco[0]) == 0 and bool((lens >= 0).all())                       # offsets are a monotone prefix
    assert bool((lens <= cb + 23).all())                                      # a feed returns at most its chunk + a held tail
    per_stream = lens.reshape(n_streams, chunks).sum(1)
    # identical streams (same base unit) must give identical results wherever they sit in the batch
    idx_t = torch.from_numpy(idx).to(dev)
    ref_len = torch.zeros(base.shape[0], dtype=torch.int64, device=dev).scatter_(0, idx_t, per_stream)
    assert torch.equal(ref_len[idx_t], per_stream)
    # oracle on the base pool gives every stream's expected output
    flat = np.ascontiguousarray(base).reshape(-1)
    co = np.arange(0, flat.size + 1, cb, dtype=np.int64)
    so = np.arange(0, base.shape[0] * chunks + 1, chunks, dtype=np.int64)
    e_out, e_oco, e_st = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), flat, co, so, None, c_oracle.FLAG_FLUSH)
    e_len = (e_oco[1:] - e_oco[:-1]).reshape(base.shape[0], chunks)
    assert np.array_equal(per_stream.cpu().numpy(), e_len.sum(1)[idx])
    assert int(oco[-1]) == int(e_len.sum(1)[idx].sum())
    got_lens = lens.reshape(n_streams, chunks)
    rng = np.random.default_rng(11)
    sample = rng.integers(0, n_streams, 3000)
    sl = got_lens[torch.from_numpy(sample).to(dev)].cpu().numpy()
    assert np.array_equal(sl, e_len[idx[sample]])
    oco_h = oco.cpu().numpy()
    for s in sample[:400]:
        b = idx[s]
        a0, a1 = oco_h[s * chunks], oco_h[(s + 1) * chunks]
        assert bytes(out[a0:a1].cpu().numpy()) == bytes(e_out[e_oco[b * chunks]: e_oco[(b + 1) * chunks]])
    depth = st.cpu().numpy().reshape(-1).view(c_oracle.STATE_DTYPE)["depth"]
    assert np.array_equal(depth, e_st["depth"][idx])


@pytest.mark.parametrize("path", ["token", "general"])
def test_chained_slices_and_pipelined_host_feeder(qb, path_env, path):
    """QB200_CHAIN: a batch fed slice by slice (what the host-buffer session does to overlap PCIe copies with compute)
    gives the same bytes and the same per-chunk offsets as one call, whichever implementation takes the first slice."""
    os.environ["QB200_PATH"] = path
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    base = synth.cfg2_base(64)
    idx = synth.gather_units(base, 640, seed=8)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    chunk_off = np.arange(0, data.size + 1, 64, dtype=np.int64)
    stream_off = np.arange(0, 640 * 256 + 1, 256, dtype=np.int64)
    exp_out, exp_oco, exp_st = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), data, chunk_off, stream_off, None, 0)
    feeder = qb.batch.HostStreamFeeder(qb.batch.tag_table(synth.TAGS5), dev, data.size, len(chunk_off) - 1, 640)
    pin = lambda a: torch.from_numpy(np.array(a, copy=True)).pin_memory()  # noqa: E731
    for slices in (1, 5, 8):
        feeder.h_out.zero_()
        feeder.h_oco.zero_()
        feeder.run(pin(data), pin(chunk_off), pin(stream_off), flush=False, slices=slices)
        torch.cuda.synchronize()
        assert np.array_equal(feeder.h_oco[: len(exp_oco)].numpy(), exp_oco), slices
        assert np.array_equal(feeder.h_out[: int(exp_oco[-1])].numpy(), exp_out), slices
        st = feeder.d_state[:640].cpu().numpy().reshape(-1).view(c_oracle.STATE_DTYPE)
        assert np.array_equal(st["depth"], exp_st["depth"]) and np.array_equal(st["held_len"], exp_st["held_len"])


def test_compact_host_feeder_32bit_lengths(qb):
    """run_compact: int32 chunk lengths up, int32 output lengths down, offsets built and narrowed on the device
    (qb200_offsets_from_lengths32 / qb200_lengths32_from_offsets).  Ragged chunks (empty ones included), several slice
    counts; bytes, per-chunk lengths and carried state equal the oracle's."""
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    base = synth.cfg2_base(64)
    idx = synth.gather_units(base, 640, seed=9)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    per = data.size // 640
    rng = np.random.default_rng(12)
    cuts = np.sort(rng.integers(0, per + 1, size=(640, 255)), axis=1)
    bounds = np.concatenate([np.zeros((640, 1), np.int64), cuts, np.full((640, 1), per, np.int64)], axis=1) + (np.arange(640) * per)[:, None]
    chunk_len = np.diff(bounds, axis=1).astype(np.int32).reshape(-1)
    chunk_off = np.concatenate([[0], np.cumsum(chunk_len, dtype=np.int64)])
    stream_off = np.arange(0, 640 * 256 + 1, 256, dtype=np.int64)
    stream_byte_off = np.arange(0, 641, dtype=np.int64) * per
    exp_out, exp_oco, exp_st = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), data, chunk_off, stream_off, None, 0)
    feeder = qb.batch.HostStreamFeeder(qb.batch.tag_table(synth.TAGS5), dev, data.size, len(chunk_len), 640)
    pin = lambda a: torch.from_numpy(np.array(a, copy=True)).pin_memory()  # noqa: E731
    for slices in (1, 5, 8):
        feeder.h_out.zero_()
        h2d, d2h = feeder.run_compact(pin(data), pin(chunk_len), pin(stream_off), pin(stream_byte_off), flush=False, slices=slices)
        torch.cuda.synchronize()
        assert int(feeder.h_ends[-1]) == int(exp_oco[-1]), slices
        assert np.array_equal(feeder.h_olen[: len(chunk_len)].numpy(), np.diff(exp_oco).astype(np.int32)), slices
        assert np.array_equal(feeder.h_out[: int(exp_oco[-1])].numpy(), exp_out), slices
        st = feeder.d_state[:640].cpu().numpy().reshape(-1).view(c_oracle.STATE_DTYPE)
        assert np.array_equal(st["depth"], exp_st["depth"]) and np.array_equal(st["held_len"], exp_st["held_len"])
        assert h2d < data.size + 8 * len(chunk_off) and d2h < int(exp_oco[-1]) + 8 * len(chunk_off)


def test_chained_slices_and_pipelined_host_strip_concat(qb):
    """QB200_CHAIN on strip + concat: the host-buffer session processes the batch in slices of requests on three streams;
    bytes and request offsets equal those of one call (ragged requests, multi-byte separator, both join forms)."""
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    units = synth.cfg5_units(6 << 20, seed=31)
    data, in_off = synth.pack(units)
    rng = np.random.default_rng(7)
    cuts = np.unique(np.concatenate([[0, len(units)], rng.integers(0, len(units), len(units) // 3)])).astype(np.int64)
    sep = synth.SEP_MULTIBYTE.encode()
    sess = qb.batch.HostStripConcat(qb.batch.tag_table(synth.TAGS5), dev, data.size, len(units), len(cuts) - 1, sep)
    pin = lambda a: torch.from_numpy(np.array(a, copy=True)).pin_memory()  # noqa: E731
    for streaming in (False, True):
        flags = c_oracle.FLAG_HIDE | (c_oracle.FLAG_STREAM_JOIN if streaming else 0)
        exp_out, exp_off, _ = c_oracle.strip_concat_batch(c_oracle.Tags(synth.TAGS5), data, in_off, cuts, sep, flags)
        for slices in (1, 3, 8):
            sess.h_out.zero_()
            sess.h_oo.zero_()
            sess.run(pin(data), pin(in_off), pin(cuts), hide=True, streaming=streaming, slices=slices)
            torch.cuda.synchronize()
            assert np.array_equal(sess.h_oo[: len(exp_off)].numpy(), exp_off), (streaming, slices)
            assert np.array_equal(sess.h_out[: int(exp_off[-1])].numpy(), exp_out), (streaming, slices)


@pytest.mark.parametrize("path", ["general"])
def test_chained_slices_strip_concat_paths(qb, path_env, path):
    os.environ["QB200_PATH"] = path
    test_chained_slices_and_pipelined_host_strip_concat(qb)


@pytest.mark.parametrize("cfg", ["cfg3", "cfg4"])
def test_cfg3_cfg4_full_size_sampled_against_oracle(qb, cfg):
    """BASELINE.json configs[2] (1M responses x 4 KiB, 4 per request, multi-byte separator) and configs[3] (8 x 32 KiB
    per request, 70 % inside <think>) at full size: every request's length is checked against the oracle's per-response
    kept lengths (a request is a sum over its responses + separators), 300 random requests byte for byte."""
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    if cfg == "cfg3":
        base, n_resp, per_req, sep = synth.cfg3_base(4096), 1_000_000, 4, synth.SEP_MULTIBYTE.encode()
    else:
        base, n_resp, per_req, sep = synth.cfg4_base(512), 131_072, 8, b"\n-------------\n"
    ub = base.shape[1]
    idx = synth.gather_units(base, n_resp, seed=1000)
    idx_t = torch.from_numpy(idx).to(dev)
    data = torch.from_numpy(base).to(dev).index_select(0, idx_t).reshape(-1).contiguous()
    in_off = torch.arange(0, n_resp * ub + 1, ub, dtype=torch.int64, device=dev)
    req_off = torch.arange(0, n_resp + 1, per_req, dtype=torch.int64, device=dev)
    sep_t = torch.frombuffer(bytearray(sep), dtype=torch.uint8).to(dev)
    out, out_off, rlen = qb.batch.strip_concat_batch(qb.batch.tag_table(synth.TAGS5), data, in_off, req_off, sep_t, hide=True,
                                                     streaming=False, want_resp_len=True)
    torch.cuda.synchronize()
    # oracle over the base pool: kept length and bytes of every distinct response
    flat = np.ascontiguousarray(base).reshape(-1)
    bo = np.arange(0, flat.size + 1, ub, dtype=np.int64)
    e_out, e_off, e_len = c_oracle.strip_concat_batch(c_oracle.Tags(synth.TAGS5), flat, bo, None, b"", c_oracle.FLAG_HIDE)
    assert np.array_equal(rlen[:n_resp].cpu().numpy(), e_len[idx])           # every response's filtered length
    n_req = n_resp // per_req
    exp_req = e_len[idx].reshape(n_req, per_req).sum(1) + (per_req - 1) * len(sep)
    got_off = out_off.cpu().numpy()
    assert got_off[0] == 0 and np.array_equal(np.diff(got_off), exp_req)      # offsets: prefix sum of the request lengths
    rng = np.random.default_rng(12)
    for q in rng.integers(0, n_req, 300):
        parts = [bytes(e_out[e_off[b]:e_off[b + 1]]) for b in idx[q * per_req:(q + 1) * per_req]]
        assert bytes(out[got_off[q]:got_off[q + 1]].cpu().numpy()) == sep.join(parts)


def test_small_grid_blocks_loop_and_take_many_tickets(qb, monkeypatch):
    """QB200_GRID_CAP=3: every warp-per-unit / lane-per-unit kernel runs with three blocks, so blocks loop over many groups
    and the count kernel's look-back sees blocks that each took dozens of tickets.  Same bytes as the oracle."""
    from oracle import c_oracle
    from quorum_b200 import synth
    monkeypatch.setenv("QB200_GRID_CAP", "3")
    dev = torch.device("cuda", 0)
    base = synth.cfg2_base(64)
    n = 20000
    idx = synth.gather_units(base, n, seed=3)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    chunk_off = np.arange(0, data.size + 1, 64, dtype=np.int64)
    stream_off = np.arange(0, n * 256 + 1, 256, dtype=np.int64)
    exp_out, exp_oco, exp_st = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), data, chunk_off, stream_off, None, 0)
    out, oco, st = qb.batch.stream_feed_batch(qb.batch.tag_table(synth.TAGS5), _t(data, dev), _t(chunk_off, dev), _t(stream_off, dev),
                                              None, flush=False)
    torch.cuda.synchronize()
    assert np.array_equal(oco.cpu().numpy(), exp_oco)
    assert np.array_equal(out[: int(exp_oco[-1])].cpu().numpy(), exp_out)
    units = synth.cfg5_units(4 << 20, seed=41)
    d2, in_off = synth.pack(units)
    req = np.arange(0, len(units) + 1, 1, dtype=np.int64)
    e_out, e_off, _ = c_oracle.strip_concat_batch(c_oracle.Tags(synth.TAGS5), d2, in_off, req, b"", c_oracle.FLAG_HIDE)
    o2, off2, _ = qb.batch.strip_concat_batch(qb.batch.tag_table(synth.TAGS5), _t(d2, dev), _t(in_off, dev), _t(req, dev), None, hide=True,
                                              streaming=False)
    torch.cuda.synchronize()
    assert np.array_equal(off2.cpu().numpy(), e_off) and np.array_equal(o2[: int(e_off[-1])].cpu().numpy(), e_out)


def test_capacity_below_the_input_size_stays_exact(qb):
    """A caller whose output is known to be short may pass out_bytes (and so out_capacity, which sizes the token-record
    tables) smaller than the input: the units beyond the stated bound must be walked from their bytes, not from record
    slots that were never written."""
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    base = synth.cfg4_base(24, 32768)                                  # 70 % of every response is removed
    data = np.ascontiguousarray(base).reshape(-1)
    in_off = np.arange(0, data.size + 1, 32768, dtype=np.int64)
    req_off = np.arange(0, base.shape[0] + 1, 8, dtype=np.int64)
    sep = b"\n--\n"
    e_out, e_off, _ = c_oracle.strip_concat_batch(c_oracle.Tags(synth.TAGS5), data, in_off, req_off, sep, c_oracle.FLAG_HIDE)
    cap = int(e_off[-1]) + 4096
    assert cap < data.size // 2
    out = torch.zeros(cap, dtype=torch.uint8, device=dev)
    got, off, _ = qb.batch.strip_concat_batch(qb.batch.tag_table(synth.TAGS5), _t(data, dev), _t(in_off, dev), _t(req_off, dev),
                                              _t(np.frombuffer(sep, dtype=np.uint8), dev), hide=True, streaming=False, out_bytes=out)
    torch.cuda.synchronize()
    assert np.array_equal(off.cpu().numpy(), e_off) and np.array_equal(got[: int(e_off[-1])].cpu().numpy(), e_out)
    # streams: half of every stream is inside thinking blocks
    sb = synth.cfg2_base(40)
    sdata = np.ascontiguousarray(sb).reshape(-1)
    chunk_off = np.arange(0, sdata.size + 1, 64, dtype=np.int64)
    stream_off = np.arange(0, 40 * 256 + 1, 256, dtype=np.int64)
    s_out, s_oco, _ = c_oracle.stream_feed_batch(c_oracle.Tags(synth.TAGS5), sdata, chunk_off, stream_off, None, c_oracle.FLAG_FLUSH)
    cap = int(s_oco[-1]) + 4096
    assert cap < sdata.size
    out = torch.zeros(cap, dtype=torch.uint8, device=dev)
    got, oco, _ = qb.batch.stream_feed_batch(qb.batch.tag_table(synth.TAGS5), _t(sdata, dev), _t(chunk_off, dev), _t(stream_off, dev),
                                             None, flush=True, out_bytes=out)
    torch.cuda.synchronize()
    assert np.array_equal(oco.cpu().numpy(), s_oco) and np.array_equal(got[: int(s_oco[-1])].cpu().numpy(), s_out)


@pytest.mark.parametrize("scanwalk", ["1", "0"])
def test_warp_walkers_for_long_responses_match_the_oracle(qb, monkeypatch, scanwalk):
    """QB200_LANE_STEPS=1 sends every response that spans more than one KiB step to the warp walkers: the scan walker
    (transition vectors composed by a warp scan, qb_scanwalk.cuh; QB200_SCANWALK=1, the default) or the run-list walker
    behind it (QB200_SCANWALK=0).  Adversarial cfg5 mix plus long responses with hundreds of blocks, both join forms."""
    from oracle import c_oracle
    from quorum_b200 import synth
    monkeypatch.setenv("QB200_LANE_STEPS", "1")
    monkeypatch.setenv("QB200_SCANWALK", scanwalk)
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(17)
    units = synth.cfg5_units(5 << 20, seed=97)
    units += [bytes(synth.make_unit(rng, 200_000, 90, 0.5)), bytes(synth.make_unit(rng, 70_000, 3, 0.7)),
              ("keep <think>drop</think> " * 3000).encode(), b"<think>" * 5000 + b"</think>", b" \n" * 40_000, b"x" * 100_000]
    data, in_off = synth.pack(units)
    req_off = np.unique(np.concatenate([[0, len(units)], rng.integers(0, len(units), len(units) // 4)])).astype(np.int64)
    sep = synth.SEP_MULTIBYTE.encode()
    table = qb.batch.tag_table(synth.TAGS5)
    for streaming in (False, True):
        flags = c_oracle.FLAG_HIDE | (c_oracle.FLAG_STREAM_JOIN if streaming else 0)
        e_out, e_off, e_len = c_oracle.strip_concat_batch(c_oracle.Tags(synth.TAGS5), data, in_off, req_off, sep, flags)
        out, off, rl = qb.batch.strip_concat_batch(table, _t(data, dev), _t(in_off, dev), _t(req_off, dev),
                                                   _t(np.frombuffer(sep, dtype=np.uint8), dev), hide=True, streaming=streaming,
                                                   want_resp_len=True)
        torch.cuda.synchronize()
        assert np.array_equal(rl.cpu().numpy()[: len(units)], e_len), streaming
        assert np.array_equal(off.cpu().numpy(), e_off), streaming
        assert np.array_equal(out[: int(e_off[-1])].cpu().numpy(), e_out), streaming
