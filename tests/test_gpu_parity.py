"""Parity of the CUDA path (through the C ABI) with the oracle and the reference's golden vectors.  Needs a GPU.

Bar: bit-exact — every output byte, every offset, every carried-state field.
"""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)["cases"]


@pytest.fixture(scope="module")
def qb():
    import quorum_b200
    from quorum_b200 import _lib
    _lib.lib()                      # fail loudly if the extension is missing
    return quorum_b200


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


def _t(a, dev):
    return torch.from_numpy(np.array(a, copy=True)).to(dev)


# ---------------------------------------------------------------------------------------------------------------
# reference-shaped API against the reference's own vectors and the golden fixtures
# ---------------------------------------------------------------------------------------------------------------

def test_reference_unit_vectors_scalar_api(qb):
    from test_oracle_golden import REF_FILTER_VECTORS
    for tags, feeds, flush in REF_FILTER_VECTORS:
        f = qb.ThinkingTagFilter(tags)
        for text, expect in feeds:
            assert f.feed(text) == expect, (tags, text)
        if flush is not None:
            assert f.flush() == flush


def test_reference_concatenate_vectors(qb):
    T4 = ["think", "reason", "reasoning", "thought"]
    parts = ["Hello there, how may I assist you today?", "I'm the second assistant, ready to help!"]
    got = qb.concatenate(parts, "\n-------------\n", tags=T4, hide_final_think=False)
    assert got == "Hello there, how may I assist you today?\n-------------\nI'm the second assistant, ready to help!"
    think = "<think>I need to calculate 2+2. That's 4.</think>The answer is 4."
    assert qb.concatenate([think, think], "\n-------------\n", tags=T4, hide_final_think=True) == \
        "The answer is 4.\n-------------\nThe answer is 4."
    assert qb.concatenate([think, think], "\n-------------\n", tags=T4, hide_final_think=True, streaming=True) == \
        "The answer is 4.\n\n-------------\nThe answer is 4."
    assert qb.strip_thinking_tags(think, T4, False) is think


def test_golden_strip(qb):
    cases = _load("strip_cases")
    groups = {}
    for c in cases:
        groups.setdefault((tuple(c["tags"]), c["hide"]), []).append(c)
    for (tags, hide), cs in groups.items():
        got = qb.strip_thinking_tags_many([c["content"] for c in cs], list(tags), hide)
        for c, g in zip(cs, got):
            assert g == c["expect"], c


def test_golden_filter_one_launch_per_case_group(qb):
    """All feeds of a case are the chunks of one stream; every case of a tag set goes through ONE launch."""
    cases = _load("filter_cases")
    groups = {}
    for c in cases:
        groups.setdefault(tuple(c["tags"]), []).append(c)
    for tags, cs in groups.items():
        bank = qb.ThinkingTagFilterBatch(list(tags), len(cs))
        outs = bank.feed_chunks([c["feeds"] for c in cs])
        states = bank.states()
        for c, o, st in zip(cs, outs, states):
            assert o == c["outs"], c
            assert st[0] == c["depth"], c
        assert bank.flush() == [c["flush"] for c in cs]


def test_golden_filter_incremental_state(qb):
    """Same cases fed one chunk per launch: the carried state (depth + held partial) must reproduce the outputs."""
    cases = _load("filter_cases")
    groups = {}
    for c in cases:
        groups.setdefault(tuple(c["tags"]), []).append(c)
    for tags, cs in groups.items():
        cs = cs[:200]
        bank = qb.ThinkingTagFilterBatch(list(tags), len(cs))
        nmax = max(len(c["feeds"]) for c in cs)
        got = [[] for _ in cs]
        for k in range(nmax):
            step = [[c["feeds"][k]] if k < len(c["feeds"]) else [] for c in cs]
            res = bank.feed_chunks(step)
            for i, r in enumerate(res):
                got[i].extend(r)
        for c, g, st in zip(cs, got, bank.states()):
            assert g == c["outs"], c
            assert st[0] == c["depth"], c


def test_golden_filter_scalar_buffer_attr(qb):
    for c in _load("filter_cases")[:150]:
        f = qb.ThinkingTagFilter(c["tags"])
        outs = [f.feed(x) for x in c["feeds"]]
        assert outs == c["outs"], c
        assert f.thinking_depth == c["depth"] and f.buffer == c["buffer"], c
        assert f.flush() == c["flush"]


def test_golden_concat(qb):
    cases = _load("concat_cases")
    groups = {}
    for c in cases:
        groups.setdefault((tuple(c["tags"]), c["sep"], c["hide_final"], c["streaming"]), []).append(c)
    for (tags, sep, hide, streaming), cs in groups.items():
        got = qb.concatenate_many([c["contents"] for c in cs], sep, tags=list(tags), hide_final_think=hide,
                                  streaming=streaming)
        for c, g in zip(cs, got):
            assert g == c["expect"], c


def test_type_errors_match_reference(qb):
    with pytest.raises(TypeError):
        qb.strip_thinking_tags(None, ["think"], True)
    assert qb.strip_thinking_tags(None, ["think"], False) is None      # oai_proxy.py:134-135
    with pytest.raises(TypeError):
        qb.ThinkingTagFilter(["think"]).feed(None)
    with pytest.raises(ValueError):
        qb.strip_thinking_tags("x", ["a|b"], True)                     # regex metacharacters: rejected, not approximated


# ---------------------------------------------------------------------------------------------------------------
# batched device API against the C oracle, buffers compared byte for byte
# ---------------------------------------------------------------------------------------------------------------

def _check_strip_concat(qb, dev, tags, data, in_off, req_off, sep, hide, streaming):
    from oracle import c_oracle
    flags = (c_oracle.FLAG_HIDE if hide else 0) | (c_oracle.FLAG_STREAM_JOIN if streaming else 0)
    exp_out, exp_off, exp_len = c_oracle.strip_concat_batch(c_oracle.Tags(tags), data, in_off, req_off, sep, flags)
    table = qb.batch.tag_table(tags)
    sep_t = _t(np.frombuffer(sep, dtype=np.uint8), dev) if sep else None
    out, out_off, rl = qb.batch.strip_concat_batch(table, _t(data, dev), _t(in_off, dev),
                                                   None if req_off is None else _t(req_off, dev), sep_t,
                                                   hide=hide, streaming=streaming, want_resp_len=True)
    torch.cuda.synchronize()
    got_off = out_off.cpu().numpy()
    assert np.array_equal(got_off, exp_off)
    assert np.array_equal(out[: int(got_off[-1])].cpu().numpy(), exp_out)
    assert np.array_equal(rl.cpu().numpy()[: len(exp_len)], exp_len)
    return exp_out


def _check_stream(qb, dev, tags, data, chunk_off, stream_off, state_in, flush):
    from oracle import c_oracle
    exp_out, exp_oco, exp_state = c_oracle.stream_feed_batch(c_oracle.Tags(tags), data, chunk_off, stream_off, state_in,
                                                             c_oracle.FLAG_FLUSH if flush else 0)
    table = qb.batch.tag_table(tags)
    st = None if state_in is None else _t(state_in.view(np.uint8).reshape(-1, 32), dev)
    out, oco, st_out = qb.batch.stream_feed_batch(table, _t(data, dev), _t(chunk_off, dev), _t(stream_off, dev), st,
                                                  flush=flush)
    torch.cuda.synchronize()
    got_oco = oco.cpu().numpy()
    assert np.array_equal(got_oco, exp_oco)
    assert np.array_equal(out[: int(got_oco[-1])].cpu().numpy(), exp_out)
    got_state = st_out.cpu().numpy().reshape(-1).view(c_oracle.STATE_DTYPE)
    assert np.array_equal(got_state["depth"], exp_state["depth"])
    assert np.array_equal(got_state["tail_len"], exp_state["tail_len"])
    assert np.array_equal(got_state["held_len"], exp_state["held_len"])
    for g, e in zip(got_state, exp_state):
        assert bytes(g["held"][: g["held_len"]]) == bytes(e["held"][: e["held_len"]])
    return exp_out, exp_oco, got_state


def _fuzz_units(seed, n_units, tags, max_items=40):
    import random
    import sys
    sys.path.insert(0, GOLDEN)
    import make_golden as mg
    rng = random.Random(seed)
    return [mg.rand_text(rng, tags, rng.randrange(0, max_items), exotic=rng.choice([0.0, 0.15, 0.5])).encode()
            for _ in range(n_units)]


@pytest.mark.parametrize("tags", [["think", "reason", "reasoning", "thought", "Thought"], ["think"], [], ["kiss", "x-1"]])
def test_fuzz_strip_concat_vs_oracle(qb, dev, tags):
    from quorum_b200 import synth
    units = _fuzz_units(11, 3000, tags)
    units[5] = b""
    units[6] = b""
    data, in_off = synth.pack(units)
    rng = np.random.default_rng(5)
    cuts = np.unique(np.concatenate([[0, len(units)], rng.integers(0, len(units), 700)]))
    req_off = cuts.astype(np.int64)
    for sep in (b"", b"\n", "\n\n—— § ——\n\n".encode()):
        for hide in (True, False):
            for streaming in (False, True):
                _check_strip_concat(qb, dev, tags, data, in_off, req_off, sep, hide, streaming)
    _check_strip_concat(qb, dev, tags, data, in_off, None, b"", True, False)


@pytest.mark.parametrize("tags", [["think", "reason", "reasoning", "thought", "Thought"], ["think"], [], [""], ["kiss", "x-1"]])
def test_fuzz_stream_vs_oracle(qb, dev, tags):
    import random
    from quorum_b200 import synth
    rng = random.Random(3)
    units = _fuzz_units(12, 2000, tags)
    # random chunkings, including empty chunks and chunk-free streams
    chunks, stream_off = [], [0]
    for u in units:
        k = rng.randrange(0, 9)
        cuts = sorted(rng.randrange(0, len(u) + 1) for _ in range(k)) if k else []
        last = 0
        if k or len(u) or rng.random() < 0.5:
            for c in cuts + [len(u)]:
                chunks.append(u[last:c])
                last = c
        stream_off.append(len(chunks))
    data, chunk_off = synth.pack(chunks)
    stream_off = np.asarray(stream_off, dtype=np.int64)
    for flush in (False, True):
        _check_stream(qb, dev, tags, data, chunk_off, stream_off, None, flush)


def test_stream_ownership_many_short_and_empty_streams(qb, dev):
    """Shapes that stress the boundary -> stream lookup of the offsets kernel: one chunk per stream (a serving step
    that feeds one delta to every stream), and long runs of chunk-free streams between the ones that have chunks."""
    import random
    from quorum_b200 import synth
    tags = synth.TAGS5
    rng = random.Random(21)
    texts = [b"a<thi", b"nk>x</think>b ", b"plain text ", b"<think>", b"</think> ok", b"", b"<", b"x" * 70]
    # (1) one chunk per stream
    chunks = [rng.choice(texts) for _ in range(6000)]
    data, chunk_off = synth.pack(chunks)
    _check_stream(qb, dev, tags, data, chunk_off, np.arange(len(chunks) + 1, dtype=np.int64), None, False)
    # (2) mostly chunk-free streams, runs of up to 200 of them in a row
    chunks, stream_off = [], [0]
    while len(stream_off) < 30000:
        for _ in range(rng.randrange(0, 200)):
            stream_off.append(len(chunks))
        for _ in range(rng.randrange(1, 4)):
            chunks.append(rng.choice(texts))
        stream_off.append(len(chunks))
    data, chunk_off = synth.pack(chunks)
    for flush in (False, True):
        _check_stream(qb, dev, tags, data, chunk_off, np.asarray(stream_off, dtype=np.int64), None, flush)


def test_stream_incremental_launches_vs_oracle(qb, dev):
    """Feed the same streams in 1, 2, 5 launches with state carried on the device: outputs concatenate identically."""
    import random
    from quorum_b200 import synth
    tags = ["think", "reason", "reasoning", "thought", "Thought"]
    rng = random.Random(9)
    units = _fuzz_units(21, 800, tags)
    per_stream_chunks = []
    for u in units:
        cuts = sorted(rng.randrange(0, len(u) + 1) for _ in range(rng.randrange(0, 12)))
        last, cs = 0, []
        for c in cuts + [len(u)]:
            cs.append(u[last:c])
            last = c
        per_stream_chunks.append(cs)
    nmax = max(len(c) for c in per_stream_chunks)
    for n_launch in (1, 2, 5):
        state = None
        got = [b""] * len(units)
        bounds = [round(i * nmax / n_launch) for i in range(n_launch + 1)]
        for li in range(n_launch):
            flat, soff = [], [0]
            for cs in per_stream_chunks:
                flat.extend(cs[bounds[li]:bounds[li + 1]])
                soff.append(len(flat))
            data, chunk_off = synth.pack(flat)
            exp_out, oco_exp, gstate = _check_stream(qb, dev, tags, data, chunk_off, np.asarray(soff, dtype=np.int64),
                                                     state, flush=(li == n_launch - 1))
            state = gstate.copy()
            for s in range(len(units)):
                got[s] += bytes(exp_out[oco_exp[soff[s]]: oco_exp[soff[s + 1]]])
        if n_launch == 1:
            base = got
        else:
            assert got == base


# ---------------------------------------------------------------------------------------------------------------
# BASELINE.json config shapes at sizes the oracle finishes in seconds
# ---------------------------------------------------------------------------------------------------------------

def test_cfg2_shape_streaming(qb, dev):
    from quorum_b200 import synth
    base = synth.cfg2_base(192)
    idx = synth.gather_units(base, 1024, seed=2)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    chunk_off = np.arange(0, data.size + 1, 64, dtype=np.int64)
    stream_off = np.arange(0, 1024 * 256 + 1, 256, dtype=np.int64)
    out, _, _ = _check_stream(qb, dev, synth.TAGS5, data, chunk_off, stream_off, None, True)
    assert 0.3 < out.size / data.size < 0.7


def test_cfg3_shape_batch_concat(qb, dev):
    from quorum_b200 import synth
    base = synth.cfg3_base(1024)
    idx = synth.gather_units(base, 8192, seed=3)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    in_off = np.arange(0, data.size + 1, 4096, dtype=np.int64)
    req_off = np.arange(0, 8192 + 1, 4, dtype=np.int64)
    _check_strip_concat(qb, dev, synth.TAGS5, data, in_off, req_off, synth.SEP_MULTIBYTE.encode(), True, False)


def test_cfg4_shape_reasoning_heavy(qb, dev):
    from quorum_b200 import synth
    base = synth.cfg4_base(128)
    idx = synth.gather_units(base, 512, seed=4)
    data = np.ascontiguousarray(base[idx]).reshape(-1)
    in_off = np.arange(0, data.size + 1, 32768, dtype=np.int64)
    req_off = np.arange(0, 512 + 1, 8, dtype=np.int64)
    for streaming in (False, True):
        sep = ("\n" if streaming else "").encode() + b"\n-------------\n"
        out = _check_strip_concat(qb, dev, ["think", "reason", "reasoning", "thought"], data, in_off, req_off, sep, True,
                                  streaming)
        assert 0.2 < out.size / data.size < 0.4


def test_cfg5_shape_adversarial(qb, dev):
    from quorum_b200 import synth
    units = synth.cfg5_units(24 << 20)
    data, in_off = synth.pack(units)
    req_off = np.arange(0, len(units) + 1, 1, dtype=np.int64)
    _check_strip_concat(qb, dev, synth.TAGS5, data, in_off, req_off, b"\n", True, False)
    # the same bytes as streams cut every 64 bytes on code-point boundaries
    chunks, stream_off = [], [0]
    for u in units[:400]:
        p = 0
        while p < len(u):
            q = min(len(u), p + 64)
            while q < len(u) and (u[q] & 0xC0) == 0x80:
                q -= 1
            chunks.append(u[p:q])
            p = q
        stream_off.append(len(chunks))
    cdata, chunk_off = synth.pack(chunks)
    _check_stream(qb, dev, synth.TAGS5, cdata, chunk_off, np.asarray(stream_off, dtype=np.int64), None, True)


def test_edge_shapes(qb, dev):
    """Empty batch, empty responses, one huge unit, unaligned ragged starts."""
    from quorum_b200 import synth
    tags = synth.TAGS5
    # empty responses / requests
    units = [b"", b"", b"x", b"", b" <think>a</think> "]
    data, in_off = synth.pack(units)
    _check_strip_concat(qb, dev, tags, np.concatenate([data, np.zeros(1, np.uint8)])[: data.size], in_off,
                        np.array([0, 0, 2, 2, 5], dtype=np.int64), b"|", True, True)
    _check_strip_concat(qb, dev, tags, data, in_off, np.array([0, 0, 2, 2, 5], dtype=np.int64), b"|", True, False)
    # one 6 MiB response with blocks far apart, and a 3 MiB stream in 3 chunks
    rng = np.random.default_rng(1)
    big = synth.make_unit(rng, 6 << 20, 5, 0.4)
    data, in_off = synth.pack([b"ab", big, b" tail "])
    _check_strip_concat(qb, dev, tags, data, in_off, None, b"", True, False)
    s = synth.make_unit(rng, 3 << 20, 7, 0.5)
    cdata, chunk_off = synth.pack([s[: 1 << 20], s[1 << 20: (2 << 20) + 5], s[(2 << 20) + 5:]])
    _check_stream(qb, dev, tags, cdata, chunk_off, np.array([0, 3], dtype=np.int64), None, False)
    # all-white-space and white-space-heavy edges
    ws = (" \t\n\x0b\x0c\r\x1c\x1d\x1e\x1f\x85\xa0\u1680\u2000\u2001\u2005\u200a\u2028\u2029\u202f\u205f\u3000" * 40).encode()
    units = [ws, ws + b"x" + ws, ws + b"<think>" + ws + b"</think>" + ws, b"\xe2\x80\x8b" + ws]
    data, in_off = synth.pack(units)
    _check_strip_concat(qb, dev, tags, data, in_off, None, b"", True, False)


def test_launch_counter_moves(qb):
    before = qb.launch_count()
    qb.strip_thinking_tags("a <think>b</think> c", ["think"])
    assert qb.launch_count() > before
