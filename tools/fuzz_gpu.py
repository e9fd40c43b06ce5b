"""Extended differential fuzz on a GPU box: the CUDA paths against the C oracle over many seeds (the pytest suite runs two
fixed seeds).  Usage: python tools/fuzz_gpu.py [first_seed] [n_seeds].  Exit code 1 on the first mismatch."""
import os
import random
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_golden as mg  # noqa: E402
import quorum_b200 as qb  # noqa: E402
from oracle import c_oracle  # noqa: E402
from quorum_b200 import synth  # noqa: E402

dev = torch.device("cuda", 0)
TAG_SETS = [["think", "reason", "reasoning", "thought", "Thought"], ["think"], [], [""], ["kiss", "x-1"], ["sick_k"]]


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a).copy()).to(dev)


def units_for(rng, tags, n, max_items):
    return [mg.rand_text(rng, tags, rng.randrange(0, max_items), exotic=rng.choice([0.0, 0.15, 0.5])).encode() for _ in range(n)]


def check_stream(seed, tags):
    rng = random.Random(seed)
    units = units_for(rng, tags, 1500, rng.choice([8, 40, 200]))
    chunks, stream_off = [], [0]
    for u in units:
        k = rng.randrange(0, 12)
        cuts = sorted(rng.randrange(0, len(u) + 1) for _ in range(k)) if k else []
        last = 0
        if k or len(u) or rng.random() < 0.5:
            for c in cuts + [len(u)]:
                chunks.append(u[last:c])
                last = c
        stream_off.append(len(chunks))
    data, chunk_off = synth.pack(chunks)
    stream_off = np.asarray(stream_off, dtype=np.int64)
    table, ot = qb.batch.tag_table(tags), c_oracle.Tags(tags)
    # two launches with carried state: first a random prefix of every stream's chunks, then the rest
    split = np.array([rng.randrange(a, b + 1) for a, b in zip(stream_off[:-1], stream_off[1:])], dtype=np.int64)
    order1 = np.concatenate([np.arange(a, s) for a, s in zip(stream_off[:-1], split)] or [np.zeros(0, np.int64)]).astype(np.int64)
    order2 = np.concatenate([np.arange(s, b) for s, b in zip(split, stream_off[1:])] or [np.zeros(0, np.int64)]).astype(np.int64)
    state, st_dev = None, None
    for order, counts, flush in ((order1, split - stream_off[:-1], False), (order2, stream_off[1:] - split, True)):
        part = [chunks[i] for i in order]
        d, co = synth.pack(part)
        so = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
        e_out, e_oco, e_st = c_oracle.stream_feed_batch(ot, d, co, so, state, c_oracle.FLAG_FLUSH if flush else 0)
        out, oco, st_dev = qb.batch.stream_feed_batch(table, t(d) if d.size else torch.zeros(16, dtype=torch.uint8, device=dev), t(co), t(so),
                                                      st_dev, flush=flush)
        torch.cuda.synchronize()
        assert np.array_equal(oco.cpu().numpy(), e_oco), ("oco", seed, tags)
        assert np.array_equal(out[: int(e_oco[-1])].cpu().numpy(), e_out), ("bytes", seed, tags)
        g = st_dev.cpu().numpy().reshape(-1).view(c_oracle.STATE_DTYPE)
        assert np.array_equal(g["depth"], e_st["depth"]) and np.array_equal(g["held_len"], e_st["held_len"]), ("state", seed, tags)
        assert np.array_equal(g["tail_len"], e_st["tail_len"]), ("tail", seed, tags)
        for a, b in zip(g, e_st):
            assert bytes(a["held"][: a["held_len"]]) == bytes(b["held"][: b["held_len"]]), ("held", seed, tags)
        state = e_st


def check_strip(seed, tags):
    rng = random.Random(seed)
    units = units_for(rng, tags, 2500, rng.choice([6, 30, 120]))
    data, in_off = synth.pack(units)
    nr = np.random.default_rng(seed)
    cuts = np.unique(np.concatenate([[0, len(units)], nr.integers(0, len(units), 600)])).astype(np.int64)
    sep = rng.choice([b"", b"\n", "\n\n—— § ——\n\n".encode()])
    streaming = rng.random() < 0.5
    flags = c_oracle.FLAG_HIDE | (c_oracle.FLAG_STREAM_JOIN if streaming else 0)
    e_out, e_off, e_len = c_oracle.strip_concat_batch(c_oracle.Tags(tags), data, in_off, cuts, sep, flags)
    sep_t = torch.frombuffer(bytearray(sep), dtype=torch.uint8).to(dev) if sep else None
    out, off, rl = qb.batch.strip_concat_batch(qb.batch.tag_table(tags), t(data), t(in_off), t(cuts), sep_t, hide=True, streaming=streaming,
                                               want_resp_len=True)
    torch.cuda.synchronize()
    assert np.array_equal(off.cpu().numpy(), e_off), ("off", seed, tags)
    assert np.array_equal(out[: int(e_off[-1])].cpu().numpy(), e_out), ("bytes", seed, tags)
    assert np.array_equal(rl[: len(units)].cpu().numpy(), e_len), ("len", seed, tags)


first = int(sys.argv[1]) if len(sys.argv) > 1 else 100
count = int(sys.argv[2]) if len(sys.argv) > 2 else 20
for seed in range(first, first + count):
    tags = TAG_SETS[seed % len(TAG_SETS)]
    check_stream(seed, tags)
    check_strip(seed, tags)
    print("seed", seed, "ok", flush=True)
print("fuzz ok:", count, "seeds")
