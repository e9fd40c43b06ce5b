"""SSE input parsing: the device parser against the reference's own code path
(src/quorum/oai_proxy.py:578-617 as restated by quorum_b200.streaming.parse_sse_chunk, which uses json.loads)."""
import json
import random

import pytest

ALPHA = ["a", "b", " ", "\n", '"', "\\", "/", "é", "日本", "\u2028", "\U0001F389", "\ud83d", "\ude00", "\x7f", "<think>", "</think>", "\t",
         "{", "}", "[", "]", ":", ",", "data: ", "\r"]


def _event(rng, content, style):
    obj = {"id": "chatcmpl-1", "object": "chat.completion.chunk", "created": 1700000000, "model": "m",
           "choices": [{"index": 0, "delta": {"content": content}, "finish_reason": None}]}
    if style == 1:
        obj["choices"][0]["delta"] = {"role": "assistant", "content": content}
    if style == 2:
        obj = {"choices": [{"delta": {"content": content}}], "usage": {"x": [1, 2.5, -3e10, True, False, None, {"y": "z"}]}}
    if style == 3:
        obj["choices"].append({"index": 1, "delta": {"content": "second choice is ignored"}})
    ensure_ascii = rng.random() < 0.5
    seps = rng.choice([None, (",", ":"), (", ", ": "), (" ,\t", " :\n ")])
    return "data: " + json.dumps(obj, ensure_ascii=ensure_ascii, separators=seps)


def _good_chunks(seed, n):
    rng = random.Random(seed)
    chunks = []
    for _ in range(n):
        evs = []
        for _ in range(rng.randrange(0, 5)):
            content = "".join(rng.choice(ALPHA) for _ in range(rng.randrange(0, 20)))
            evs.append(_event(rng, content, rng.randrange(0, 4)))
            if rng.random() < 0.1:
                evs.append("data: [DONE]")
            if rng.random() < 0.1:
                evs.append(": keep-alive comment")
            if rng.random() < 0.05:
                evs.append("")
        text = rng.choice(["", " ", "\n", "\n\n", "\u2003"]) + "\n\n".join(evs) + rng.choice(["\n\n", "\n", "", " \n\n\n", "\n\n\n"])
        chunks.append(text.encode("utf-8", "surrogatepass") if False else text.encode("utf-8", "replace"))
    return chunks


ODD = [
    b"data: {\"choices\": []}\n\n", b"data: {\"choices\": [{}]}\n\n", b"data: {\"choices\": [{\"delta\": {}}]}\n\n",
    b"data: {\"nochoices\": 1}\n\n", b"data: {}\n\n", b"data: [DONE]\n\n", b"data:[DONE]\n\n", b"data:  [DONE] \n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"a\"}}]} trailing\n\n", b"data: {\"choices\": [{\"delta\": {\"content\": \"a\"}}]\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"a\\qb\"}}]}\n\n", b"data: {\"choices\": [{\"delta\": {\"content\": \"a\nb\"}}]}\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"x\", \"content\": \"y\"}}]}\n\n",
    b"data: {\"cho\\u0069ces\": [{\"delta\": {\"content\": \"esc key\"}}]}\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"\\ud83d\\ude00 pair \\ud83d lone \\u00e9\"}}]}\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"n\"}}], \"v\": 01}\n\n", b"data: {\"choices\": [{\"delta\": {\"content\": \"n\"}}], \"v\": NaN}\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"n\"}}], \"v\": -Infinity}\n\n", b"data: {\"choices\": [{\"delta\": {\"content\": \"n\"}}],}\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"deep\"}}], \"d\": " + b"[" * 40 + b"]" * 40 + b"}\n\n",
    b"\xff\xfe broken utf8 data: {}\n\n", b"data: {\"choices\": [{\"delta\": {\"content\": \"ok\"}}]}\n\n\xc3", b"\xed\xa0\x80",
    b"event: message\ndata: {\"choices\": [{\"delta\": {\"content\": \"not at start\"}}]}\n\n", b" data: {\"choices\": []}\n\n",
    b"\n\n\ndata: {\"choices\": [{\"delta\": {\"content\": \"after three newlines\"}}]}\n\n\n\ndata: {\"choices\": [{\"delta\": {\"content\": \"b\"}}]}",
    b"", b"   ", b"\n\n", b"data: ", b"data: \"just a string\"\n\n", b"data: [1,2]\n\n", b"data: [\"choices\"]\n\n" if False else b"data: []\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": null}}]}\n\n", b"data: {\"choices\": null}\n\n", b"data: {\"choices\": 0, \"a\": \"b\"}\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"tab\\there \\/ \\b\\f\\r\\n \\\" \\\\\"}}]}\n\n",
    b"data: {\"choices\": [{\"delta\": {\"content\": \"\\u0041\\u00E9\\uFFFF\\u0000\"}}]}\n\n",
]
# these end the backend's stream in the reference (TypeError / AttributeError / KeyError out of the parsing block): the
# device must hand them to the host, which then raises the same way
RAISING = [
    b"data: {\"choices\": [5]}\n\n", b"data: 5\n\n", b"data: null\n\n", b"data: true\n\n",
    b"data: {\"choices\": {\"a\": 1}}\n\n", b"data: {\"choices\": [{\"delta\": 7}]}\n\n", b"data: {\"choices\": \"abc\"}\n\n",
]


def test_host_parser_is_the_reference_code_path():
    """parse_sse_chunk against a literal transcription of oai_proxy.py:580-617."""
    from quorum_b200.streaming import parse_sse_chunk

    def reference(chunk):
        try:
            chunk_decoded = chunk.decode()
        except UnicodeDecodeError:
            return None
        out = []
        events = chunk_decoded.strip().split("\n\n")
        for event in events:
            if not event.strip():
                continue
            if event.startswith("data: "):
                event_data = event[6:].strip()
                if event_data == "[DONE]":
                    continue
                try:
                    parsed = json.loads(event_data)
                    if "choices" in parsed and parsed["choices"]:
                        delta = parsed["choices"][0].get("delta", {})
                        if "content" in delta:
                            out.append(delta["content"])
                except json.JSONDecodeError:
                    continue
        return out
    for c in _good_chunks(1, 300) + ODD:
        assert parse_sse_chunk(c) == reference(c), c


@pytest.mark.gpu
def test_gpu_parser_matches_reference_code_path():
    import quorum_b200 as qb
    from quorum_b200.streaming import parse_sse_chunk
    chunks = _good_chunks(2, 3000) + ODD
    got = qb.sse.parse_http_chunks(chunks)
    for c, g in zip(chunks, got):
        assert g == parse_sse_chunk(c), c


@pytest.mark.gpu
def test_gpu_parser_kinds():
    """The conservative classification itself: canonical events are handled on the device, doubtful ones go to the host."""
    import numpy as np
    import torch
    import quorum_b200 as qb
    from quorum_b200 import batch
    dev = torch.device("cuda", 0)

    def kinds(chunk):
        data, off = batch.pack_host([chunk])
        out, out_off, kind, ev_first, n_host = qb.sse.parse_http_chunks_device(data.to(dev), off.to(dev))
        return kind.cpu().numpy().tolist(), int(n_host.item())
    assert kinds(b"data: {\"choices\": [{\"delta\": {\"content\": \"a\"}}]}\n\ndata: [DONE]\n\n") == ([1, 0], 0)
    assert kinds(b"data: {\"choices\": []}\n\n: comment\n\ndata: {\"x\": {\"choices\": 1}}") == ([0, 0, 0], 0)
    for c in RAISING + [b"data: {\"choices\": [{\"delta\": {\"content\": null}}]}", b"data: {\"a\": NaN}", b"data: {\"a\": 1} x",
                        b"data: {\"choices\": [{\"delta\": {\"content\": \"x\", \"content\": \"y\"}}]}"]:
        k, n = kinds(c)
        assert k == [2] and n == 1, c
    assert kinds(b"\xff")[0] == []
    # device content of a canonical batch equals json.loads, byte for byte (lone surrogates in surrogatepass form)
    rng = random.Random(9)
    contents = ["".join(rng.choice(ALPHA) for _ in range(rng.randrange(0, 40))) for _ in range(500)]
    chunk = "\n\n".join("data: " + json.dumps({"choices": [{"delta": {"content": t}}]}) for t in contents).encode()
    data, off = batch.pack_host([chunk])
    out, out_off, kind, ev_first, n_host = qb.sse.parse_http_chunks_device(data.to(dev), off.to(dev))
    assert int(n_host.item()) == 0 and kind.cpu().numpy().tolist() == [1] * 500
    pieces = batch.split_host(out, out_off)
    # (json.loads(json.dumps(t)) rather than t: two adjacent lone surrogates come back as one code point)
    assert pieces == [json.loads(json.dumps(t)).encode("utf-8", "surrogatepass") for t in contents]


@pytest.mark.gpu
def test_gpu_parser_raising_events_raise_like_the_reference():
    import quorum_b200 as qb
    from quorum_b200.streaming import parse_sse_chunk
    for c in RAISING:
        with pytest.raises(Exception) as e_ref:
            parse_sse_chunk(c)
        with pytest.raises(type(e_ref.value)):
            qb.sse.parse_http_chunks([c])


@pytest.mark.gpu
def test_gpu_parser_mutation_fuzz():
    """Random byte edits of valid events: whatever json.loads makes of the result (accept, reject, raise later), the batch
    parser must do the same.  The device may only answer for an event it is certain about."""
    import quorum_b200 as qb
    from quorum_b200.streaming import parse_sse_chunk
    rng = random.Random(77)
    junk = [b'"', b"\\", b"{", b"}", b"[", b"]", b",", b":", b" ", b"\n", b"\n\n", b"0", b"-", b".", b"e", b"null", b"true", b"\\u12", b"\\ud83d",
            b"\xc3\xa9", b"\xff", b"choices", b"\"content\":", b"\"delta\":", b"data: ", b"NaN", b"\t", b"\x00", b"\x1f"]
    chunks = []
    for base in _good_chunks(5, 4000):
        b = bytearray(base)
        for _ in range(rng.randrange(1, 4)):
            if not b:
                break
            p = rng.randrange(0, len(b))
            op = rng.randrange(0, 3)
            if op == 0:
                del b[p:p + rng.randrange(1, 4)]
            elif op == 1:
                b[p:p] = rng.choice(junk)
            else:
                b[p:p + 1] = rng.choice(junk)
        chunks.append(bytes(b))

    def host(c):
        try:
            return ("ok", parse_sse_chunk(c))
        except Exception as e:  # noqa: BLE001 - the reference lets these end the backend's stream
            return ("raise", type(e).__name__)
    expected = [host(c) for c in chunks]
    clean = [c for c, e in zip(chunks, expected) if e[0] == "ok"]
    got = qb.sse.parse_http_chunks(clean)
    for c, g, e in zip(clean, got, [e for e in expected if e[0] == "ok"]):
        assert g == e[1], c
    for c, e in zip(chunks, expected):
        if e[0] == "raise":
            with pytest.raises(Exception) as ei:
                qb.sse.parse_http_chunks([c])
            assert type(ei.value).__name__ == e[1], c


@pytest.mark.gpu
def test_device_stream_pipeline_matches_the_reference_loop():
    """HTTP chunks in, SSE events out, three device steps in a row: against the reference's per-chunk loop replayed with its
    own pieces (parse_sse_chunk = :578-617, the oracle filter = :278-350, json.dumps framing = :629-652)."""
    import quorum_b200 as qb
    from oracle import quorum_oracle as o
    from quorum_b200.streaming import parse_sse_chunk
    rng = random.Random(13)
    tags = ["think", "reason", "reasoning", "thought"]
    n, created = 40, 1712000000
    backend = [s % 3 for s in range(n)]
    pieces = ["Hello ", "<thi", "nk>", "secret", "</th", "ink>", " world", "é日本", "<", "reason>", "</reason>", "\n", '"q"', "\\", "x<y "]
    # per stream a sequence of deltas, cut into HTTP chunks of 0-3 events, some with comments / [DONE] / broken JSON
    streams = []
    for s in range(n):
        deltas = [rng.choice(pieces) for _ in range(rng.randrange(5, 40))]
        chunks, k = [], 0
        while k < len(deltas):
            m = rng.randrange(0, 4)
            evs = ["data: " + json.dumps({"choices": [{"index": 0, "delta": {"content": t}}]}, ensure_ascii=rng.random() < 0.5)
                   for t in deltas[k:k + m]]
            if rng.random() < 0.1:
                evs.append(": ping")
            if rng.random() < 0.05:
                evs.append("data: {\"choices\": [{\"delta\": {\"content\": \"broken")
            if rng.random() < 0.05:
                evs.append("data: {\"choices\": [{\"delta\": {\"content\": \"dup\", \"content\": \"last wins\"}}]}")
            chunks.append(("\n\n".join(evs) + "\n\n").encode())
            k += m
        chunks.append(b"data: [DONE]\n\n")
        streams.append(chunks)
    pipe = qb.sse.DeviceStreamPipeline(tags, backend, created)
    filters = [o.ThinkingTagFilter(tags) for _ in range(n)]
    content = ["" for _ in range(n)]
    for step in range(max(len(c) for c in streams)):
        batch_chunks = [c[step] if step < len(c) and rng.random() < 0.9 else None for c in streams]
        for s, c in enumerate(batch_chunks):  # a chunk that was withheld this step comes in the next one
            if c is None and step < len(streams[s]):
                streams[s].insert(step, None)
        got = pipe.step(batch_chunks)
        for s, c in enumerate(batch_chunks):
            exp = []
            if c is not None:
                for delta in parse_sse_chunk(c) or []:
                    safe = filters[s].feed(delta)
                    content[s] += safe
                    if safe:
                        ev = {"id": f"chatcmpl-parallel-{backend[s]}", "object": "chat.completion.chunk", "created": created,
                              "model": "parallel-proxy", "choices": [{"index": 0, "delta": {"content": safe}, "finish_reason": None}]}
                        exp.append(f"data: {json.dumps(ev)}\n\n".encode())
            assert got[s] == exp, (step, s, c)
    assert pipe.finish() == content
