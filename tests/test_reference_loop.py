"""Rows a5, a8 and f2 of SURVEY.md section 8 pinned to the reference EXECUTED: the unmodified proxy (staged copy,
oracle/_ref) serves streaming and non-streaming requests over mocked backends, and the repo's drop-ins must reproduce
what it sent, byte for byte (the clock-dependent `created` field normalised):

  * the reference's own loop  src/quorum/oai_proxy.py:489-885 (progress_streaming_aggregator: per-backend filters
    :543-545, decode -> split -> `data: ` -> `[DONE]` -> json.loads -> feed -> content += -> emit :578-659, flush
    :739-741, final join :759-764 + 841, error event :863-881) against
      (i)  quorum_b200.streaming.ConcatenateStreamSession  (host parsing, GPU filter, GPU join)
      (ii) quorum_b200.sse.DeviceStreamPipeline            (GPU parsing, GPU filter, GPU framing) + frame_final_event
  * the flag plumbing  :1049-1075 (streaming) and :1166-1199 + 1286 (non-streaming) against
    ConcatenateStrategy.from_config built from the same config dict the proxy reads.

The HTTP chunks are the generated / odd / raising sets of tests/test_sse_parsing.py plus streams whose tags straddle
events and chunks, a backend whose iterator raises mid-stream, a backend that answers 500, and a `content: null` delta.
"""
import itertools
import json
import os
import random
import re
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = os.path.join(ROOT, "oracle", "_ref", "src")
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

TAGS = ["think", "reason", "reasoning", "thought"]


def _config(strategy_block, n_backends=3):
    return {
        "primary_backends": [{"name": f"LLM{i + 1}", "url": f"http://test{i + 1}.example.com/v1", "model": f"m{i + 1}"}
                             for i in range(n_backends)],
        "iterations": {"aggregation": {"strategy": "concatenate"}},
        "strategy": {"concatenate": strategy_block},
        "settings": {"timeout": 30},
    }


class _StreamResp:
    """A backend's SSE stream given as HTTP chunks; `raise_at` = index of the chunk whose read raises."""

    def __init__(self, chunks, raise_at=None):
        self.status_code, self.headers = 200, {"content-type": "text/event-stream"}
        self._chunks, self._raise_at = chunks, raise_at

    async def aiter_bytes(self):
        for k, c in enumerate(self._chunks):
            if k == self._raise_at:
                raise RuntimeError("connection reset by peer")
            yield c
        if self._raise_at is not None and self._raise_at >= len(self._chunks):
            raise RuntimeError("connection reset by peer")

    def json(self):
        return {"stream": True}


class _Resp:
    def __init__(self, payload, status=200):
        self.status_code, self._p, self.headers = status, payload, {"content-type": "application/json"}

    async def aread(self):
        return json.dumps(self._p).encode()

    def json(self):
        return self._p


def _completion(content, k):
    return {"id": f"chatcmpl-{k}", "object": "chat.completion", "created": 1677652288, "model": "gpt-4o-mini",
            "choices": [{"index": 0, "message": {"role": "assistant", "content": content}, "finish_reason": "stop"}],
            "usage": {"prompt_tokens": 8 + k, "completion_tokens": 20 + k, "total_tokens": 28 + 2 * k}}


def _split_events(body: bytes):
    """The SSE events of a streamed response, `created` normalised to 0 (the reference stamps int(loop.time()))."""
    body = re.sub(rb'"created": \d+', b'"created": 0', body)
    assert body.endswith(b"\n\n")
    return [e + b"\n\n" for e in body[:-2].split(b"\n\n")]


def _by_id(events):
    """{event id: [events in order]} plus the order in which ids first appear."""
    groups, order = {}, []
    for e in events:
        if e == b"data: [DONE]\n\n":
            key = "[DONE]"
        else:
            key = json.loads(e[6:])["id"]
        if key not in groups:
            order.append(key)
        groups.setdefault(key, []).append(e)
    return groups, order


@pytest.fixture(scope="module")
def ref():
    if not os.path.isfile(os.path.join(REF_SRC, "quorum", "oai_proxy.py")):
        pytest.skip("staged reference (oracle/_ref) not present")
    import logging
    logging.disable(logging.CRITICAL)
    sys.path.insert(0, REF_SRC)
    import quorum.oai_proxy as proxy
    from fastapi.testclient import TestClient
    return proxy, TestClient(proxy.app)


def _delta_chunks(rng, deltas, style=0):
    """`deltas` as the HTTP chunks of one backend: 0-3 events per chunk, with the noise real streams carry."""
    chunks, k = [b'data: {"choices":[{"delta":{"role":"assistant"}}]}\n\n'], 0
    while k < len(deltas):
        m = rng.randrange(0, 4)
        evs = ["data: " + json.dumps({"id": "x", "choices": [{"index": 0, "delta": {"content": t}, "finish_reason": None}]},
                                     ensure_ascii=rng.random() < 0.5) for t in deltas[k:k + m]]
        if rng.random() < 0.15:
            evs.append(": keep-alive")
        if rng.random() < 0.1:
            evs.append('data: {"choices": [{"delta": {"content": "broken')
        if rng.random() < 0.1:
            evs.append('data: {"choices": [{"delta": {"content": "dup", "content": "last wins"}}]}')
        chunks.append((rng.choice(["", "\n", " "]) + "\n\n".join(evs) + "\n\n").encode())
        k += m
    chunks += [b'data: {"choices":[{"delta":{},"finish_reason":"stop"}]}\n\n', b"data: [DONE]\n\n"]
    return chunks


PIECES = ["Hello ", "<thi", "nk>", "secret", "</th", "ink>", " world", "é日本", "<", "reason>", "</reason>", "\n", '"q"', "\\",
          "x<y ", "<THINK>", "</Thought>", "<thought>", "  ", " ", "<reasoning>deep</reasoning>", "</think>", "🎉"]


def _scenarios():
    """(name, per-backend spec) with spec = dict(chunks=[...], raise_at=None, status=200)."""
    from test_sse_parsing import ODD, RAISING, _good_chunks
    rng = random.Random(2024)

    def deltas(n):
        return [rng.choice(PIECES) for _ in range(n)]
    out = []
    out.append(("tags straddle events", [dict(chunks=_delta_chunks(rng, deltas(60))), dict(chunks=_delta_chunks(rng, deltas(45))),
                                         dict(chunks=_delta_chunks(rng, deltas(30)))]))
    out.append(("one backend raises mid-stream, one answers 500",
                [dict(chunks=_delta_chunks(rng, deltas(40)), raise_at=7), dict(chunks=[], status=500),
                 dict(chunks=_delta_chunks(rng, ["<think>a</think>", "kept ", "<think>unclosed"]))]))
    out.append(("generated chunks in every json.dumps style", [dict(chunks=_good_chunks(31, 40)), dict(chunks=_good_chunks(32, 25)),
                                                               dict(chunks=_good_chunks(33, 10))]))
    out.append(("odd chunks", [dict(chunks=list(ODD[:len(ODD) // 2])), dict(chunks=list(ODD[len(ODD) // 2:])),
                               dict(chunks=_delta_chunks(rng, deltas(12)))]))
    # every raising payload ends exactly its own backend's stream, after the events that came before it in the chunk
    for k, bad in enumerate(RAISING + [b'data: {"choices": [{"delta": {"content": null}}]}\n\n',
                                       b'data: {"choices": [{"delta": {"content": 5}}]}\n\n']):
        good = b'data: {"choices": [{"delta": {"content": "before <think>x</think>"}}]}\n\n'
        after = b'data: {"choices": [{"delta": {"content": " after"}}]}\n\n'
        out.append((f"raising payload {k}", [dict(chunks=[good + bad + after, after]),
                                             dict(chunks=_delta_chunks(rng, deltas(8))), dict(chunks=[good, after])]))
    out.append(("nobody has content", [dict(chunks=[b"data: [DONE]\n\n"]), dict(chunks=[], status=503),
                                       dict(chunks=[b'data: {"choices": [5]}\n\n'])]))
    # mutation set: random byte edits of valid chunks (what json.loads makes of them decides)
    junk = [b'"', b"\\", b"{", b"}", b"[", b"]", b",", b":", b" ", b"\n\n", b"0", b"null", b"\\ud83d", b"\xc3\xa9", b"\xff", b"choices",
            b'"content":', b"data: ", b"NaN", b"\x00"]
    muts = []
    for base in _good_chunks(77, 90):
        b = bytearray(base)
        for _ in range(rng.randrange(1, 3)):
            if b:
                p = rng.randrange(0, len(b))
                if rng.random() < 0.5:
                    del b[p:p + rng.randrange(1, 4)]
                else:
                    b[p:p] = rng.choice(junk)
        muts.append(bytes(b))
    out.append(("mutated chunks", [dict(chunks=muts[:30]), dict(chunks=muts[30:60]), dict(chunks=muts[60:])]))
    return out


def _run_reference_stream(ref, monkeypatch, config, specs, body):
    proxy, client = ref
    import httpx

    async def mock_post(self, url, *args, **kwargs):
        i = int(re.search(r"test(\d+)\.example\.com", str(url)).group(1)) - 1
        sp = specs[i]
        if sp.get("status", 200) != 200:
            return _Resp({"error": {"message": "backend down"}}, sp["status"])
        return _StreamResp(sp["chunks"], sp.get("raise_at"))
    monkeypatch.setattr(proxy, "config", config)
    monkeypatch.setattr(httpx.AsyncClient, "post", mock_post)
    r = client.post("/chat/completions", json=dict(body, stream=True), headers={"Authorization": "Bearer k"})
    assert r.status_code == 200
    return _split_events(r.content)


def _feed_rounds(specs, feed, abandon):
    """Round k hands chunk k of every live backend to `feed`; a backend whose iterator raises at k is abandoned there."""
    n = len(specs)
    events = [[] for _ in range(n)]
    live = [sp.get("status", 200) == 200 for sp in specs]
    rounds = max((len(sp["chunks"]) + 1 for sp in specs), default=0)
    for k in range(rounds):
        chunks = []
        for i, sp in enumerate(specs):
            if live[i] and sp.get("raise_at") is not None and k == min(sp["raise_at"], len(sp["chunks"])):
                abandon(i)
                live[i] = False
            chunks.append(sp["chunks"][k] if live[i] and k < len(sp["chunks"]) else None)
        for i, evs in enumerate(feed(chunks)):
            events[i].extend(evs)
    return events


def _check_stream(expected, per_backend, tail, suppress):
    import quorum_b200.streaming as st
    groups, order = _by_id(expected)
    assert expected[0] == st.role_event(0) and order[0] == "chatcmpl-parallel"
    for i, evs in enumerate(per_backend):
        assert groups.get(f"chatcmpl-parallel-{i}", []) == evs, f"backend {i}"
        if suppress:
            assert evs == []
    n_tail = len(tail)
    assert expected[-n_tail:] == tail
    assert len(expected) == 1 + sum(len(e) for e in per_backend) + n_tail


FLAGS = list(itertools.product([True, False], repeat=4))  # hide_intermediate, hide_final, skip_final, suppress


@pytest.mark.parametrize("flags", FLAGS, ids=lambda f: "".join("HFSU"[k] if v else "-" for k, v in enumerate(f)))
def test_streaming_flag_matrix_against_the_reference_endpoint(ref, monkeypatch, flags):
    """All 16 combinations of the four streaming flags on one busy scenario; the strategy object is built by
    ConcatenateStrategy.from_config from the config dict the proxy itself reads (oai_proxy.py:1049-1075)."""
    import quorum_b200 as qb
    import quorum_b200.streaming as st
    hide_i, hide_f, skip, suppress = flags
    block = {"separator": "\n-------------\n", "hide_intermediate_think": hide_i, "hide_final_think": hide_f,
             "thinking_tags": TAGS, "skip_final_aggregation": skip, "suppress_individual_responses": suppress}
    config = _config(block)
    name, specs = _scenarios()[1] if skip else _scenarios()[0]
    body = {"messages": [{"role": "user", "content": "q"}]}
    expected = _run_reference_stream(ref, monkeypatch, config, specs, body)
    strategy = qb.ConcatenateStrategy.from_config(config, request_body=body)
    # (i) host parsing + GPU filter + GPU join
    sess = qb.ConcatenateStreamSession(strategy, len(specs))
    per = _feed_rounds(specs, lambda ch: sess.feed_http_chunks_events(ch, 0), sess.abandon_backend)
    _check_stream(expected, per, sess.finish_events(0), suppress)
    # (ii) everything on the device
    pipe = qb.sse.DeviceStreamPipeline(strategy.thinking_tags, list(range(len(specs))), 0,
                                       hide_intermediate_think=strategy.hide_intermediate_think,
                                       suppress_individual_responses=strategy.suppress_individual_responses)
    per = _feed_rounds(specs, lambda ch: pipe.step(ch, created=0), pipe.abandon)
    _check_stream(expected, per, qb.sse.tail_events_device(strategy, pipe.finish(), 0), suppress)
    # the per-request override of suppress_individual_responses (oai_proxy.py:1072-1075)
    body2 = dict(body, suppress_individual_responses=not suppress)
    expected2 = _run_reference_stream(ref, monkeypatch, config, specs, body2)
    strategy2 = qb.ConcatenateStrategy.from_config(config, request_body=body2)
    assert strategy2.suppress_individual_responses == (not suppress)
    sess2 = qb.ConcatenateStreamSession(strategy2, len(specs))
    per2 = _feed_rounds(specs, lambda ch: sess2.feed_http_chunks_events(ch, 0), sess2.abandon_backend)
    _check_stream(expected2, per2, sess2.finish_events(0), not suppress)
    assert st.DONE_EVENT == expected[-1]


def test_streaming_scenarios_against_the_reference_endpoint(ref, monkeypatch):
    """Every scenario (odd, raising, mutated chunks; failing backends) with the reference's shipped flag defaults and with
    everything hidden; config keys left out fall back to the reference's defaults in both implementations."""
    import quorum_b200 as qb
    for block in ({}, {"separator": " § ", "hide_intermediate_think": True, "hide_final_think": True,
                       "thinking_tags": ["think", "reason", "reasoning", "thought", "Thought"]}):
        config = _config(block)
        for name, specs in _scenarios():
            body = {"messages": [{"role": "user", "content": "q"}]}
            expected = _run_reference_stream(ref, monkeypatch, config, specs, body)
            strategy = qb.ConcatenateStrategy.from_config(config, request_body=body)
            sess = qb.ConcatenateStreamSession(strategy, len(specs))
            per = _feed_rounds(specs, lambda ch: sess.feed_http_chunks_events(ch, 0), sess.abandon_backend)
            _check_stream(expected, per, sess.finish_events(0), False)
            sess_d = qb.ConcatenateStreamSession(strategy, len(specs), device_parse=True)   # host session, device parser
            per = _feed_rounds(specs, lambda ch: sess_d.feed_http_chunks_events(ch, 0), sess_d.abandon_backend)
            _check_stream(expected, per, sess_d.finish_events(0), False)
            pipe = qb.sse.DeviceStreamPipeline(strategy.thinking_tags, list(range(len(specs))), 0,
                                               hide_intermediate_think=strategy.hide_intermediate_think)
            per = _feed_rounds(specs, lambda ch: pipe.step(ch, created=0), pipe.abandon)
            _check_stream(expected, per, qb.sse.tail_events_device(strategy, pipe.finish(), 0), False)


NONSTREAM_CONTENTS = [
    "<think>Let me think... 2+2</think>The answer is 4.",
    "  <reason>because</reason>\n The answer is 4. <thought>unclosed",
    "",
    "<THINK>upper</think> é日本 <reasoning>r</reasoning> done  ",
]


@pytest.mark.parametrize("block", [
    {},                                                                     # every key left to the reference's defaults
    {"separator": "\n-------------\n", "hide_final_think": True},
    {"separator": "\n\n—— § ——\n\n", "hide_final_think": True, "hide_intermediate_think": False, "skip_final_aggregation": True,
     "thinking_tags": ["think", "reason", "reasoning", "thought", "Thought"]},
    {"separator": "", "hide_final_think": False, "skip_final_aggregation": True, "thinking_tags": ["think"]},
], ids=["defaults", "hide", "multibyte-skip", "nosep"])
def test_nonstream_combine_against_the_reference_endpoint(ref, monkeypatch, block):
    """ConcatenateStrategy.combine vs the endpoint's combined message (oai_proxy.py:1166-1199 + 1286): the non-streaming
    branch does not read skip_final_aggregation, keeps empty contents, leaves failed backends out."""
    import quorum_b200 as qb
    import httpx
    proxy, client = ref
    config = _config(block, n_backends=4)
    status = [200, 200, 500, 200]

    async def mock_post(self, url, *args, **kwargs):
        i = int(re.search(r"test(\d+)\.example\.com", str(url)).group(1)) - 1
        if status[i] != 200:
            return _Resp({"error": {"message": "down"}}, status[i])
        return _Resp(_completion(NONSTREAM_CONTENTS[i], i))
    monkeypatch.setattr(proxy, "config", config)
    monkeypatch.setattr(httpx.AsyncClient, "post", mock_post)
    r = client.post("/chat/completions", json={"messages": [{"role": "user", "content": "q"}]}, headers={"Authorization": "Bearer k"})
    assert r.status_code == 200
    expected = r.json()["choices"][0]["message"]["content"]
    strategy = qb.ConcatenateStrategy.from_config(config)
    ok = [c for c, s in zip(NONSTREAM_CONTENTS, status) if s == 200]
    assert strategy.combine(ok) == expected
    assert strategy.combine_many([ok, ok[:1], []])[:2] == [expected, strategy.combine(ok[:1])]


@pytest.mark.parametrize("hide_final", [None, True, False], ids=["shipped-default", "hide", "keep"])
def test_aggregate_prompt_from_config_against_the_reference_endpoint(ref, monkeypatch, hide_final):
    """AggregatePromptBuilder.from_config vs the prompt the reference's endpoint really sends to the aggregator backend
    (oai_proxy.py:1166-1199 filter with hide_final_think from the ACTIVE strategy block, :1204-1268 + 406-423 prompt),
    with the aggregate block of the reference's shipped config.yaml (template with doubled braces included)."""
    import quorum_b200 as qb
    import httpx
    proxy, client = ref
    aggregate = {
        "source_backends": ["LLM1", "LLM2", "LLM3"], "aggregator_backend": "LLM1", "intermediate_separator": "\n\n---\n\n",
        "include_source_names": hide_final is True, "source_label_format": "Response from {backend_name}:\n",
        "prompt_template": "You have received the following responses regarding the user's query:\n\n{{intermediate_results}}\n\n"
                           "Synthesize these responses into a single, comprehensive answer.\n",
        "strip_intermediate_thinking": True, "hide_aggregator_thinking": True,
        "thinking_tags": ["think", "reason", "reasoning", "thought", "Thought"], "include_original_query": True,
        "query_format": "Original query: {query}\n\n",
    }
    if hide_final is not None:
        aggregate["hide_final_think"] = hide_final
    config = _config({"separator": "\n\nSEPARATOR\n\n", "hide_intermediate_think": False, "hide_final_think": False,
                      "thinking_tags": ["think"], "skip_final_aggregation": True})
    config["iterations"]["aggregation"]["strategy"] = "aggregate"
    config["strategy"]["aggregate"] = aggregate
    contents = NONSTREAM_CONTENTS[:3]
    query = "What is 2+2? <think>not a tag here</think>"
    seen = []

    async def mock_post(self, url, *args, **kwargs):
        body = json.loads(kwargs["content"])
        text = body["messages"][0]["content"]
        if text != query:                      # the aggregator call: its single message is the prompt
            seen.append(text)
            return _Resp(_completion("synthesised", 9))
        i = int(re.search(r"test(\d+)\.example\.com", str(url)).group(1)) - 1
        return _Resp(_completion(contents[i], i))
    monkeypatch.setattr(proxy, "config", config)
    monkeypatch.setattr(httpx.AsyncClient, "post", mock_post)
    r = client.post("/chat/completions", json={"messages": [{"role": "user", "content": query}]}, headers={"Authorization": "Bearer k"})
    assert r.status_code == 200 and len(seen) == 1
    builder = qb.AggregatePromptBuilder.from_config(config)
    assert builder.hide_final_think == bool(hide_final)
    assert builder.build_prompts([contents], [query]) == seen
