"""BASELINE.json configs[0] through the reference's own endpoint: the unmodified proxy (staged copy, oracle/_ref) serves a
request over two mocked backends, once as it is and once with quorum_b200.install() rebinding its filter - the responses
must be identical.  Mirrors the reference's tests/test_parallel_backends.py (mocked httpx.AsyncClient.post, patched
config), non-streaming and streaming."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = os.path.join(ROOT, "oracle", "_ref", "src")

CONTENTS = [
    "<think>Let me think about this... 2+2 is a basic arithmetic operation</think>The answer is 4.",
    "<think>First, let me consider this carefully</think>The answer is 4.<reason>This is because 2+2 has always equaled 4 in "
    "base-10 arithmetic</reason>",
    "  plain answer with é and 日本語, an unclosed <think>tail",
    "<THINK>upper</think> mixed <reasoning>r</reasoning> done <thought>t</Thought> \n",
]
CONFIG = {
    "primary_backends": [{"name": "LLM1", "url": "http://test1.example.com/v1", "model": "gpt-4-1"},
                         {"name": "LLM2", "url": "http://test2.example.com/v1", "model": "gpt-4-2"}],
    "iterations": {"aggregation": {"strategy": "concatenate"}},
    "strategy": {"concatenate": {"separator": "\n-------------\n", "hide_intermediate_think": True, "hide_final_think": True,
                                 "thinking_tags": ["think", "reason", "reasoning", "thought"], "skip_final_aggregation": False}},
    "settings": {"timeout": 30},
}


def _completion(content, k):
    return {"id": f"chatcmpl-{k}", "object": "chat.completion", "created": 1677652288, "model": "gpt-4o-mini", "system_fingerprint": "fp",
            "choices": [{"index": 0, "message": {"role": "assistant", "content": content}, "logprobs": None, "finish_reason": "stop"}],
            "usage": {"prompt_tokens": 8 + k, "completion_tokens": 20 + k, "total_tokens": 28 + 2 * k}}


class _Resp:
    def __init__(self, payload):
        self.status_code, self._p, self.headers = 200, payload, {"content-type": "application/json"}

    async def aread(self):
        return json.dumps(self._p).encode()

    def json(self):
        return self._p


class _StreamResp:
    """A backend's SSE stream: the content cut into deltas so that tags straddle events."""

    def __init__(self, content):
        self.status_code, self.headers = 200, {"content-type": "text/event-stream"}
        cuts = list(range(0, len(content), 7)) + [len(content)]
        self._events = [b'data: {"choices":[{"delta":{"role":"assistant"}}]}\n\n']
        self._events += [f"data: {json.dumps({'choices': [{'delta': {'content': content[a:b]}}]})}\n\n".encode() for a, b in zip(cuts, cuts[1:])]
        self._events += [b'data: {"choices":[{"delta":{},"finish_reason":"stop"}]}\n\n', b"data: [DONE]\n\n"]

    async def aiter_bytes(self):
        for e in self._events:
            yield e

    def __aiter__(self):
        self._i = 0
        return self

    async def __anext__(self):
        if self._i >= len(self._events):
            raise StopAsyncIteration()
        self._i += 1
        return self._events[self._i - 1]

    def json(self):
        return {"stream": True}


def _normalise(body: bytes):
    """Events of a streamed response with the one clock-dependent field removed."""
    out = []
    for ev in body.decode().strip().split("\n\n"):
        if ev.startswith("data: ") and ev[6:].strip() != "[DONE]":
            d = json.loads(ev[6:])
            d.pop("created", None)
            out.append(d)
        else:
            out.append(ev)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("pair", [(0, 1), (2, 3), (1, 2)])
def test_endpoint_identical_with_and_without_install(pair, monkeypatch):
    if not os.path.isfile(os.path.join(REF_SRC, "quorum", "oai_proxy.py")):
        pytest.skip("staged reference (oracle/_ref) not present")
    import logging
    logging.disable(logging.CRITICAL)
    sys.path.insert(0, REF_SRC)
    import httpx
    from fastapi.testclient import TestClient
    import quorum.oai_proxy as proxy
    import quorum_b200

    streaming = {"on": False}

    async def mock_post(self, url, *args, **kwargs):
        k = pair[0] if "test1.example.com" in str(url) else pair[1]
        return _StreamResp(CONTENTS[k]) if streaming["on"] else _Resp(_completion(CONTENTS[k], k))

    monkeypatch.setattr(proxy, "config", CONFIG)
    monkeypatch.setattr(httpx.AsyncClient, "post", mock_post)
    client = TestClient(proxy.app)
    body = {"messages": [{"role": "user", "content": "What is 2+2?"}]}
    hdr = {"Authorization": "Bearer test-key"}

    def run():
        streaming["on"] = False
        plain = client.post("/chat/completions", json=body, headers=hdr)
        streaming["on"] = True
        streamed = client.post("/chat/completions", json=dict(body, stream=True), headers=hdr)
        return plain.status_code, plain.json(), streamed.status_code, _normalise(streamed.content)

    expected = run()                                   # the reference as it is
    previous = quorum_b200.install(proxy)              # proxy.strip_thinking_tags / proxy.ThinkingTagFilter -> GPU
    try:
        assert proxy.strip_thinking_tags is quorum_b200.strip_thinking_tags
        got = run()
    finally:
        quorum_b200.uninstall(previous, proxy)
    assert expected[0] == 200 and expected[2] == 200
    assert got == expected
    assert "<think>" not in json.dumps(expected[1]) or 2 in pair  # (content 2 ends in an unclosed tag, which the regex keeps)
