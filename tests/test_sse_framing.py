"""SSE output framing: the oracle and the CUDA path against CPython's json.dumps, which is
what the reference calls (src/quorum/oai_proxy.py:646, 860)."""
import json
import random

import numpy as np
import pytest

ALPHA = [chr(c) for c in list(range(0, 0x30)) + [0x5C, 0x7E, 0x7F, 0x80, 0xE9, 0x7FF, 0x800, 0x2028, 0xD7FF, 0xE000, 0xFFFF,
                                                 0x10000, 0x1F389, 0x10FFFF]] + ["a", '"', "\\", "日", "\ud800", "\udfff"]


def _reference_event(i, created, text):
    """Restates oai_proxy.py:629-646 literally."""
    stream_event = {
        "id": f"chatcmpl-parallel-{i}",
        "object": "chat.completion.chunk",
        "created": created,
        "model": "parallel-proxy",
        "choices": [{"index": 0, "delta": {"content": text}, "finish_reason": None}],
    }
    return f"data: {json.dumps(stream_event)}\n\n".encode()


def _reference_final(created, text):
    """Restates oai_proxy.py:847-860 literally."""
    final_event = {
        "id": "chatcmpl-parallel-final",
        "object": "chat.completion.chunk",
        "created": created,
        "model": "parallel-proxy",
        "choices": [{"index": 0, "delta": {"content": text}, "finish_reason": "stop"}],
    }
    return f"data: {json.dumps(final_event)}\n\n".encode()


def _texts(seed, n):
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        out.append("".join(rng.choice(ALPHA) for _ in range(rng.randrange(0, 60))))
    out[3] = ""
    out[7] = "plain ascii only, no escapes at all"
    return out


def test_oracle_json_escape_matches_cpython():
    from oracle import c_oracle
    for t in _texts(1, 5000):
        assert c_oracle.json_escape(t.encode("utf-8", "surrogatepass")) == json.dumps(t)[1:-1].encode()


def test_envelope_split_reproduces_json_dumps():
    from quorum_b200.sse import event_envelope, backend_event_id, FINAL_EVENT_ID
    from oracle import c_oracle
    for t in _texts(2, 300):
        p, s = event_envelope(backend_event_id(4), 1712345, None)
        assert p + c_oracle.json_escape(t.encode("utf-8", "surrogatepass")) + s == _reference_event(4, 1712345, t)
        p, s = event_envelope(FINAL_EVENT_ID, 9, "stop")
        assert p + c_oracle.json_escape(t.encode("utf-8", "surrogatepass")) + s == _reference_final(9, t)


@pytest.mark.gpu
def test_gpu_framing_matches_reference_events():
    import quorum_b200 as qb
    texts = _texts(3, 4000) + ["x" * 5000 + "é" * 3000 + "\n" * 100]
    rng = random.Random(4)
    backend = [rng.randrange(0, 5) for _ in texts]
    got = qb.sse.frame_events(texts, backend, 1700000000)
    for t, i, g in zip(texts, backend, got):
        assert g == (_reference_event(i, 1700000000, t) if t else b""), (t, i)
    got = qb.sse.frame_events(texts[:50], backend[:50], 5, skip_empty=False)
    for t, i, g in zip(texts[:50], backend[:50], got):
        assert g == _reference_event(i, 5, t)
    assert qb.sse.frame_final_event(texts[10] + "\n-------------\n" + texts[11], 77) == \
        _reference_final(77, texts[10] + "\n-------------\n" + texts[11])


@pytest.mark.gpu
def test_gpu_framing_vs_oracle_batch_buffers():
    import torch
    import quorum_b200 as qb
    from oracle import c_oracle
    from quorum_b200 import synth
    dev = torch.device("cuda", 0)
    units = synth.cfg5_units(2 << 20, seed=9)
    chunks = []
    for u in units:
        p = 0
        while p < len(u):
            q = min(len(u), p + 64)
            while q < len(u) and (u[q] & 0xC0) == 0x80:
                q -= 1
            chunks.append(u[p:q])
            p = q
    chunks[5] = b""
    data, off = synth.pack(chunks)
    ip = (np.arange(len(chunks)) % 3).astype(np.int32)
    env = [qb.sse.event_envelope(qb.sse.backend_event_id(i), 123) for i in range(3)]
    exp_out, exp_off = c_oracle.sse_frame_batch(data, off, ip, [e[0] for e in env], env[0][1], True)
    t = lambda a: torch.from_numpy(np.array(a, copy=True)).to(dev)  # noqa: E731
    out, out_off = qb.sse.frame_events_device(t(data), t(off), t(ip), [e[0] for e in env], env[0][1], skip_empty=True)
    torch.cuda.synchronize()
    assert np.array_equal(out_off.cpu().numpy(), exp_off)
    assert np.array_equal(out[: int(exp_off[-1])].cpu().numpy(), exp_out)
