"""Aggregate-strategy prompt assembly and usage sums.

tests/golden/aggregate_cases.json holds prompts produced by the reference's own aggregate_responses
(src/quorum/oai_proxy.py:374-423, aggregator call stubbed; generator: tests/golden/make_golden.py --aggregate-only)."""
import json
import os
import random

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _cases():
    with open(os.path.join(HERE, "golden", "aggregate_cases.json")) as f:
        return json.load(f)["cases"]


def test_oracle_matches_reference_prompts():
    from oracle import quorum_oracle as o
    cases = _cases()
    assert len(cases) >= 300
    for c in cases:
        got = o.aggregate_prompt(c["responses"], c["query"], c["sep"], c["include_query"], c["query_format"], c["include_names"],
                                 c["label_format"], c["template"])
        assert got == c["expect"], c


def test_oracle_usage_sums():
    from oracle import quorum_oracle as o
    assert o.usage_sums([[1, 2, 3], [10, 20, 30]]) == [11, 22, 33]
    assert o.usage_sums([]) == [0, 0, 0]


@pytest.mark.gpu
def test_gpu_prompts_match_reference_prompts():
    """The reference passes already filtered answers to aggregate_responses; hide_final_think=False (the reference's
    default) keeps the builder's filter out of the way so the golden prompts apply as they are (grouped by configuration: one batch each)."""
    import quorum_b200 as qb
    groups = {}
    for c in _cases():
        key = (c["sep"], c["include_query"], c["query_format"], c["include_names"], c["label_format"], c["template"])
        groups.setdefault(key, []).append(c)
    for key, cs in groups.items():
        b = qb.AggregatePromptBuilder(separator=key[0], include_original_query=key[1], query_format=key[2], include_source_names=key[3],
                                      source_label_format=key[4], prompt_template=key[5], hide_final_think=False)
        got = b.build_prompts([c["responses"] for c in cs], [c["query"] for c in cs])
        for c, g in zip(cs, got):
            assert g == c["expect"], c


@pytest.mark.gpu
def test_gpu_prompts_with_filter_match_oracle():
    import quorum_b200 as qb
    from oracle import quorum_oracle as o
    rng = random.Random(5)
    tags = ["think", "reason", "reasoning", "thought"]
    pieces = ["<think>", "</think>", "<reason>", "</reason>", "text ", "日本語", " \n", "<thi", "x<y", "é"]
    reqs, queries = [], []
    for _ in range(200):
        reqs.append(["".join(rng.choice(pieces) for _ in range(rng.randrange(0, 30))) for _ in range(rng.randrange(1, 6))])
        queries.append("".join(rng.choice(pieces) for _ in range(rng.randrange(0, 5))))
    b = qb.AggregatePromptBuilder.from_config({"strategy": {"aggregate": {"include_source_names": True, "thinking_tags": tags,
                                                                           "hide_final_think": True,
                                                                           "prompt_template": "T: {intermediate_results} :end"}},
                                               "iterations": {"aggregation": {"strategy": "aggregate"}}})
    got = b.build_prompts(reqs, queries)
    for cs, q, g in zip(reqs, queries, got):
        filtered = [o.strip_thinking_tags(c, tags, True) for c in cs]
        assert g == o.aggregate_prompt(filtered, q, "\n\n---\n\n", True, "Original query: {query}\n\n", True,
                                       "Response from {backend_name}:\n", "T: {responses} :end")


@pytest.mark.gpu
def test_gpu_usage_sums():
    import torch
    from quorum_b200 import aggregate
    from oracle import quorum_oracle as o
    rng = random.Random(6)
    usage, ro = [], [0]
    for _ in range(1000):
        k = rng.randrange(0, 9)
        usage += [[rng.randrange(0, 10**6), rng.randrange(0, 10**6), rng.randrange(0, 10**7)] for _ in range(k)]
        ro.append(ro[-1] + k)
    dev = torch.device("cuda", 0)
    u = torch.tensor(usage, dtype=torch.int64).reshape(-1, 3).to(dev)
    got = aggregate.usage_sums_device(u, torch.tensor(ro, dtype=torch.int64).to(dev)).cpu().tolist()
    for q in range(1000):
        assert got[q] == o.usage_sums(usage[ro[q]:ro[q + 1]])
