#!/usr/bin/env python3
"""Generate tests/golden/*.json by running the UNMODIFIED reference (staged by oracle/make_ref.py).

Run in the build container only (needs /root/reference or an already staged oracle/_ref):

    python tests/golden/make_golden.py            # rewrite the committed fixtures
    python tests/golden/make_golden.py --fuzz N   # additionally differential-fuzz the oracle, N cases

Fixtures (all strings are JSON str; lone surrogates never generated):
  strip_cases.json   [{tags, hide, content, expect}]                    <- strip_thinking_tags
  filter_cases.json  [{tags, feeds:[..], outs:[..], flush, depth, buffer}] <- ThinkingTagFilter
  concat_cases.json  [{tags, sep, hide_final, streaming, contents, expect}] <- the two join forms
  fold_tables.json   the CPython facts the oracle restates (checked exhaustively over all code points)
The interpreter version is recorded in every file (reference pins 3.13; this image has 3.12.3).
"""
import argparse
import json
import logging
import os
import random
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def load_reference():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import make_ref
    if not make_ref.stage():
        raise SystemExit("reference not available: need /root/reference or a staged oracle/_ref")
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "src"))
    logging.disable(logging.CRITICAL)
    import quorum.oai_proxy as q
    return q


TAGS5 = ["think", "reason", "reasoning", "thought", "Thought"]
TAGS4 = ["think", "reason", "reasoning", "thought"]
TAG_SETS = [TAGS5, TAGS4, ["think"], ["think", "reason"], ["a"], ["kiss", "x-1"], [], [""], ["sick_k"]]

FOLD = {"i": ["i", "I", "İ", "ı"], "k": ["k", "K", "K"], "s": ["s", "S", "ſ"]}
WS = [" ", "\n", "\t", "\r", "\x0b", "\x0c", "\x1c", "\x1f", "\x85", "\xa0", " ", " ", " ",
      " ", " ", " ", " ", "　"]
WORDS = ["Hello", "world", "The answer is 4.", "héllo", "日本語", "🎉", "a", "x", "İ", "ı", "ſ", "K", "​",
         "̇", "é", "<", ">", "/", "</", "<<", ">>", "<x>", "</x>", "<div>", "&lt;", "⁠", "﻿"]


def spell(name, rng, exotic):
    out = []
    for ch in name:
        lo = ch.lower()
        if lo in FOLD and rng.random() < exotic:
            out.append(rng.choice(FOLD[lo]))
        elif rng.random() < 0.3:
            out.append(ch.upper())
        else:
            out.append(ch)
    return "".join(out)


def rand_text(rng, tags, n_items, exotic=0.15):
    names = [t for t in tags] or [""]
    parts = []
    for _ in range(n_items):
        r = rng.random()
        if r < 0.22:
            parts.append("<" + spell(rng.choice(names), rng, exotic) + ">")
        elif r < 0.44:
            parts.append("</" + spell(rng.choice(names), rng, exotic) + ">")
        elif r < 0.52:
            full = rng.choice(["<", "</"]) + spell(rng.choice(names), rng, exotic) + ">"
            parts.append(full[: rng.randrange(1, len(full))])          # partial tag
        elif r < 0.56:
            parts.append("<" + spell(rng.choice(names), rng, exotic) + rng.choice([" >", "X>", "/>", ""]))
        elif r < 0.72:
            parts.append(rng.choice(WS))
        else:
            parts.append(rng.choice(WORDS))
    return "".join(parts)


def rand_chunks(rng, text):
    if not text:
        return [""] if rng.random() < 0.5 else []
    cuts = sorted(set(rng.randrange(0, len(text) + 1) for _ in range(rng.randrange(0, 8))))
    pieces, last = [], 0
    for c in cuts + [len(text)]:
        pieces.append(text[last:c])
        last = c
    if rng.random() < 0.2:
        pieces.insert(rng.randrange(len(pieces) + 1), "")
    return pieces


def run_filter(q, tags, feeds):
    f = q.ThinkingTagFilter(tags)
    outs = [f.feed(x) for x in feeds]
    depth_before_flush, buffer_before_flush = f.thinking_depth, f.buffer
    fl = f.flush()
    return outs, fl, depth_before_flush, buffer_before_flush


def ref_concat(q, contents, sep, tags, hide_final, streaming):
    if streaming:   # oai_proxy.py:759-764, 841
        filtered = [q.strip_thinking_tags(t, tags, hide_intermediate=hide_final) for t in contents if t]
        return f"\n{sep}".join(filtered)
    processed = [q.strip_thinking_tags(c, tags, hide_intermediate=hide_final) for c in contents]   # :1194-1199
    return sep.join(processed)                                                                      # :1286


def fold_tables():
    import _sre
    out = {"python": sys.version}
    for p in "iks":
        rx = re.compile(p, re.I)
        out["regex_" + p] = [c for c in range(0x110000) if rx.fullmatch(chr(c))]
    for p in "abcdefghjlmnopqrtuvwxyz0123456789_-":
        rx = re.compile(re.escape(p), re.I)
        extra = [c for c in range(128, 0x110000) if rx.fullmatch(chr(c))]
        assert not extra, (p, extra)
    out["simple_lower_to_ascii"] = [[c, _sre.unicode_tolower(c)] for c in range(128, 0x110000)
                                    if _sre.unicode_tolower(c) < 128]
    out["str_lower_ascii_start"] = [[c, [ord(x) for x in chr(c).lower()]] for c in range(128, 0x110000)
                                    if ord(chr(c).lower()[0]) < 128]
    out["isspace"] = [c for c in range(0x110000) if chr(c).isspace()]
    return out


def aggregate_cases(q, rng):
    """Prompts built by the reference's own aggregate_responses (oai_proxy.py:374-423): the aggregator call is replaced
    by a stub that captures the request body, everything before it is the unmodified code."""
    import asyncio
    captured = {}

    async def fake_call_backend(backend, body, headers, timeout=None):
        captured["prompt"] = json.loads(body)["messages"][0]["content"]
        return {"status_code": 200, "content": {"choices": [{"message": {"content": "synth"}}]}}

    q.call_backend = fake_call_backend
    seps = ["\n\n---\n\n", "\n", "", " § "]
    templates = ["You have received the following responses regarding the user's query:\n\n{responses}\n\nProvide a concise synthesis of these responses.",
                 "{responses}", "no placeholder at all", "A {responses} B {responses} C", "é{responses}日本"]
    qfmts = ["Original query: {query}\n\n", "{query}", "Q={query!r}; "]
    lfmts = ["Response from {backend_name}:\n", "[{backend_name}] ", ""]
    cases = []
    for _ in range(300):
        tags = rng.choice(TAG_SETS)
        responses = [rand_text(rng, tags, rng.randrange(0, 8)) if rng.random() > 0.1 else "" for _ in range(rng.randrange(0, 7))]
        case = {"responses": responses, "query": rand_text(rng, tags, rng.randrange(0, 4)), "sep": rng.choice(seps),
                "include_query": rng.random() < 0.7, "query_format": rng.choice(qfmts), "include_names": rng.random() < 0.5,
                "label_format": rng.choice(lfmts), "template": rng.choice(templates)}
        captured.clear()
        asyncio.run(q.aggregate_responses(responses, {"model": "m", "url": "http://x"}, case["query"], case["sep"], case["include_query"],
                                          case["query_format"], case["include_names"], case["label_format"], case["template"],
                                          {"Authorization": "Bearer t"}))
        case["expect"] = captured["prompt"]
        cases.append(case)
    return cases


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--aggregate-only", action="store_true", help="only (re)write aggregate_cases.json")
    ap.add_argument("--fuzz", type=int, default=0)
    ap.add_argument("--seed", type=int, default=20261106)
    ap.add_argument("--no-write", action="store_true")
    args = ap.parse_args()
    q = load_reference()
    rng = random.Random(args.seed)
    meta = {"python": sys.version, "generator": "tests/golden/make_golden.py", "seed": args.seed}
    if args.aggregate_only:
        cases = aggregate_cases(q, random.Random(args.seed + 1))
        with open(os.path.join(HERE, "aggregate_cases.json"), "w") as f:
            json.dump({"meta": meta, "cases": cases}, f, ensure_ascii=True, separators=(",", ":"))
        print("wrote", len(cases), "aggregate cases")
        return

    # hand-written seeds (observed behaviour) + random cases
    hand = [
        "A <think>x<think>y</think>z</think> B", "A <think>x<reason>y</think>z</reason> B", "A <think>unclosed B",
        "A </think> stray B", "A <think>x</reason> B </think> C", "A <reasoning>x</reason>y</reasoning> B",
        "a<<think>x</think>>b", "<think>x</think></think>y<think>z</think>w", "  \n<think>x</think>\n lead/trail ws \t\n",
        "A <thi", "A <think", "A <", "A <x", "A <thinkX", " 　 héllo <think>日本語</think> 🎉  \x85", "",
        "<thınk>x</think>y", "<thİnk>x</think>y", "<thinK>x</think>y", "<reaſon>x</reason>y",
        "<reaſon>x</reaſon>y", "<think>a</thınk>b</think>c", "<THINK>\n\n</think>", "<think></think>",
        " <think>x</think> ", "<think>x</think> \xa0 <reason>y</reason> ", "<Thought>t</thought>ok",
        "<reasoning>r</reason>", "<reason>r</reasoning>s</reason>t", " ", "\n\n", "<think>", "</think>", "<think></think><think>",
    ]
    strip_cases = []
    for tags in TAG_SETS:
        for c in hand:
            for hide in (True, False):
                strip_cases.append({"tags": tags, "hide": hide, "content": c,
                                    "expect": q.strip_thinking_tags(c, tags, hide_intermediate=hide)})
    for _ in range(1500):
        tags = rng.choice(TAG_SETS)
        c = rand_text(rng, tags, rng.randrange(0, 24))
        strip_cases.append({"tags": tags, "hide": True, "content": c, "expect": q.strip_thinking_tags(c, tags, True)})

    hand_feeds = [
        ["Hello <thi", "nk>secret</th", "ink> World"], ["a<reaſ", "on>x</reason>b"], ["a<reaſon>x</reason>b"],
        ["a<thinK", ">x</think>b"], ["a<thİ", "nk>x</think>b"], ["<think>a<thİ", "nk>b</think>c</think>d"],
        ["a<", "t", "h", "i", "n", "k", ">", "x", "<", "/", "t", "h", "i", "n", "k", ">", "b"], ["a<thi", "x"], ["a<thi", ""],
        ["a<thi", "<thi", "nk>"], ["<think>x</think></think>y<think>z</think>w"], ["A </think> stray B"], ["A <", ""],
        ["<think>", "</thi", "nk", ">", "ok"], ["x</thi", "nk>y"], ["<think><think>", "</think>a", "</think>b"],
        ["<THINK>a</ThInK", ">b<", "think>c"], [], [""], ["<"], ["<think"], ["<think>x</think", ""],
    ]
    filter_cases = []
    for tags in TAG_SETS:
        for feeds in hand_feeds:
            outs, fl, d, b = run_filter(q, tags, feeds)
            filter_cases.append({"tags": tags, "feeds": feeds, "outs": outs, "flush": fl, "depth": d, "buffer": b})
    for _ in range(1500):
        tags = rng.choice(TAG_SETS)
        feeds = rand_chunks(rng, rand_text(rng, tags, rng.randrange(0, 24)))
        outs, fl, d, b = run_filter(q, tags, feeds)
        filter_cases.append({"tags": tags, "feeds": feeds, "outs": outs, "flush": fl, "depth": d, "buffer": b})

    concat_cases = []
    seps = ["\n", "\n-------------\n", "\n\n—— § ——\n\n", "", "\n\nSEPARATOR\n\n"]
    for _ in range(400):
        tags = rng.choice(TAG_SETS)
        contents = [rand_text(rng, tags, rng.randrange(0, 10)) if rng.random() > 0.15 else ""
                    for _ in range(rng.randrange(1, 9))]
        for streaming in (False, True):
            sep, hide_final = rng.choice(seps), rng.random() < 0.7
            concat_cases.append({"tags": tags, "sep": sep, "hide_final": hide_final, "streaming": streaming,
                                 "contents": contents,
                                 "expect": ref_concat(q, contents, sep, tags, hide_final, streaming)})

    if not args.no_write:
        for name, data in (("strip_cases", strip_cases), ("filter_cases", filter_cases), ("concat_cases", concat_cases)):
            with open(os.path.join(HERE, name + ".json"), "w") as f:
                json.dump({"meta": meta, "cases": data}, f, ensure_ascii=True, separators=(",", ":"))
        with open(os.path.join(HERE, "fold_tables.json"), "w") as f:
            json.dump(fold_tables(), f, separators=(",", ":"))
        print("wrote", len(strip_cases), len(filter_cases), len(concat_cases), "cases")

    if args.fuzz:
        from oracle import quorum_oracle as o
        bad = 0
        for k in range(args.fuzz):
            tags = rng.choice(TAG_SETS)
            text = rand_text(rng, tags, rng.randrange(0, 40), exotic=rng.choice([0.0, 0.15, 0.5]))
            a, b = q.strip_thinking_tags(text, tags, True), o.strip_thinking_tags(text, tags, True)
            if a != b:
                bad += 1
                print("STRIP MISMATCH", tags, repr(text), repr(a), repr(b))
            feeds = rand_chunks(rng, text)
            ra = run_filter(q, tags, feeds)
            f = o.ThinkingTagFilter(tags)
            outs = [f.feed(x) for x in feeds]
            d, bf = f.thinking_depth, f.buffer
            rb = (outs, f.flush(), d, bf)
            if ra != rb:
                bad += 1
                print("FILTER MISMATCH", tags, feeds, ra, rb)
            if bad > 10:
                break
        print("fuzz", args.fuzz, "mismatches", bad)
        sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
