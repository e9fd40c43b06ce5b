"""Seeded synthetic workloads of the shapes BASELINE.json names.

numpy only; the same bytes go to the GPU path and to the CPU oracle / reference.  Large batches are built from a
pool of distinct base units gathered with repetition, which changes nothing for a byte scanner (every unit is read
from its own address) and keeps generation time in seconds.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np

TAGS5 = ["think", "reason", "reasoning", "thought", "Thought"]          # reference config.yaml:37
SEP_MULTIBYTE = "\n\n—— § ——\n\n"

_WORDS = [w.encode() for w in (
    "the of and to in is that it was for on are as with his they at be this from have or by one had not but what all "
    "were when we there can an your which their said if do will each about how up out them then she many some so these "
    "would other into has more her two like him see time could no make than first been its who now people my made over "
    "did down only way find use may water long little very after words called just where most know").split()]
_MB = ["é", "ü", "ñ", "ß", "Ω", "ж", "中", "日本", "語", "한", "€", "—", "“", "”", "🎉", "🚀", "𝛑"]
_MBB = [m.encode() for m in _MB]


def _spell(rng: np.random.Generator, name: str) -> bytes:
    return "".join(c.upper() if rng.random() < 0.25 else c for c in name).encode()


_POOLS: dict = {}


def _pool(mb_frac: float) -> bytes:
    """A fixed 1 MiB pool of words, spaces, newlines and ~mb_frac multi-byte characters (valid UTF-8)."""
    key = round(mb_frac, 3)
    if key not in _POOLS:
        rng = np.random.default_rng(int(key * 1000) + 77)
        out = bytearray()
        while len(out) < (1 << 20):
            if rng.random() < mb_frac:
                out += _MBB[int(rng.integers(len(_MBB)))]
            else:
                out += _WORDS[int(rng.integers(len(_WORDS)))]
            out += b"\n" if rng.random() < 0.05 else b" "
        _POOLS[key] = bytes(out)
    return _POOLS[key]


def _fill_text(rng: np.random.Generator, n: int, mb_frac: float = 0.10) -> bytearray:
    """n bytes of text cut from the pool at a random offset, trimmed to code-point boundaries, space padded."""
    pool = _pool(mb_frac)
    out = bytearray()
    while len(out) < n:
        a = int(rng.integers(0, len(pool) - 1))
        while (pool[a] & 0xC0) == 0x80:
            a += 1
        b = min(len(pool), a + (n - len(out)))
        while b < len(pool) and b > a and (pool[b] & 0xC0) == 0x80:
            b -= 1
        if b == len(pool):
            b = a + max(0, min(n - len(out), len(pool) - a - 8))
            while b > a and (pool[b] & 0xC0) == 0x80:
                b -= 1
        out += pool[a:b] if b > a else b" "
        if len(out) < n:
            out += b" "
    del out[n:]
    k = n
    # the cut at n may have split a character that started before n
    i = n - 1
    while i >= 0 and (out[i] & 0xC0) == 0x80:
        i -= 1
    if i >= 0 and out[i] >= 0xC0:
        need = 2 if out[i] < 0xE0 else 3 if out[i] < 0xF0 else 4
        if i + need > n:
            for j in range(i, n):
                out[j] = 0x20
    return out


def _fix_boundaries(buf: bytearray, step: int) -> None:
    """Make sure no multi-byte character straddles a multiple of `step` (chunk cuts land on code points):
    a straddling character is replaced by ASCII 'x' bytes."""
    for cut in range(step, len(buf), step):
        if (buf[cut] & 0xC0) == 0x80:
            a = cut
            while (buf[a] & 0xC0) == 0x80:
                a -= 1
            b = cut
            while b < len(buf) and (buf[b] & 0xC0) == 0x80:
                b += 1
            for i in range(a, b):
                buf[i] = 0x78


def _place(buf: bytearray, pos: int, piece: bytes) -> None:
    buf[pos:pos + len(piece)] = piece


def _repair_after_overwrite(buf: bytearray, a: int, b: int) -> None:
    """Overwriting [a,b) may have cut a multi-byte character at either edge: blank the orphaned bytes."""
    i = b
    while i < len(buf) and (buf[i] & 0xC0) == 0x80:
        buf[i] = 0x20
        i += 1
    i = a - 1
    while i >= 0 and (buf[i] & 0xC0) == 0x80:
        i -= 1
    if i >= 0 and buf[i] >= 0xC0:
        need = 2 if buf[i] < 0xE0 else 3 if buf[i] < 0xF0 else 4
        if i + need > a:
            for j in range(i, a):
                buf[j] = 0x20


def make_unit(rng: np.random.Generator, n: int, blocks: int, inside_frac: float, tags=TAGS5, chunk: int = 0,
              straddle: bool = False, unclosed: bool = False, end_partial: bool = False, lead_block: bool = False,
              ws_edges: bool = True) -> bytes:
    """One response / stream of exactly n bytes with `blocks` thinking blocks covering ~inside_frac of the bytes."""
    buf = _fill_text(rng, n)
    if blocks > 0 and n >= 64:
        inside = int(n * inside_frac)
        # split `inside` bytes over the blocks, place them at increasing random positions
        sizes = np.maximum(24, (rng.dirichlet(np.ones(blocks)) * inside).astype(int)) if not lead_block else None
        if lead_block:
            sizes = np.array([inside] + [int(rng.integers(24, 64)) for _ in range(blocks - 1)])
        free = n - int(sizes.sum()) - 8
        if free > blocks:
            gaps = (rng.dirichlet(np.ones(blocks + 1)) * free).astype(int)
            if lead_block:
                gaps[0] = int(rng.integers(0, 4))
            pos = 0
            for bi in range(blocks):
                pos += int(gaps[bi])
                size = int(sizes[bi])
                name = tags[int(rng.integers(len(tags)))]
                op = b"<" + _spell(rng, name) + b">"
                cl = b"</" + _spell(rng, name) + b">"
                a, b = pos, pos + size
                if chunk and straddle and bi == 0:
                    # force the open tag to straddle a chunk boundary at a random interior offset
                    cut = ((a // chunk) + 1) * chunk
                    if cut + len(op) < b - len(cl) - 1:
                        a = cut - int(rng.integers(1, len(op)))
                    cut2 = (((b - len(cl)) // chunk)) * chunk
                    if cut2 - len(cl) > a + len(op) + 1 and cut2 + len(cl) < n:
                        b = cut2 + int(rng.integers(1, len(cl)))
                _place(buf, a, op)
                last = bi == blocks - 1
                if not (unclosed and last):
                    _place(buf, b - len(cl), cl)
                _repair_after_overwrite(buf, a, a + len(op))
                if not (unclosed and last):
                    _repair_after_overwrite(buf, b - len(cl), b)
                pos = b
    if end_partial:
        part = b"<" + _spell(rng, tags[0])
        part = part[: int(rng.integers(1, len(part) + 1))]
        _place(buf, n - len(part), part)
        _repair_after_overwrite(buf, n - len(part), n)
    if not ws_edges and n > 0:
        pass
    if chunk:
        _fix_boundaries(buf, chunk)
    assert len(buf) == n
    return bytes(buf)


def gather_units(base: np.ndarray, n_units: int, seed: int) -> np.ndarray:
    """[n_units, unit_bytes] gathered with repetition from the pool `base` ([n_base, unit_bytes])."""
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, base.shape[0], size=n_units)
    idx[: min(n_units, base.shape[0])] = np.arange(min(n_units, base.shape[0]))   # every base unit appears
    return idx


def cfg2_base(n_base: int = 1024, chunks: int = 256, chunk_bytes: int = 64, seed: int = 1002) -> np.ndarray:
    """Streaming config: streams of `chunks` x `chunk_bytes`, 4 thinking blocks ~50 % of bytes, 5 tags in random case,
    first block's tags straddle chunk boundaries, 1 % unclosed, 1 % end on a held partial."""
    rng = np.random.default_rng(seed)
    n = chunks * chunk_bytes
    rows = []
    for i in range(n_base):
        rows.append(np.frombuffer(make_unit(rng, n, 4, 0.5, chunk=chunk_bytes, straddle=True,
                                            unclosed=(i % 100 == 7), end_partial=(i % 100 == 13)), dtype=np.uint8))
    return np.stack(rows)


def cfg3_base(n_base: int = 4096, resp_bytes: int = 4096, seed: int = 1003) -> np.ndarray:
    """1M x 4 KB config: 0-3 blocks per response, ~30 % of bytes inside."""
    rng = np.random.default_rng(seed)
    rows = []
    for _ in range(n_base):
        b = int(rng.integers(0, 4))
        rows.append(np.frombuffer(make_unit(rng, resp_bytes, b, 0.3 * 4 / 3 if b else 0.0), dtype=np.uint8))
    return np.stack(rows)


def cfg4_base(n_base: int = 512, resp_bytes: int = 32768, seed: int = 1004) -> np.ndarray:
    """8 x 32 KB reasoning-heavy config: one leading <think> block = 70 % of bytes + up to 2 short later blocks."""
    rng = np.random.default_rng(seed)
    rows = []
    for _ in range(n_base):
        rows.append(np.frombuffer(make_unit(rng, resp_bytes, 1 + int(rng.integers(0, 3)), 0.7, tags=["think"],
                                            lead_block=True), dtype=np.uint8))
    return np.stack(rows)


_FOLD = {"i": ["i", "I", "İ", "ı"], "k": ["k", "K", "K"], "s": ["s", "S", "ſ"]}
_WS_MB = ["\x85", "\xa0", "\u1680", "\u2000", "\u2003", "\u200a", "\u2028", "\u2029", "\u202f", "\u205f", "\u3000"]


def cfg5_units(total_bytes: int, seed: int = 1005, min_len: int = 64, max_len: int = 65536) -> List[bytes]:
    """Adversarial mix with log-uniform lengths: nested / crossed / unclosed / stray tags, fold characters, multi-byte
    white space at the edges, plain text."""
    rng = np.random.default_rng(seed)
    units: List[bytes] = []
    acc = 0

    def exotic(name: str) -> str:
        return "".join(_FOLD[c][int(rng.integers(len(_FOLD[c])))] if c in _FOLD and rng.random() < 0.5 else
                       (c.upper() if rng.random() < 0.3 else c) for c in name)

    while acc < total_bytes:
        n = int(np.exp(rng.uniform(np.log(min_len), np.log(max_len))))
        kind = rng.random()
        parts: List[str] = []
        size = 0
        while size < n:
            r = rng.random()
            t = TAGS5[int(rng.integers(len(TAGS5)))]
            if kind < 0.25:      # nested same/different names
                piece = (f"<{exotic(t)}>a<{exotic(TAGS5[int(rng.integers(5))])}>b</{exotic(t)}>c</{exotic(t)}> "
                         if r < 0.2 else _fill_text(rng, int(rng.integers(8, 200))).decode())
            elif kind < 0.40:    # unclosed
                piece = f"<{exotic(t)}>never closed " if r < 0.1 else _fill_text(rng, int(rng.integers(8, 200))).decode()
            elif kind < 0.50:    # stray closers
                piece = f"</{exotic(t)}> " if r < 0.15 else _fill_text(rng, int(rng.integers(8, 200))).decode()
            elif kind < 0.60:    # crossed names
                u = TAGS5[int(rng.integers(len(TAGS5)))]
                piece = (f"<{exotic(t)}>x<{exotic(u)}>y</{exotic(t)}>z</{exotic(u)}> " if r < 0.2
                         else _fill_text(rng, int(rng.integers(8, 200))).decode())
            elif kind < 0.80:    # heavy multi-byte, fold characters, partial tags
                piece = (_fill_text(rng, int(rng.integers(8, 120)), mb_frac=0.5).decode() if r < 0.7 else
                         f"<{exotic(t)}>" if r < 0.8 else f"</{exotic(t)}>" if r < 0.9 else f"<{exotic(t)[:2]}")
            else:
                piece = _fill_text(rng, int(rng.integers(8, 400))).decode()
            parts.append(piece)
            size += len(piece.encode())
        s = "".join(parts)
        if kind >= 0.60 and kind < 0.80:
            s = _WS_MB[int(rng.integers(len(_WS_MB)))] + " " + s + " " + _WS_MB[int(rng.integers(len(_WS_MB)))]
        b = s.encode()
        units.append(b)
        acc += len(b)
    return units


def pack(units: List[bytes]) -> Tuple[np.ndarray, np.ndarray]:
    off = np.zeros(len(units) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(u) for u in units])
    data = np.frombuffer(b"".join(units), dtype=np.uint8) if off[-1] else np.zeros(0, dtype=np.uint8)
    return data, off
