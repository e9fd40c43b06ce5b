"""CPU oracle for the quorum hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product path (quorum_b200/) never does and has no CPU fallback.

This is a byte-level (UTF-8) sequential restatement of the reference's two filters and two joins:

  * strip_thinking_tags  — /root/reference/src/quorum/oai_proxy.py:120-139
  * ThinkingTagFilter    — /root/reference/src/quorum/oai_proxy.py:262-371
  * non-stream join      — /root/reference/src/quorum/oai_proxy.py:1194-1199, 1286
  * stream-final join    — /root/reference/src/quorum/oai_proxy.py:759-764, 841

The arithmetic of the reference lives in CPython's `re` (_sre) and `str` (reference pins Python 3.13,
.python-version:1; this image runs 3.12.3).  The fold tables restated below were checked exhaustively
over all 0x110000 code points against CPython 3.12.3 in this container (tests/golden/make_golden.py
re-checks them whenever /root/reference is present):

  * re.IGNORECASE on an ASCII pattern letter p matches p, upper(p) and additionally
      'i': U+0130 (C4 B0), U+0131 (C4 B1);  'k': U+212A (E2 84 AA);  's': U+017F (C5 BF)
  * the backreference \\1 under IGNORECASE compares simple-lower-case per code point:
      U+0130 -> 'i', U+212A -> 'k';  U+0131 and U+017F only equal themselves
  * str.lower() (hold-back test, oai_proxy.py:303,317): ASCII upper->lower, U+212A -> 'k';
      U+0130 -> 'i'+U+0307 (two code points, never a tag prefix); U+0131/U+017F unchanged
  * str.strip() removes the 29 str.isspace() code points listed in WS_SEQS

Parity pinning: tests/test_oracle_golden.py checks this module against every exact-string vector
of the reference's tests/test_thinking_tag_filter.py and tests/test_parallel_backends.py:52-56, and
against tests/golden/*.json, which were produced by running the unmodified reference here
(tests/golden/make_golden.py).

Inputs are `bytes`.  `str` helpers encode with errors="surrogatepass" so that any Python str
round-trips; all byte sequences the automata look at are ASCII or the specific multi-byte forms
above, so every other byte passes through untouched.
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence, Tuple

LT, GT, SL = 0x3C, 0x3E, 0x2F

# str.isspace() code points in UTF-8 (oai_proxy.py:139 `.strip()`)
WS_SEQS: Tuple[bytes, ...] = tuple(
    [bytes([b]) for b in (0x09, 0x0A, 0x0B, 0x0C, 0x0D, 0x1C, 0x1D, 0x1E, 0x1F, 0x20)]
    + [b"\xc2\x85", b"\xc2\xa0", b"\xe1\x9a\x80"]
    + [bytes([0xE2, 0x80, x]) for x in range(0x80, 0x8B)]
    + [b"\xe2\x80\xa8", b"\xe2\x80\xa9", b"\xe2\x80\xaf", b"\xe2\x81\x9f", b"\xe3\x80\x80"]
)

_REGEX_SPECIAL = set(b".^$*+?{}[]\\|()")
MAX_TAG_LEN = 22          # same limits as the C-ABI tag table (include/quorum_b200.h)
MAX_TAGS = 16
MAX_PARTIAL_BYTES = 23    # longest viable partial "</name" in bytes, fold forms included


class TagTable:
    """Distinct lower-cased tag names + the flags both automata need.

    Mirrors how the reference builds its patterns: `"|".join(tags)` interpolated unescaped into
    `<(...)>` (oai_proxy.py:137-138, 272-274).  Because names are restricted to printable ASCII
    without regex metacharacters and without '<', '>', '/', alternation order and duplicates
    (e.g. "thought" / "Thought", config.yaml:37) cannot change what matches.
    An EMPTY list gives the pattern `<()>`, i.e. one empty name, but an empty `allowed_tags`
    list for the hold-back loop (oai_proxy.py:301,315) — `hold_enabled` records that.
    """

    def __init__(self, tags: Sequence[str]):
        names: List[bytes] = []
        for t in tags:
            if not isinstance(t, str):
                raise TypeError("tag names must be str")
            try:
                b = t.encode("ascii")
            except UnicodeEncodeError:
                raise ValueError(f"unsupported tag name {t!r}: ASCII only")
            for ch in b:
                if ch < 0x20 or ch > 0x7E or ch in _REGEX_SPECIAL or ch in (LT, GT, SL):
                    raise ValueError(f"unsupported tag name {t!r}: regex metacharacter or '<', '>', '/'")
            b = b.lower()
            if len(b) > MAX_TAG_LEN:
                raise ValueError(f"tag name {t!r} longer than {MAX_TAG_LEN}")
            if b not in names:
                names.append(b)
        self.hold_enabled = len(tags) > 0
        if not names:
            names = [b""]
        if len(names) > MAX_TAGS:
            raise ValueError(f"more than {MAX_TAGS} distinct tag names")
        for n in names:
            worst = 2 + sum(3 if c == 0x6B else 2 if c in (0x69, 0x73) else 1 for c in n)
            if worst > MAX_PARTIAL_BYTES:
                raise ValueError(f"tag name {n!r}: partial tag can exceed {MAX_PARTIAL_BYTES} bytes")
        self.names = names


def _match_name(buf: bytes, pos: int, end: int, name: bytes, hold_fold: bool) -> Optional[Tuple[int, int]]:
    """Match `name` at buf[pos:end).  Returns (pos_after_name, variant_mask) or None.

    hold_fold=False: re.IGNORECASE rules (tag recognition).  variant_mask has one bit per 'i'/'s'
    letter of the name, set when the text used U+0131 / U+017F (the forms that only equal
    themselves under the backreference).
    hold_fold=True: `candidate.lower()` rules (ASCII case + Kelvin sign only).
    """
    mask = 0
    bit = 0
    for p in name:
        if pos >= end:
            return None
        b = buf[pos]
        is_letter = 0x61 <= p <= 0x7A
        if is_letter and (b | 0x20) == p and (0x41 <= (b & 0xDF) <= 0x5A):
            pos += 1
        elif not is_letter and b == p:
            pos += 1
        elif p == 0x6B and buf[pos:pos + 3] == b"\xe2\x84\xaa" and pos + 3 <= end:
            pos += 3
        elif not hold_fold and p == 0x69 and pos + 2 <= end and buf[pos] == 0xC4 and buf[pos + 1] in (0xB0, 0xB1):
            if buf[pos + 1] == 0xB1:
                mask |= 1 << bit
            pos += 2
        elif not hold_fold and p == 0x73 and pos + 2 <= end and buf[pos] == 0xC5 and buf[pos + 1] == 0xBF:
            mask |= 1 << bit
            pos += 2
        else:
            return None
        if p in (0x69, 0x73):
            bit += 1
    return pos, mask


def match_tag(buf: bytes, pos: int, end: int, table: TagTable, close: bool) -> Optional[Tuple[int, int, int]]:
    """`<name>` (close=False) or `</name>` (close=True) starting at pos, fully inside [pos, end).

    Returns (end_after_'>', name_index, variant_mask) or None.  (name_index, variant_mask) is the
    backreference class: two captured names compare equal under `\\1`+IGNORECASE iff both agree.
    """
    if pos >= end or buf[pos] != LT:
        return None
    q = pos + 1
    if close:
        if q >= end or buf[q] != SL:
            return None
        q += 1
    for idx, name in enumerate(table.names):
        m = _match_name(buf, q, end, name, False)
        if m is not None and m[0] < end and buf[m[0]] == GT:
            return m[0] + 1, idx, m[1]
    return None


def _letter_forms(p: int, hold_fold: bool) -> Tuple[bytes, ...]:
    """Every UTF-8 form that matches pattern byte p (a lower-cased tag-name byte)."""
    if not (0x61 <= p <= 0x7A):
        return (bytes([p]),)
    forms = [bytes([p]), bytes([p - 0x20])]
    if p == 0x6B:
        forms.append(b"\xe2\x84\xaa")              # U+212A: matches under both fold rules
    if not hold_fold:
        if p == 0x69:
            forms += [b"\xc4\xb0", b"\xc4\xb1"]    # U+0130, U+0131: re.IGNORECASE only
        if p == 0x73:
            forms.append(b"\xc5\xbf")               # U+017F: re.IGNORECASE only
    return tuple(forms)


def is_tag_prefix(buf: bytes, pos: int, end: int, table: TagTable, hold_fold: bool, allow_close: bool) -> bool:
    """True iff buf[pos:end) (which starts with '<' and is not a complete tag) is a BYTE prefix of
    some complete tag: "<name>" (and "</name>" when allow_close) with every letter in any of its
    matching forms.  A multi-byte form cut by `end` still counts (the str API can never produce
    that cut; the byte-level ABI defines it this way so that a flat scan of a stream and this
    incremental restatement agree on every input)."""
    if pos >= end or buf[pos] != LT:
        return False
    for close in ((False, True) if allow_close else (False,)):
        q = pos + 1
        if close:
            if q == end:
                return True
            if buf[q] != SL:
                continue
            q += 1
        for name in table.names:
            r = q
            ok = True
            for p in name:
                if r == end:
                    break
                matched = False
                for f in _letter_forms(p, hold_fold):
                    avail = buf[r:min(end, r + len(f))]
                    if avail == f:
                        r += len(f)
                        matched = True
                        break
                    if len(avail) < len(f) and f.startswith(avail):
                        return True                 # form cut by `end`
                if not matched:
                    ok = False
                    break
            if ok and r == end:
                return True
    return False


def holdable(buf: bytes, pos: int, end: int, table: TagTable) -> bool:
    """Depth-0 hold-back test: `full_tag.startswith(candidate.lower())` for any allowed tag
    (oai_proxy.py:300-304, 313-319).  Never true when the configured tag list is empty."""
    return table.hold_enabled and is_tag_prefix(buf, pos, end, table, True, False)


def viable_partial(buf: bytes, pos: int, end: int, table: TagTable) -> bool:
    """Inside a thinking block the reference keeps its whole buffer (oai_proxy.py:332-334); only a
    trailing partial open/close tag that more text could still complete can matter later."""
    return is_tag_prefix(buf, pos, end, table, False, True)


# ----------------------------------------------------------------------------------------------
# A.1  strip_thinking_tags  (oai_proxy.py:120-139)
# ----------------------------------------------------------------------------------------------

def _ws_len_at(buf: bytes, pos: int, end: int) -> int:
    for s in WS_SEQS:
        if pos + len(s) <= end and buf[pos:pos + len(s)] == s:
            return len(s)
    return 0


def strip_ws(buf: bytes) -> bytes:
    """str.strip() on UTF-8 bytes: drop leading/trailing str.isspace() code points."""
    n = len(buf)
    # byte p is "white" iff it lies inside a WS sequence found at p, p-1 or p-2
    white = bytearray(n)
    for p in range(n):
        l = _ws_len_at(buf, p, n)
        for k in range(l):
            white[p + k] = 1
    a = 0
    while a < n and white[a]:
        a += 1
    b = n
    while b > a and white[b - 1]:
        b -= 1
    return bytes(buf[a:b])


def strip_thinking_tags_bytes(content: bytes, table: TagTable, hide_intermediate: bool = True) -> bytes:
    """re.sub("<(t1|t2|..)>.*?</\\1>", "", content, IGNORECASE|DOTALL).strip()  (oai_proxy.py:137-139).

    Leftmost open tag that has a LATER closer of the same backreference class deletes through the
    FIRST such closer; regions do not nest; an open tag without such a closer is literal text and
    scanning resumes one byte later; stray closers are literal.  With the flag off the input is
    returned unchanged — not even stripped (oai_proxy.py:134-135).
    """
    if not hide_intermediate:
        return bytes(content)
    n = len(content)
    out = bytearray()
    i = 0
    dead = set()  # classes with no closer at or after the current position (pure optimisation)
    while i < n:
        if content[i] == LT:
            m = match_tag(content, i, n, table, False)
            if m is not None and (m[1], m[2]) not in dead:
                j = m[0]
                found = -1
                while j < n:
                    if content[j] == LT:
                        c = match_tag(content, j, n, table, True)
                        if c is not None and (c[1], c[2]) == (m[1], m[2]):
                            found = c[0]
                            break
                    j += 1
                if found >= 0:
                    i = found
                    continue
                dead.add((m[1], m[2]))
        out.append(content[i])
        i += 1
    return strip_ws(bytes(out))


# ----------------------------------------------------------------------------------------------
# A.2  ThinkingTagFilter  (oai_proxy.py:262-371)
# ----------------------------------------------------------------------------------------------

class ThinkingTagFilterBytes:
    """Incremental depth-counting filter with partial-tag hold-back.

    State carried between feeds: `depth` (thinking_depth, oai_proxy.py:276) and `held`, the tail the
    reference keeps in `self.buffer`: at depth 0 a holdable partial open tag (oai_proxy.py:311-323);
    at depth>0 the reference keeps everything since the last consumed tag (oai_proxy.py:332-334), of
    which only a still-viable partial tag at the very end can influence later output — that is what
    `held` keeps.  `ref_buffer` additionally mirrors the reference's `self.buffer` exactly.
    """

    def __init__(self, table: TagTable):
        self.table = table
        self.depth = 0
        self.held = b""
        self.ref_buffer = b""

    def feed(self, text: bytes) -> bytes:
        t = self.table
        buf = self.held + bytes(text)
        full_ref = self.ref_buffer + bytes(text)
        ref_base = len(full_ref) - len(buf)     # buf is a suffix of the reference's buffer
        n = len(buf)
        out = bytearray()
        out_pos: List[int] = []                 # source index of every emitted byte
        i = 0
        last_tag_end = 0
        while i < n:
            if buf[i] == LT:
                m = match_tag(buf, i, n, t, False)          # open tag: consumed at ANY depth
                if m is not None:                            # (oai_proxy.py:288-296, 343-346)
                    self.depth += 1
                    i = last_tag_end = m[0]
                    continue
                if self.depth > 0:                           # close tag: only inside a block
                    m = match_tag(buf, i, n, t, True)        # (oai_proxy.py:335-342)
                    if m is not None:
                        self.depth -= 1
                        i = last_tag_end = m[0]
                        continue
            if self.depth == 0:
                out.append(buf[i])
                out_pos.append(i)
            i += 1
        p = buf.rfind(b"<", last_tag_end)
        if self.depth == 0:
            if p >= 0 and holdable(buf, p, n, t):                  # oai_proxy.py:311-323
                self.held = buf[p:]
                while out_pos and out_pos[-1] >= p:
                    out_pos.pop()
                    out.pop()
            else:
                self.held = b""
            self.ref_buffer = self.held
        else:
            self.held = buf[p:] if (p >= 0 and viable_partial(buf, p, n, t)) else b""
            self.ref_buffer = full_ref[ref_base + last_tag_end:] if last_tag_end > 0 else full_ref
        return bytes(out)

    def flush(self) -> bytes:
        """oai_proxy.py:352-371 — always returns b"": unclosed content and a held partial are dropped.
        `depth` is left as it is (the reference does not reset thinking_depth)."""
        self.held = b""
        self.ref_buffer = b""
        return b""


# ----------------------------------------------------------------------------------------------
# A.3  joins
# ----------------------------------------------------------------------------------------------

def concatenate_bytes(contents: Iterable[bytes], separator: bytes, table: TagTable,
                      hide_final_think: bool, streaming: bool) -> bytes:
    """Non-stream: sep.join(strip(c) for c in contents) — empties kept (oai_proxy.py:1194-1199, 1286).
    Stream final: ("\\n"+sep).join(strip(c) for c in contents if c) — contents empty BEFORE the
    filter are dropped, ones that become empty after it are kept (oai_proxy.py:759-764, 841)."""
    if streaming:
        items = [strip_thinking_tags_bytes(c, table, hide_final_think) for c in contents if len(c) > 0]
        return (b"\n" + separator).join(items)
    return separator.join(strip_thinking_tags_bytes(c, table, hide_final_think) for c in contents)


# ----------------------------------------------------------------------------------------------
# str-level helpers with the reference's signatures
# ----------------------------------------------------------------------------------------------

def _enc(s: str) -> bytes:
    return s.encode("utf-8", "surrogatepass")


def _dec(b: bytes) -> str:
    return b.decode("utf-8", "surrogatepass")


def strip_thinking_tags(content: str, tags: List[str], hide_intermediate: bool = True) -> str:
    if not hide_intermediate:
        return content
    if not isinstance(content, str):
        raise TypeError("expected string or bytes-like object, got %r" % type(content).__name__)
    return _dec(strip_thinking_tags_bytes(_enc(content), TagTable(tags), True))


class ThinkingTagFilter:
    def __init__(self, tags):
        self.allowed_tags = [tag.lower() for tag in tags]
        self._f = ThinkingTagFilterBytes(TagTable(tags))

    @property
    def thinking_depth(self) -> int:
        return self._f.depth

    @property
    def buffer(self) -> str:
        return _dec(self._f.ref_buffer)

    def feed(self, text: str) -> str:
        return _dec(self._f.feed(_enc(text)))

    def flush(self) -> str:
        return _dec(self._f.flush())


def concatenate(contents: List[str], separator: str, tags: List[str], hide_final_think: bool,
                streaming: bool) -> str:
    return _dec(concatenate_bytes([_enc(c) for c in contents], _enc(separator), TagTable(tags),
                                  hide_final_think, streaming))


def aggregate_prompt(source_responses: List[str], user_query: str, separator: str, include_original_query: bool = True,
                     query_format: str = "Original query: {query}\n\n", include_source_names: bool = False,
                     source_label_format: str = "Response from {backend_name}:\n",
                     prompt_template: str = "You have received the following responses regarding the user's query:\n\n{responses}\n\nProvide a concise synthesis of these responses.") -> str:
    """The aggregator prompt of src/quorum/oai_proxy.py:406-423 (row f3), restated: labels, join, template substitution."""
    formatted = []
    for i, response in enumerate(source_responses):  # :407-415
        if include_source_names:
            formatted.append(source_label_format.format(backend_name=f"LLM{i + 1}") + response)
        else:
            formatted.append(response)
    intermediate = separator.join(formatted)  # :417
    prompt = ""
    if include_original_query:  # :419-421
        prompt += query_format.format(query=user_query)
    prompt += prompt_template.replace("{responses}", intermediate)  # :423
    return prompt


def usage_sums(usages: List[List[int]]) -> List[int]:
    """oai_proxy.py:1298-1311 (row f4): the three sums over the successful responses of one request."""
    return [sum(u[0] for u in usages), sum(u[1] for u in usages), sum(u[2] for u in usages)]
