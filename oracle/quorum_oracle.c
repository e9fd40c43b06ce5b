/* CPU oracle for the quorum hot path, plain C — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (quorum_b200/) never links or calls it and has no CPU fallback.
 *
 * Sequential byte-level (UTF-8) restatement of
 *   strip_thinking_tags      /root/reference/src/quorum/oai_proxy.py:120-139
 *   ThinkingTagFilter        /root/reference/src/quorum/oai_proxy.py:262-371
 *   non-stream join          /root/reference/src/quorum/oai_proxy.py:1194-1199, 1286
 *   stream-final join        /root/reference/src/quorum/oai_proxy.py:759-764, 841
 * The arithmetic lives in CPython's re/_sre and str (reference pins Python 3.13; fold tables
 * verified exhaustively on 3.12.3, see oracle/quorum_oracle.py header and tests/golden/fold_tables.json).
 * Parity pinning: tests/test_oracle_golden.py runs this library against the reference's own
 * exact-string test vectors and against tests/golden/ JSON fixtures (outputs of the unmodified reference).
 *
 * The batch entry points take the same buffers as the CUDA C-ABI (include/quorum_b200.h) but in host
 * memory, so the tests can compare output buffers byte for byte.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define QO_MAX_TAGS 16
#define QO_MAX_TAG_LEN 22
#define QO_HELD_MAX 23

#define QO_FLAG_HIDE 1u        /* run the filter + strip (hide_final_think / hide_intermediate True) */
#define QO_FLAG_STREAM_JOIN 2u /* a7: drop contents empty BEFORE filtering (oai_proxy.py:763) */
#define QO_FLAG_FLUSH 4u       /* streaming: apply flush() after the last chunk (oai_proxy.py:739) */

typedef struct {
    int32_t n;
    int32_t hold_enabled;
    uint8_t len[QO_MAX_TAGS];
    uint8_t name[QO_MAX_TAGS][QO_MAX_TAG_LEN + 2];
} qo_tags;

typedef struct {
    int32_t depth;
    int32_t tail_len; /* bytes after the last tag consumed by this call, -1 if none was consumed */
    uint8_t held_len;
    uint8_t held[QO_HELD_MAX];
} qo_stream_state;

static int is_regex_special(uint8_t c) { return c != 0 && strchr(".^$*+?{}[]\\|()", c) != NULL; }

/* names: concatenated tag bytes, name_off[n_tags+1].  Returns 0 or a negative error code. */
int qo_tags_init(qo_tags *t, const uint8_t *names, const int32_t *name_off, int32_t n_tags) {
    memset(t, 0, sizeof(*t));
    t->hold_enabled = n_tags > 0;
    for (int i = 0; i < n_tags; i++) {
        int len = name_off[i + 1] - name_off[i];
        if (len < 0 || len > QO_MAX_TAG_LEN) return -2;
        uint8_t low[QO_MAX_TAG_LEN + 2];
        int worst = 2;
        for (int k = 0; k < len; k++) {
            uint8_t c = names[name_off[i] + k];
            if (c < 0x20 || c > 0x7E || is_regex_special(c) || c == '<' || c == '>' || c == '/') return -3;
            if (c >= 'A' && c <= 'Z') c = (uint8_t)(c + 32);
            low[k] = c;
            worst += c == 'k' ? 3 : (c == 'i' || c == 's') ? 2 : 1;
        }
        if (worst > QO_HELD_MAX) return -4;
        int dup = 0;
        for (int j = 0; j < t->n; j++)
            if (t->len[j] == len && memcmp(t->name[j], low, (size_t)len) == 0) dup = 1;
        if (dup) continue;
        if (t->n == QO_MAX_TAGS) return -5;
        memcpy(t->name[t->n], low, (size_t)len);
        t->len[t->n++] = (uint8_t)len;
    }
    if (t->n == 0) { t->n = 1; t->len[0] = 0; } /* "<()>" : one empty name */
    return 0;
}

/* One pattern letter p at buf[pos..end).  Returns bytes consumed (0 = no match); *variant set to 1 when the
 * text used U+0131 / U+017F.  *cut set to 1 when a multi-byte form is cut by `end` (prefix still viable). */
static int match_letter(const uint8_t *buf, int64_t pos, int64_t end, uint8_t p, int hold_fold, int *variant, int *cut) {
    *variant = 0;
    *cut = 0;
    if (pos >= end) return 0;
    uint8_t b = buf[pos];
    if (p >= 'a' && p <= 'z') {
        if (b == p || b == (uint8_t)(p - 32)) return 1;
        if (p == 'k' && b == 0xE2) {
            if (pos + 1 >= end) { *cut = 1; return 0; }
            if (buf[pos + 1] != 0x84) return 0;
            if (pos + 2 >= end) { *cut = 1; return 0; }
            return buf[pos + 2] == 0xAA ? 3 : 0;
        }
        if (!hold_fold && p == 'i' && b == 0xC4) {
            if (pos + 1 >= end) { *cut = 1; return 0; }
            if (buf[pos + 1] == 0xB0) return 2;
            if (buf[pos + 1] == 0xB1) { *variant = 1; return 2; }
            return 0;
        }
        if (!hold_fold && p == 's' && b == 0xC5) {
            if (pos + 1 >= end) { *cut = 1; return 0; }
            if (buf[pos + 1] == 0xBF) { *variant = 1; return 2; }
            return 0;
        }
        return 0;
    }
    return b == p ? 1 : 0;
}

/* "<name>" / "</name>" fully inside [pos,end).  Returns end offset after '>' or -1; class = idx<<16 | variant mask. */
static int64_t match_tag(const qo_tags *t, const uint8_t *buf, int64_t pos, int64_t end, int close, uint32_t *cls) {
    if (pos >= end || buf[pos] != '<') return -1;
    int64_t q = pos + 1;
    if (close) {
        if (q >= end || buf[q] != '/') return -1;
        q++;
    }
    for (int i = 0; i < t->n; i++) {
        int64_t r = q;
        uint32_t mask = 0;
        int bit = 0, ok = 1;
        for (int k = 0; k < t->len[i]; k++) {
            uint8_t p = t->name[i][k];
            int var, cut;
            int c = match_letter(buf, r, end, p, 0, &var, &cut);
            if (!c) { ok = 0; break; }
            if (p == 'i' || p == 's') { if (var) mask |= 1u << bit; bit++; }
            r += c;
        }
        if (ok && r < end && buf[r] == '>') {
            *cls = ((uint32_t)i << 24) | mask;
            return r + 1;
        }
    }
    return -1;
}

/* buf[pos..end) starts with '<', is not a complete tag: is it a byte prefix of some complete tag? */
static int is_tag_prefix(const qo_tags *t, const uint8_t *buf, int64_t pos, int64_t end, int hold_fold, int allow_close) {
    if (pos >= end || buf[pos] != '<') return 0;
    for (int close = 0; close <= (allow_close ? 1 : 0); close++) {
        int64_t q = pos + 1;
        if (close) {
            if (q == end) return 1;
            if (buf[q] != '/') continue;
            q++;
        }
        for (int i = 0; i < t->n; i++) {
            int64_t r = q;
            int ok = 1;
            for (int k = 0; k < t->len[i]; k++) {
                if (r == end) break;
                int var, cut;
                int c = match_letter(buf, r, end, t->name[i][k], hold_fold, &var, &cut);
                if (cut) return 1;
                if (!c) { ok = 0; break; }
                r += c;
            }
            if (ok && r == end) return 1;
        }
    }
    return 0;
}

static int ws_len_at(const uint8_t *b, int64_t p, int64_t end) {
    uint8_t c = b[p];
    if ((c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x20)) return 1;
    if (c == 0xC2 && p + 1 < end && (b[p + 1] == 0x85 || b[p + 1] == 0xA0)) return 2;
    if (p + 2 < end) {
        uint8_t d = b[p + 1], e = b[p + 2];
        if (c == 0xE1 && d == 0x9A && e == 0x80) return 3;
        if (c == 0xE2 && d == 0x80 && ((e >= 0x80 && e <= 0x8A) || e == 0xA8 || e == 0xA9 || e == 0xAF)) return 3;
        if (c == 0xE2 && d == 0x81 && e == 0x9F) return 3;
        if (c == 0xE3 && d == 0x80 && e == 0x80) return 3;
    }
    return 0;
}

/* byte p of b[0..n) lies inside a white-space sequence that starts at p, p-1 or p-2 */
static int is_white_byte(const uint8_t *b, int64_t p, int64_t n) {
    for (int back = 0; back < 3 && p - back >= 0; back++)
        if (ws_len_at(b, p - back, n) > back) return 1;
    return 0;
}

/* A.1  strip_thinking_tags (oai_proxy.py:120-139).  out must hold n bytes.  Returns output length. */
int64_t qo_strip(const qo_tags *t, const uint8_t *in, int64_t n, int hide, uint8_t *out) {
    if (!hide) { /* oai_proxy.py:134-135: unchanged, not stripped */
        memcpy(out, in, (size_t)n);
        return n;
    }
    uint32_t dead[64];
    int ndead = 0;
    int64_t i = 0, o = 0;
    while (i < n) {
        const uint8_t *lt = memchr(in + i, '<', (size_t)(n - i));
        int64_t p = lt ? (int64_t)(lt - in) : n;
        memcpy(out + o, in + i, (size_t)(p - i)); /* literal run up to the next '<' */
        o += p - i;
        i = p;
        if (i >= n) break;
        uint32_t cls;
        int64_t oe = match_tag(t, in, i, n, 0, &cls);
        int live = 0;
        if (oe >= 0) {
            int isdead = 0;
            for (int k = 0; k < ndead; k++)
                if (dead[k] == cls) isdead = 1;
            if (!isdead) {
                int64_t j = oe;
                while (j < n) { /* first closer of the same backreference class at or after oe */
                    const uint8_t *c = memchr(in + j, '<', (size_t)(n - j));
                    if (!c) break;
                    j = (int64_t)(c - in);
                    uint32_t ccls;
                    int64_t ce = match_tag(t, in, j, n, 1, &ccls);
                    if (ce >= 0 && ccls == cls) {
                        i = ce;
                        live = 1;
                        break;
                    }
                    j++;
                }
                if (!live && ndead < 64) dead[ndead++] = cls;
            }
        }
        if (!live) out[o++] = in[i++]; /* literal '<', resume one byte later */
    }
    /* .strip(): white-ness is decided on the filtered text */
    int64_t a = 0, b = o;
    while (a < b && is_white_byte(out, a, o)) a++;
    while (b > a && is_white_byte(out, b - 1, o)) b--;
    memmove(out, out + a, (size_t)(b - a));
    return b - a;
}

/* A.2  ThinkingTagFilter.feed (oai_proxy.py:278-350).  out must hold held_len + n bytes.  Returns output length. */
int64_t qo_filter_feed(const qo_tags *t, qo_stream_state *st, const uint8_t *text, int64_t n, uint8_t *out) {
    int64_t hl = st->held_len, m = hl + n;
    uint8_t stackbuf[4096];
    uint8_t *buf = m <= (int64_t)sizeof(stackbuf) ? stackbuf : (uint8_t *)malloc((size_t)m);
    memcpy(buf, st->held, (size_t)hl);
    memcpy(buf + hl, text, (size_t)n);
    int64_t i = 0, o = 0, last_tag_end = 0, last_lt = -1, o_at_last_lt = 0;
    int32_t depth = st->depth;
    while (i < m) {
        if (buf[i] == '<') {
            uint32_t cls;
            int64_t e = match_tag(t, buf, i, m, 0, &cls); /* open: consumed at any depth (:288-296, :343-346) */
            if (e >= 0) { depth++; i = last_tag_end = e; last_lt = -1; continue; }
            if (depth > 0) {
                e = match_tag(t, buf, i, m, 1, &cls); /* close: only inside a block (:335-342) */
                if (e >= 0) { depth--; i = last_tag_end = e; last_lt = -1; continue; }
            }
            last_lt = i;
            o_at_last_lt = o;
        }
        if (depth == 0) out[o++] = buf[i];
        i++;
    }
    st->tail_len = last_tag_end > 0 ? (int32_t)(m - last_tag_end) : -1;
    st->held_len = 0;
    if (last_lt >= 0) {
        if (depth == 0) {
            if (t->hold_enabled && is_tag_prefix(t, buf, last_lt, m, 1, 0)) { /* :311-323 */
                st->held_len = (uint8_t)(m - last_lt);
                o = o_at_last_lt;
            }
        } else if (is_tag_prefix(t, buf, last_lt, m, 0, 1)) { /* :332-334, only this tail can matter */
            st->held_len = (uint8_t)(m - last_lt);
        }
        if (st->held_len) memcpy(st->held, buf + last_lt, st->held_len);
    }
    st->depth = depth;
    if (buf != stackbuf) free(buf);
    return o;
}

/* A.2 flush (oai_proxy.py:352-371): always returns "", drops the held partial, keeps depth. */
void qo_filter_flush(qo_stream_state *st) { st->held_len = 0; }

/* a1 + a6/a7 over a batch.  Same argument meaning as qb200_strip_concat_batch, host memory. */
int qo_strip_concat_batch(const qo_tags *t, const uint8_t *in_bytes, const int64_t *in_off, int64_t n_resp,
                          const int64_t *req_off, int64_t n_req, const uint8_t *sep, int32_t sep_len, uint32_t flags,
                          uint8_t *out_bytes, int64_t out_capacity, int64_t *out_off, int64_t *resp_out_len) {
    int64_t o = 0;
    (void)n_resp;
    int hide = (flags & QO_FLAG_HIDE) != 0;
    for (int64_t q = 0; q < n_req; q++) {
        int64_t r0 = req_off ? req_off[q] : q, r1 = req_off ? req_off[q + 1] : q + 1;
        out_off[q] = o;
        int first = 1;
        for (int64_t r = r0; r < r1; r++) {
            int64_t len = in_off[r + 1] - in_off[r];
            if ((flags & QO_FLAG_STREAM_JOIN) && len == 0) { /* oai_proxy.py:763 `if text` */
                if (resp_out_len) resp_out_len[r] = 0;
                continue;
            }
            if (!first) {
                if (o + sep_len > out_capacity) return -10;
                memcpy(out_bytes + o, sep, (size_t)sep_len);
                o += sep_len;
            }
            first = 0;
            if (o + len > out_capacity) return -10;
            int64_t k = qo_strip(t, in_bytes + in_off[r], len, hide, out_bytes + o);
            if (resp_out_len) resp_out_len[r] = k;
            o += k;
        }
    }
    out_off[n_req] = o;
    return 0;
}

/* a3/a4 over a batch of streams.  Same argument meaning as qb200_stream_feed_batch, host memory. */
int qo_stream_feed_batch(const qo_tags *t, const uint8_t *chunk_bytes, const int64_t *chunk_off, int64_t n_chunks,
                         const int64_t *stream_off, int64_t n_streams, const qo_stream_state *state_in,
                         qo_stream_state *state_out, uint32_t flags, uint8_t *out_bytes, int64_t out_capacity,
                         int64_t *out_chunk_off) {
    int64_t o = 0;
    for (int64_t s = 0; s < n_streams; s++) {
        qo_stream_state st;
        if (state_in) st = state_in[s];
        else memset(&st, 0, sizeof(st)), st.tail_len = -1;
        int32_t tail = -1;
        for (int64_t c = stream_off[s]; c < stream_off[s + 1]; c++) {
            int64_t len = chunk_off[c + 1] - chunk_off[c];
            out_chunk_off[c] = o;
            if (o + len + st.held_len > out_capacity) return -10;
            o += qo_filter_feed(t, &st, chunk_bytes + chunk_off[c], len, out_bytes + o);
            /* tail_len over the whole call: restart at a consumed tag, else keep growing */
            tail = st.tail_len >= 0 ? st.tail_len : (tail >= 0 ? tail + (int32_t)len : -1);
        }
        st.tail_len = tail;
        if (flags & QO_FLAG_FLUSH) qo_filter_flush(&st);
        if (state_out) state_out[s] = st;
    }
    out_chunk_off[n_chunks] = o;
    return 0;
}

/* CPython json.dumps(str) body (ensure_ascii=True) on UTF-8 bytes, without the surrounding quotes.
 * Restates json.encoder.py_encode_basestring_ascii, the string encoder behind the reference's
 * f"data: {json.dumps(stream_event)}\n\n" (oai_proxy.py:646, 860).  out must hold 6*n bytes.  Returns length. */
int64_t qo_json_escape(const uint8_t *in, int64_t n, uint8_t *out) {
    static const char hex[] = "0123456789abcdef";
    int64_t o = 0;
    for (int64_t i = 0; i < n;) {
        uint32_t c = in[i];
        if (c < 0x80) {
            i++;
            const char *sh = c == '"' ? "\\\"" : c == '\\' ? "\\\\" : c == '\n' ? "\\n" : c == '\r' ? "\\r" : c == '\t' ? "\\t"
                           : c == '\b' ? "\\b" : c == '\f' ? "\\f" : NULL;
            if (sh) { out[o++] = (uint8_t)sh[0]; out[o++] = (uint8_t)sh[1]; continue; }
            if (c >= 0x20 && c < 0x7F) { out[o++] = (uint8_t)c; continue; }
        } else if (c < 0xE0) {
            c = ((c & 0x1F) << 6) | (i + 1 < n ? (in[i + 1] & 0x3F) : 0);
            i += 2;
        } else if (c < 0xF0) {
            c = ((c & 0x0F) << 12) | ((i + 1 < n ? (in[i + 1] & 0x3F) : 0) << 6) | (i + 2 < n ? (in[i + 2] & 0x3F) : 0);
            i += 3;
        } else {
            c = ((c & 0x07) << 18) | ((i + 1 < n ? (in[i + 1] & 0x3F) : 0) << 12) | ((i + 2 < n ? (in[i + 2] & 0x3F) : 0) << 6) |
                (i + 3 < n ? (in[i + 3] & 0x3F) : 0);
            i += 4;
        }
        uint32_t units[2];
        int nu = 1;
        if (c >= 0x10000) {
            c -= 0x10000;
            units[0] = 0xD800 + ((c >> 10) & 0x3FF);
            units[1] = 0xDC00 + (c & 0x3FF);
            nu = 2;
        } else {
            units[0] = c;
        }
        for (int k = 0; k < nu; k++) {
            out[o++] = '\\';
            out[o++] = 'u';
            out[o++] = (uint8_t)hex[(units[k] >> 12) & 15];
            out[o++] = (uint8_t)hex[(units[k] >> 8) & 15];
            out[o++] = (uint8_t)hex[(units[k] >> 4) & 15];
            out[o++] = (uint8_t)hex[units[k] & 15];
        }
    }
    return o;
}

/* SSE framing of a batch: item k -> prefix[item_prefix[k]] + escape(content_k) + suffix (host memory twin of
 * qb200_sse_frame_batch).  skip_empty: items with empty content produce nothing (oai_proxy.py:625-628). */
int qo_sse_frame_batch(const uint8_t *in_bytes, const int64_t *in_off, int64_t n_items, const int32_t *item_prefix,
                       const uint8_t *prefix_bytes, const int64_t *prefix_off, const uint8_t *suffix, int32_t suffix_len,
                       int skip_empty, uint8_t *out_bytes, int64_t out_capacity, int64_t *out_off) {
    int64_t o = 0;
    for (int64_t k = 0; k < n_items; k++) {
        out_off[k] = o;
        int64_t len = in_off[k + 1] - in_off[k];
        if (skip_empty && len == 0) continue;
        int p = item_prefix ? item_prefix[k] : 0;
        int64_t pl = prefix_off[p + 1] - prefix_off[p];
        if (o + pl + 6 * len + suffix_len > out_capacity) return -10;
        memcpy(out_bytes + o, prefix_bytes + prefix_off[p], (size_t)pl);
        o += pl;
        o += qo_json_escape(in_bytes + in_off[k], len, out_bytes + o);
        memcpy(out_bytes + o, suffix, (size_t)suffix_len);
        o += suffix_len;
    }
    out_off[n_items] = o;
    return 0;
}

int qo_sizeof_tags(void) { return (int)sizeof(qo_tags); }
int qo_sizeof_state(void) { return (int)sizeof(qo_stream_state); }
