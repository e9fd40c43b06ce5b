// SSE input parsing: what the reference does with every HTTP chunk of a backend's stream
// before the filter sees a delta (src/quorum/oai_proxy.py:578-617):
//     chunk.decode()  ->  .strip().split("\n\n")  ->  event.startswith("data: ")  ->  event[6:].strip()  ->  "[DONE]"
//     ->  json.loads  ->  parsed["choices"][0].get("delta", {})["content"]
// Scalar work per chunk / per event, so a lane takes a chunk (split) or an event (parse): 32 of them per warp.
//
// The JSON side is deliberately conservative.  An event gets kind 1 (a delta: its unescaped content is written) or kind 0
// (no delta: not "data: ", blank, [DONE], a well-formed object without choices[0].delta.content, an empty choices list)
// only when that is certain under CPython's json module; everything else - malformed JSON, escapes in the watched keys,
// duplicate watched keys, a watched value of another type, NaN/Infinity, nesting deeper than kJsonDepth - gets kind 2 and
// the host decides with json.loads itself (quorum_b200/streaming.py), because there the reference's behaviour is Python's
// dynamic typing (TypeError / AttributeError end the backend's stream, JSONDecodeError skips the event).
#pragma once
#include "qb_common.cuh"

namespace qb {

constexpr int kJsonDepth = 16;
enum : uint8_t { kEvNone = 0, kEvDelta = 1, kEvHost = 2 };

// CPython's strict UTF-8 decoder: no overlong forms, no surrogates, nothing above U+10FFFF, no truncated sequence.
__device__ inline bool utf8_valid(const uint8_t* p, int64_t n) {
    int64_t i = 0;
    while (i < n) {
        const uint8_t c = p[i];
        if (c < 0x80) {
            i++;
            continue;
        }
        if (c < 0xC2 || c > 0xF4) return false;
        const int need = c < 0xE0 ? 1 : c < 0xF0 ? 2 : 3;
        if (i + need >= n) return false;  // truncated
        const uint8_t c1 = p[i + 1];
        if ((c1 & 0xC0) != 0x80) return false;
        if (c == 0xE0 && c1 < 0xA0) return false;
        if (c == 0xED && c1 >= 0xA0) return false;
        if (c == 0xF0 && c1 < 0x90) return false;
        if (c == 0xF4 && c1 >= 0x90) return false;
        for (int k = 2; k <= need; k++)
            if ((p[i + k] & 0xC0) != 0x80) return false;
        i += need + 1;
    }
    return true;
}

// str.strip() bounds of text [s, e) of buffer p: returns the new [s, e)
__device__ inline void strip_bounds(const uint8_t* p, int64_t& s, int64_t& e) {
    const PtrAcc B{p + s};
    int64_t a = 0, b = e - s;
    while (a < b) {
        const int l = ws_len_at(B, a, b);
        if (l == 0) break;
        a += l;
    }
    while (b > a && is_white_byte(B, b - 1, b)) b--;  // every byte of a white-space code point tests white
    e = s + b;
    s += a;
}

// The events of one HTTP chunk: callback(k, start, end) for every piece of chunk.decode().strip().split("\n\n"), in
// order.  Returns the number of pieces (>= 1), or 0 when the chunk does not decode (the reference drops it whole,
// :580-588).  One pass over the bytes does both the UTF-8 check (VALIDATE; CPython's strict decoder, see utf8_valid) and
// the split; the strip only looks at the two ends.
template <bool VALIDATE, class F>
__device__ inline int64_t for_each_event(const uint8_t* p, int64_t s0, int64_t e0, F&& cb) {
    int64_t s = s0, e = e0;
    strip_bounds(p, s, e);  // on bytes that do not decode this may be off, and the result is discarded below
    int64_t count = 0, start = s, i = VALIDATE ? s0 : s;
    const int64_t stop = VALIDATE ? e0 : e;
    while (i < stop) {
        const uint8_t c = p[i];
        if (c < 0x80) {
            if (c == '\n' && i >= start && i + 1 < e && p[i + 1] == '\n') {  // inside the stripped text: a separator
                cb(count, start, i);
                count++;
                start = i + 2;
                i += 2;
            } else {
                i++;
            }
            continue;
        }
        if (!VALIDATE) {
            i++;
            continue;
        }
        if (c < 0xC2 || c > 0xF4) return 0;
        const int need = c < 0xE0 ? 1 : c < 0xF0 ? 2 : 3;
        if (i + need >= e0) return 0;  // truncated
        const uint8_t c1 = p[i + 1];
        if ((c1 & 0xC0) != 0x80) return 0;
        if ((c == 0xE0 && c1 < 0xA0) || (c == 0xED && c1 >= 0xA0) || (c == 0xF0 && c1 < 0x90) || (c == 0xF4 && c1 >= 0x90)) return 0;
        for (int k = 2; k <= need; k++)
            if ((p[i + k] & 0xC0) != 0x80) return 0;
        i += need + 1;
    }
    cb(count, start, e);
    return count + 1;
}

__global__ void sse_split_count_kernel(const uint8_t* __restrict__ http, const int64_t* __restrict__ off, int64_t n_http,
                                       int64_t* __restrict__ counts) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_http) return;
    counts[c] = for_each_event<true>(http, off[c], off[c + 1], [](int64_t, int64_t, int64_t) {});
}

__global__ void sse_split_write_kernel(const uint8_t* __restrict__ http, const int64_t* __restrict__ off, int64_t n_http,
                                       const int64_t* __restrict__ ev_first, int64_t* __restrict__ ev_s, int64_t* __restrict__ ev_e) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_http) return;
    const int64_t base = ev_first[c];
    if (ev_first[c + 1] == base) return;  // did not decode: no events
    for_each_event<false>(http, off[c], off[c + 1], [&](int64_t k, int64_t a, int64_t b) {
        ev_s[base + k] = a;
        ev_e[base + k] = b;
    });
}

// ---- JSON ----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool json_ws(uint8_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
__device__ __forceinline__ int hexval(uint8_t c) {
    if (c >= '0' && c <= '9') return c - '0';
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    if (c >= 'A' && c <= 'F') return c - 'A' + 10;
    return -1;
}
// A string starting at the opening quote p[i]: returns the index after the closing quote, or -1 (not certain it is
// valid).  has_esc: it contains a backslash.
__device__ inline int64_t json_string(const uint8_t* p, int64_t i, int64_t e, bool& has_esc) {
    has_esc = false;
    i++;
    while (i < e) {
        const uint8_t c = p[i];
        if (c == '"') return i + 1;
        if (c < 0x20) return -1;
        if (c == '\\') {
            has_esc = true;
            if (i + 1 >= e) return -1;
            const uint8_t d = p[i + 1];
            if (d == 'u') {
                if (i + 5 >= e) return -1;
                for (int k = 2; k < 6; k++)
                    if (hexval(p[i + k]) < 0) return -1;
                i += 6;
            } else if (d == '"' || d == '\\' || d == '/' || d == 'b' || d == 'f' || d == 'n' || d == 'r' || d == 't') {
                i += 2;
            } else {
                return -1;
            }
        } else {
            i++;
        }
    }
    return -1;
}
// number / true / false / null starting at p[i]: index after it, or -1
__device__ inline int64_t json_scalar(const uint8_t* p, int64_t i, int64_t e) {
    auto lit = [&](const char* w, int n) -> int64_t {
        if (i + n > e) return -1;
        for (int k = 0; k < n; k++)
            if (p[i + k] != (uint8_t)w[k]) return -1;
        return i + n;
    };
    const uint8_t c = p[i];
    if (c == 't') return lit("true", 4);
    if (c == 'f') return lit("false", 5);
    if (c == 'n') return lit("null", 4);
    int64_t j = i;
    if (j < e && p[j] == '-') j++;
    if (j >= e) return -1;
    if (p[j] == '0') {
        j++;
    } else if (p[j] >= '1' && p[j] <= '9') {
        while (j < e && p[j] >= '0' && p[j] <= '9') j++;
    } else {
        return -1;
    }
    if (j < e && p[j] == '.') {
        j++;
        if (j >= e || p[j] < '0' || p[j] > '9') return -1;
        while (j < e && p[j] >= '0' && p[j] <= '9') j++;
    }
    if (j < e && (p[j] == 'e' || p[j] == 'E')) {
        j++;
        if (j < e && (p[j] == '+' || p[j] == '-')) j++;
        if (j >= e || p[j] < '0' || p[j] > '9') return -1;
        while (j < e && p[j] >= '0' && p[j] <= '9') j++;
    }
    if (j < e) {  // what follows must be able to follow a value
        const uint8_t d = p[j];
        if (!(json_ws(d) || d == ',' || d == ']' || d == '}')) return -1;
    }
    return j;
}

// event_data [s, e) -> kind; for kEvDelta, [cs, ce) is the raw content string between its quotes
__device__ inline uint8_t json_find_content(const uint8_t* p, int64_t s, int64_t e, int64_t& cs, int64_t& ce) {
    enum { R_OTHER = 0, R_TOP, R_CHOICES, R_CHOICE0, R_DELTA, R_CONTENT };
    uint8_t is_arr[kJsonDepth], role[kJsonDepth], seen[kJsonDepth];
    int depth = 0;
    int64_t i = s;
    bool found = false;
    auto ws = [&]() {
        while (i < e && json_ws(p[i])) i++;
    };
    auto key_is = [&](int64_t a, int64_t b, const char* w, int n) -> bool {
        if (b - a != n) return false;
        for (int k = 0; k < n; k++)
            if (p[a + k] != (uint8_t)w[k]) return false;
        return true;
    };
    int vrole = R_TOP;
    enum { M_VALUE, M_AFTER, M_KEY0, M_KEY, M_ELEM0 } mode = M_VALUE;
    for (;;) {
        ws();
        if (mode == M_VALUE) {
            if (i >= e) return kEvHost;
            const uint8_t c = p[i];
            if (c == '{' || c == '[') {
                // a watched value of the wrong container type: Python's typing decides, not us
                if ((vrole == R_TOP || vrole == R_CHOICE0 || vrole == R_DELTA) && c != '{') return kEvHost;
                if (vrole == R_CHOICES && c != '[') return kEvHost;
                if (vrole == R_CONTENT) return kEvHost;
                if (depth == kJsonDepth) return kEvHost;
                is_arr[depth] = c == '[';
                role[depth] = (uint8_t)vrole;
                seen[depth] = 0;
                depth++;
                i++;
                mode = c == '[' ? M_ELEM0 : M_KEY0;
            } else {
                if (vrole == R_TOP || vrole == R_CHOICES || vrole == R_CHOICE0 || vrole == R_DELTA) return kEvHost;
                if (c == '"') {
                    bool esc;
                    const int64_t j = json_string(p, i, e, esc);
                    if (j < 0) return kEvHost;
                    if (vrole == R_CONTENT) {
                        cs = i + 1;
                        ce = j - 1;
                        found = true;
                    }
                    i = j;
                } else {
                    if (vrole == R_CONTENT) return kEvHost;  // null, a number ...: feed(c) raises in the reference
                    const int64_t j = json_scalar(p, i, e);
                    if (j < 0) return kEvHost;
                    i = j;
                }
                mode = M_AFTER;
            }
        } else if (mode == M_KEY0 || mode == M_KEY) {
            if (i >= e) return kEvHost;
            if (p[i] == '}' && mode == M_KEY0) {
                depth--;
                i++;
                mode = M_AFTER;
                continue;
            }
            if (p[i] != '"') return kEvHost;
            bool esc;
            const int64_t j = json_string(p, i, e, esc);
            if (j < 0) return kEvHost;
            const int r = role[depth - 1];
            vrole = R_OTHER;
            if (r == R_TOP || r == R_CHOICE0 || r == R_DELTA) {
                if (esc) return kEvHost;  // could spell a watched key
                const char* w = r == R_TOP ? "choices" : r == R_CHOICE0 ? "delta" : "content";
                const int n = r == R_TOP ? 7 : r == R_CHOICE0 ? 5 : 7;
                if (key_is(i + 1, j - 1, w, n)) {
                    if (seen[depth - 1]) return kEvHost;  // duplicate key: the last one wins in Python
                    seen[depth - 1] = 1;
                    vrole = r == R_TOP ? R_CHOICES : r == R_CHOICE0 ? R_DELTA : R_CONTENT;
                }
            }
            i = j;
            ws();
            if (i >= e || p[i] != ':') return kEvHost;
            i++;
            mode = M_VALUE;
        } else if (mode == M_ELEM0) {
            if (i >= e) return kEvHost;
            if (p[i] == ']') {
                depth--;
                i++;
                mode = M_AFTER;
                continue;
            }
            vrole = role[depth - 1] == R_CHOICES ? R_CHOICE0 : R_OTHER;
            mode = M_VALUE;
        } else {  // M_AFTER
            if (depth == 0) break;
            if (i >= e) return kEvHost;
            const uint8_t c = p[i];
            if (is_arr[depth - 1]) {
                if (c == ',') {
                    i++;
                    vrole = R_OTHER;  // only element 0 of choices is watched
                    mode = M_VALUE;
                } else if (c == ']') {
                    depth--;
                    i++;
                } else {
                    return kEvHost;
                }
            } else {
                if (c == ',') {
                    i++;
                    mode = M_KEY;
                } else if (c == '}') {
                    depth--;
                    i++;
                } else {
                    return kEvHost;
                }
            }
        }
    }
    ws();
    if (i != e) return kEvHost;  // extra data
    return found ? kEvDelta : kEvNone;
}

// unescaped length / bytes of a raw JSON string body [s, e) (already validated by json_string)
template <bool WRITE>
__device__ inline int64_t json_unescape(const uint8_t* p, int64_t s, int64_t e, uint8_t* out) {
    int64_t n = 0;
    auto put = [&](uint8_t b) {
        if (WRITE) out[n] = b;
        n++;
    };
    auto put_cp = [&](uint32_t cp) {  // UTF-8, lone surrogates in their 3-byte form (str.encode("utf-8", "surrogatepass"))
        if (cp < 0x80) {
            put((uint8_t)cp);
        } else if (cp < 0x800) {
            put((uint8_t)(0xC0 | (cp >> 6)));
            put((uint8_t)(0x80 | (cp & 0x3F)));
        } else if (cp < 0x10000) {
            put((uint8_t)(0xE0 | (cp >> 12)));
            put((uint8_t)(0x80 | ((cp >> 6) & 0x3F)));
            put((uint8_t)(0x80 | (cp & 0x3F)));
        } else {
            put((uint8_t)(0xF0 | (cp >> 18)));
            put((uint8_t)(0x80 | ((cp >> 12) & 0x3F)));
            put((uint8_t)(0x80 | ((cp >> 6) & 0x3F)));
            put((uint8_t)(0x80 | (cp & 0x3F)));
        }
    };
    auto hex4 = [&](int64_t i) -> uint32_t {
        return (uint32_t)((hexval(p[i]) << 12) | (hexval(p[i + 1]) << 8) | (hexval(p[i + 2]) << 4) | hexval(p[i + 3]));
    };
    int64_t i = s;
    while (i < e) {
        const uint8_t c = p[i];
        if (c != '\\') {
            put(c);
            i++;
            continue;
        }
        const uint8_t d = p[i + 1];
        if (d == 'u') {
            uint32_t cp = hex4(i + 2);
            i += 6;
            // json.decoder: a high surrogate followed by \uDC00-\uDFFF is one code point
            if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 <= e && p[i] == '\\' && p[i + 1] == 'u') {
                const uint32_t lo = hex4(i + 2);
                if (lo >= 0xDC00 && lo <= 0xDFFF) {
                    cp = 0x10000 + (((cp - 0xD800) << 10) | (lo - 0xDC00));
                    i += 6;
                }
            }
            put_cp(cp);
        } else {
            put(d == 'b' ? '\b' : d == 'f' ? '\f' : d == 'n' ? '\n' : d == 'r' ? '\r' : d == 't' ? '\t' : d);
            i += 2;
        }
    }
    return n;
}

// lane per event: classify, and for a delta remember the raw content span and its unescaped length
__global__ void sse_parse_kernel(const uint8_t* __restrict__ http, const int64_t* __restrict__ ev_s, const int64_t* __restrict__ ev_e,
                                 int64_t n_events, uint8_t* __restrict__ kind, int64_t* __restrict__ cs_out,
                                 int64_t* __restrict__ ce_out, int64_t* __restrict__ len_out, unsigned long long* __restrict__ n_host) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_events) return;
    int64_t s = ev_s[k], e = ev_e[k];
    uint8_t kd = kEvNone;
    int64_t cs = 0, ce = 0, len = 0;
    int64_t bs = s, be = e;
    strip_bounds(http, bs, be);
    if (be > bs && e - s >= 6 && http[s] == 'd' && http[s + 1] == 'a' && http[s + 2] == 't' && http[s + 3] == 'a' && http[s + 4] == ':' &&
        http[s + 5] == ' ') {
        int64_t ds = s + 6, de = e;
        strip_bounds(http, ds, de);
        const bool done = de - ds == 6 && http[ds] == '[' && http[ds + 1] == 'D' && http[ds + 2] == 'O' && http[ds + 3] == 'N' &&
                          http[ds + 4] == 'E' && http[ds + 5] == ']';
        if (!done) {
            kd = json_find_content(http, ds, de, cs, ce);
            if (kd == kEvDelta) len = json_unescape<false>(http, cs, ce, nullptr);
        }
    }
    kind[k] = kd;
    cs_out[k] = cs;
    ce_out[k] = ce;
    len_out[k] = len;
    if (kd == kEvHost) atomicAdd(n_host, 1ull);
}

__global__ void sse_unescape_kernel(const uint8_t* __restrict__ http, int64_t n_events, const uint8_t* __restrict__ kind,
                                    const int64_t* __restrict__ cs, const int64_t* __restrict__ ce, const int64_t* __restrict__ out_off,
                                    uint8_t* __restrict__ out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_events || kind[k] != kEvDelta) return;
    json_unescape<true>(http, cs[k], ce[k], out + out_off[k]);
}

}  // namespace qb
