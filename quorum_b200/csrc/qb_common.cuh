// Device-side building blocks shared by every kernel of the path: the tag table, tag / partial-tag matchers,
// white-space classification, warp-level helpers.  Semantics follow SURVEY.md Appendix A; each function names the
// reference lines (src/quorum/oai_proxy.py) whose behaviour it reproduces.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace qb {

// Programmatic dependent launch: a kernel launched with the stream-serialization attribute may be scheduled while the
// kernel before it on the stream is still draining; everything it reads from that kernel comes after this wait (a no-op
// for a kernel launched the ordinary way).  (An explicit griddepcontrol.launch_dependents right behind the wait, so that
// the next kernel's blocks take their slots while this one's last blocks run: cfg2 0.478 -> 0.485 ms, cfg3 1.889 -> 1.886.)
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }


constexpr int kMaxTags = 16;
constexpr int kNameStride = 24;
constexpr int kHeldMax = 23;
constexpr int kMaxCls = 128;
constexpr unsigned kFull = 0xffffffffu;

// Passed to kernels by value (__grid_constant__): no device allocation for the tag table.
struct DevTags {
    int32_t n;             // distinct lower-cased names (>= 1; an empty configured list gives one empty name)
    int32_t hold_enabled;  // 0 when the configured list was empty (oai_proxy.py:301, 315 loop over nothing)
    int32_t ncls;          // total backreference classes = sum over names of 2^(#i + #s letters)
    int32_t pad;
    uint8_t len[kMaxTags];
    uint8_t cls_base[kMaxTags];
    uint8_t name[kMaxTags][kNameStride];
    // fast matcher (ASCII spellings): the 24 bytes after "<" or "</" are compared as six 32-bit words against
    // name + '>' :  ((text | om) & lm) == pk.   om = 0x20 at letter positions, lm = 0xFF over the pattern length.
    uint32_t pk[kMaxTags][6], om[kMaxTags][6], lm[kMaxTags][6];
    uint32_t first_ok[8];  // bytes that may follow '<' in a tag ('/', first letters in both cases, fold lead bytes, '>' for "")
};

struct StreamState {  // == qb200_stream_state
    int32_t depth;
    int32_t tail_len;
    uint8_t held_len;
    uint8_t held[kHeldMax];
};
static_assert(sizeof(StreamState) == 32, "state layout");

// Byte accessors.  Positions are relative to the start of a unit (a response or a stream's data of this call).
struct PtrAcc {
    const uint8_t* p;
    __device__ __forceinline__ uint8_t operator()(int64_t x) const { return p[x]; }
};
// Logical text = held bytes (positions -h .. -1) followed by the unit's bytes (positions 0 ..).
struct HeldAcc {
    const uint8_t* p;
    const uint8_t* held;
    int h;
    __device__ __forceinline__ uint8_t operator()(int64_t x) const { return x < 0 ? held[h + x] : p[x]; }
};

// One pattern letter `pl` (lower-cased tag byte) against the text at x, text ends at `end`.
// Returns bytes consumed (0 = no match).  var=1 when the text used U+0131 / U+017F (forms that only equal
// themselves under the backreference); exotic=1 when it used a form str.lower() does not fold (U+0130, U+0131,
// U+017F); cut=1 when a multi-byte form is cut by `end` (the prefix is still viable).
// hold_fold=false: re.IGNORECASE (tag recognition, oai_proxy.py:139, 273-274);
// hold_fold=true : candidate.lower() (hold-back test, oai_proxy.py:303, 317).
template <class A>
__device__ __forceinline__ int match_letter(const A& B, int64_t x, int64_t end, uint8_t pl, bool hold_fold, int& var,
                                            int& exotic, int& cut) {
    var = 0;
    exotic = 0;
    cut = 0;
    if (x >= end) return 0;
    uint8_t b = B(x);
    if (pl >= 'a' && pl <= 'z') {
        if (b == pl || b == (uint8_t)(pl - 32)) return 1;
        if (pl == 'k' && b == 0xE2) {
            if (x + 1 >= end) { cut = 1; return 0; }
            if (B(x + 1) != 0x84) return 0;
            if (x + 2 >= end) { cut = 1; return 0; }
            return B(x + 2) == 0xAA ? 3 : 0;
        }
        if (!hold_fold && pl == 'i' && b == 0xC4) {
            if (x + 1 >= end) { cut = 1; return 0; }
            uint8_t c = B(x + 1);
            if (c == 0xB0) { exotic = 1; return 2; }
            if (c == 0xB1) { exotic = 1; var = 1; return 2; }
            return 0;
        }
        if (!hold_fold && pl == 's' && b == 0xC5) {
            if (x + 1 >= end) { cut = 1; return 0; }
            if (B(x + 1) == 0xBF) { exotic = 1; var = 1; return 2; }
            return 0;
        }
        return 0;
    }
    return b == pl ? 1 : 0;
}

// "<name>" (close=false) or "</name>" (close=true) starting at x and complete inside [x, end).
// Returns the position after '>' or -1.  cls = dense backreference class (name, U+0131/U+017F variant mask);
// exotic_pos = position of the first letter written in a form the hold-back fold rejects, or -1.
template <class A>
__device__ int64_t match_tag(const DevTags& tg, const A& B, int64_t x, int64_t end, bool close, int& cls,
                             int64_t& exotic_pos) {
    exotic_pos = -1;
    if (x >= end || B(x) != '<') return -1;
    int64_t q = x + 1;
    if (close) {
        if (q >= end || B(q) != '/') return -1;
        q++;
    }
    for (int i = 0; i < tg.n; i++) {
        int64_t r = q, ex = -1;
        int mask = 0, bit = 0;
        bool ok = true;
        const int len = tg.len[i];
        for (int k = 0; k < len; k++) {
            uint8_t pl = tg.name[i][k];
            int var, exo, cut;
            int c = match_letter(B, r, end, pl, false, var, exo, cut);
            if (!c) { ok = false; break; }
            if (pl == 'i' || pl == 's') { mask |= var << bit; bit++; }
            if (exo && ex < 0) ex = r;
            r += c;
        }
        if (ok && r < end && B(r) == '>') {
            cls = tg.cls_base[i] + mask;
            exotic_pos = ex;
            return r + 1;
        }
    }
    return -1;
}

// Fast tag recognition on a shared-memory copy (at least 36 readable bytes after x).  Same result as match_tag for
// open and close tags together: returns the tag length in bytes (0 = not a tag); is_close, cls, exrel (offset of the
// first U+0130/U+0131/U+017F form, 0xFF if none) as in match_tag.  ASCII spellings take the word-compare path; anything
// with a non-ASCII byte in the window falls back to the exact byte-wise matcher.
// kDefer: a window that needs the exact matcher returns -1 instead of running it (the tokenizer collects such candidates
// and matches them 32 at a time).
template <bool kDefer = false>
__device__ __forceinline__ int fast_match_tag(const DevTags& tg, const uint8_t* D, int x, int end, int& is_close, int& cls,
                                              int& exrel) {
    is_close = 0;
    cls = 0;
    exrel = 0xFF;
    if (x + 1 >= end) return 0;
    const uint8_t c1 = D[x + 1];
    if (!((tg.first_ok[c1 >> 5] >> (c1 & 31)) & 1u)) return 0;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(D + ((x + 1) & ~3));
    const int sh = ((x + 1) & 3) * 8;
    uint32_t r0 = w[0], r1 = w[1], r2 = w[2], r3 = w[3], r4 = w[4], r5 = w[5], r6 = w[6], r7 = w[7];
    uint32_t b0 = __funnelshift_r(r0, r1, sh), b1 = __funnelshift_r(r1, r2, sh), b2 = __funnelshift_r(r2, r3, sh),
             b3 = __funnelshift_r(r3, r4, sh), b4 = __funnelshift_r(r4, r5, sh), b5 = __funnelshift_r(r5, r6, sh),
             b6 = __funnelshift_r(r6, r7, sh);
    if ((b0 & 0xffu) == '/') {
        is_close = 1;
        b0 = __funnelshift_r(b0, b1, 8);
        b1 = __funnelshift_r(b1, b2, 8);
        b2 = __funnelshift_r(b2, b3, 8);
        b3 = __funnelshift_r(b3, b4, 8);
        b4 = __funnelshift_r(b4, b5, 8);
        b5 = __funnelshift_r(b5, b6, 8);
    }
    const int avail = end - x - 1 - is_close;  // bytes available for name + '>'
    for (int i = 0; i < tg.n; i++) {
        if (((b0 | tg.om[i][0]) & tg.lm[i][0]) != tg.pk[i][0]) continue;
        const int L = tg.len[i] + 1;
        if (L > avail) continue;
        if (L > 4) {
            uint32_t d = (((b1 | tg.om[i][1]) & tg.lm[i][1]) ^ tg.pk[i][1]) | (((b2 | tg.om[i][2]) & tg.lm[i][2]) ^ tg.pk[i][2]);
            if (L > 12)
                d |= (((b3 | tg.om[i][3]) & tg.lm[i][3]) ^ tg.pk[i][3]) | (((b4 | tg.om[i][4]) & tg.lm[i][4]) ^ tg.pk[i][4]) |
                     (((b5 | tg.om[i][5]) & tg.lm[i][5]) ^ tg.pk[i][5]);
            if (d) continue;
        }
        cls = tg.cls_base[i];
        return 1 + is_close + L;
    }
    // A spelling the word compare cannot see uses U+0130/U+0131 (C4 B0/B1), U+017F (C5 BF) or U+212A (E2 84 AA): only a
    // window that holds one of those lead bytes goes to the exact matcher (other multi-byte text cannot be part of a tag)
    auto has_byte = [](uint32_t w, uint32_t pat) { const uint32_t v = w ^ pat; return (v - 0x01010101u) & ~v & 0x80808080u; };
    // (w, nx: a word and the one after it; E2 counts only when 84 follows, which keeps typographic punctuation E2 80 xx out)
    auto fold_lead = [&](uint32_t w, uint32_t nx) {
        return has_byte(w & 0xFEFEFEFEu, 0xC4C4C4C4u) | (has_byte(w, 0xE2E2E2E2u) & has_byte(__funnelshift_r(w, nx, 8), 0x84848484u));
    };
    if (((b0 | b1 | b2 | b3 | b4 | b5) & 0x80808080u) &&
        (fold_lead(b0, b1) | fold_lead(b1, b2) | fold_lead(b2, b3) | fold_lead(b3, b4) | fold_lead(b4, b5) | fold_lead(b5, 0x84848484u))) {
        if (kDefer) return -1;
        const PtrAcc A{D};
        int64_t ex;
        int64_t tend = match_tag(tg, A, x, end, is_close != 0, cls, ex);
        if (tend >= 0) {
            if (ex >= 0) exrel = (int)(ex - x);
            return (int)(tend - x);
        }
    }
    return 0;
}

// Text [x, end) starts with '<' and is not a complete tag: is it a byte prefix of some complete tag?
// hold_fold=true, allow_close=false : depth-0 hold-back test (oai_proxy.py:311-323)
// hold_fold=false, allow_close=true : tail worth carrying inside a thinking block (oai_proxy.py:332-334)
template <class A>
__device__ bool is_tag_prefix(const DevTags& tg, const A& B, int64_t x, int64_t end, bool hold_fold, bool allow_close) {
    if (x >= end || B(x) != '<') return false;
    if (end - x > kHeldMax) return false;
    for (int close = 0; close <= (allow_close ? 1 : 0); close++) {
        int64_t q = x + 1;
        if (close) {
            if (q == end) return true;
            if (B(q) != '/') continue;
            q++;
        }
        for (int i = 0; i < tg.n; i++) {
            int64_t r = q;
            bool ok = true;
            const int len = tg.len[i];
            for (int k = 0; k < len; k++) {
                if (r == end) break;
                int var, exo, cut;
                int c = match_letter(B, r, end, tg.name[i][k], hold_fold, var, exo, cut);
                if (cut) return true;
                if (!c) { ok = false; break; }
                r += c;
            }
            if (ok && r == end) return true;
        }
    }
    return false;
}

// Length of the str.isspace() code point that starts at p (0 if none); the sequence must fit in [p, end).
template <class A>
__device__ __forceinline__ int ws_len_at(const A& B, int64_t p, int64_t end) {
    uint8_t c = B(p);
    if ((c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x20)) return 1;
    if (c < 0xC2 || c > 0xE3) return 0;
    if (c == 0xC2) return (p + 1 < end && (B(p + 1) == 0x85 || B(p + 1) == 0xA0)) ? 2 : 0;
    if (p + 2 < end) {
        uint8_t d = B(p + 1), e = B(p + 2);
        if (c == 0xE1) return (d == 0x9A && e == 0x80) ? 3 : 0;
        if (c == 0xE2) {
            if (d == 0x80 && ((e >= 0x80 && e <= 0x8A) || e == 0xA8 || e == 0xA9 || e == 0xAF)) return 3;
            return (d == 0x81 && e == 0x9F) ? 3 : 0;
        }
        if (c == 0xE3) return (d == 0x80 && e == 0x80) ? 3 : 0;
    }
    return 0;
}

// Byte p of the text [0, n) belongs to a white-space code point (str.strip(), oai_proxy.py:139).
template <class A>
__device__ __forceinline__ bool is_white_byte(const A& B, int64_t p, int64_t n) {
    uint8_t c = B(p);
    if (c < 0x80) return (c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x20);
#pragma unroll
    for (int back = 0; back < 3; back++)
        if (p - back >= 0 && ws_len_at(B, p - back, n) > back) return true;
    return false;
}

// 4-bit mask of the bytes of w equal to '<'.
__device__ __forceinline__ uint32_t lt_nibble(uint32_t w) {
    uint32_t v = w ^ 0x3C3C3C3Cu;
    uint32_t t = (v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    t = ~(t | v | 0x7F7F7F7Fu);        // 0x80 in every byte of v that is zero
    return (t * 0x00204081u) >> 28;    // gather bits 7,15,23,31 -> 0..3
}

// non-zero iff some byte of w equals '<' (exact as a boolean; the borrow only smears above a real hit)
__device__ __forceinline__ uint32_t lt_any(uint32_t w) {
    const uint32_t v = w ^ 0x3C3C3C3Cu;
    return (v - 0x01010101u) & ~v & 0x80808080u;
}
__device__ __forceinline__ uint32_t lt_any32(const uint4& a, const uint4& b) {
    return lt_any(a.x) | lt_any(a.y) | lt_any(a.z) | lt_any(a.w) | lt_any(b.x) | lt_any(b.y) | lt_any(b.z) | lt_any(b.w);
}

__device__ __forceinline__ uint32_t lt_mask16(uint4 v) {
    return lt_nibble(v.x) | (lt_nibble(v.y) << 4) | (lt_nibble(v.z) << 8) | (lt_nibble(v.w) << 12);
}

// '<' bitmap of the 32-byte word `w` (absolute word index on the flat buffer `buf`, 16-byte aligned base,
// `total` readable bytes).  Bit b <-> byte 32*w + b.
__device__ __forceinline__ uint32_t load_lt_word(const uint8_t* buf, int64_t total, int64_t w) {
    int64_t a = w << 5;
    if (a + 32 <= total) {
        const uint4* p = reinterpret_cast<const uint4*>(buf + a);
        uint4 lo = p[0], hi = p[1];  // generic loads: the buffer may be global or shared memory
        return lt_mask16(lo) | (lt_mask16(hi) << 16);
    }
    uint32_t m = 0;
    for (int b = 0; b < 32 && a + b < total; b++)
        if (buf[a + b] == '<') m |= 1u << b;
    return m;
}

// bits b of a 32-byte word starting at unit-relative position wb whose position wb+b is >= p
__device__ __forceinline__ uint32_t mask_ge(int64_t wb, int64_t p) {
    int64_t d = p - wb;
    return d <= 0 ? 0xffffffffu : (d >= 32 ? 0u : (0xffffffffu << (int)d));
}

// Warp-cooperative byte copy dst[0..n) = src[0..n) (src and dst uniform across the warp).  16-byte stores on the aligned
// middle of dst; the source is read as aligned 16-byte vectors and realigned with a word select + funnel shift.
// [.., src_hi) bounds what may be read.
template <bool PREFETCH = false>
__device__ inline void warp_copy(uint8_t* dst, const uint8_t* src, int64_t n, const uint8_t* src_hi, int lane) {
    if (n <= 0) return;
    int64_t head = (16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15;
    if (head > n) head = n;
    if (lane < head) dst[lane] = src[lane];
    dst += head;
    src += head;
    n -= head;
    const int shift = (int)(reinterpret_cast<uintptr_t>(src) & 15);
    const uint4* s128 = reinterpret_cast<const uint4*>(src - shift);
    int64_t nvec = n >> 4;
    // the realigning read touches the next 16-byte vector when shift != 0: keep that inside [.., src_hi)
    if (shift != 0) {
        const int64_t room = (src_hi - reinterpret_cast<const uint8_t*>(s128)) >> 4;  // whole vectors readable
        if (nvec + 1 > room) nvec = room > 0 ? room - 1 : 0;
    }
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    const int wsel = shift >> 2, sh = (shift & 3) * 8;
    auto realign = [&](const uint4& A, const uint4& B) -> uint4 {
        uint32_t w0, w1, w2, w3, w4;
        switch (wsel) {
            case 0: w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; break;
            case 1: w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; break;
            case 2: w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; break;
            default: w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; break;
        }
        uint4 o;
        o.x = __funnelshift_r(w0, w1, sh);
        o.y = __funnelshift_r(w1, w2, sh);
        o.z = __funnelshift_r(w2, w3, sh);
        o.w = __funnelshift_r(w3, w4, sh);
        return o;
    };
    int64_t v = lane;
    if (shift == 0) {
        for (; v + 96 < nvec; v += 128) {  // four independent rounds in flight
            const uint4 a0 = s128[v], a1 = s128[v + 32], a2 = s128[v + 64], a3 = s128[v + 96];
            d4[v] = a0;
            d4[v + 32] = a1;
            d4[v + 64] = a2;
            d4[v + 96] = a3;
        }
        for (; v < nvec; v += 32) d4[v] = s128[v];
    } else {
        for (; v + 32 < nvec; v += 64) {  // two rounds (four loads) in flight before the first store waits
            const uint4 a0 = s128[v], b0 = s128[v + 1], a1 = s128[v + 32], b1 = s128[v + 33];
            d4[v] = realign(a0, b0);
            d4[v + 32] = realign(a1, b1);
        }
        for (; v < nvec; v += 32) {
            const uint4 a0 = s128[v], b0 = s128[v + 1];
            d4[v] = realign(a0, b0);
        }
    }
    for (int64_t i = (nvec << 4) + lane; i < n; i += 32) dst[i] = src[i];
}

// Lean variant for the kept runs of a unit (lengths below 2 GiB, typically 0.1-4 KiB): 32-bit arithmetic, and the word
// select of the realignment hoisted out of the loop (one loop per word offset) - the general warp_copy above spends
// ~300 warp instructions per call before and after its loops, which is most of the cost of a 1 KiB run.
template <int W>
__device__ __forceinline__ uint4 realign_w(const uint4& A, const uint4& B, int sh) {
    uint32_t w0, w1, w2, w3, w4;
    if (W == 0) { w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; }
    else if (W == 1) { w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; }
    else if (W == 2) { w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; }
    else { w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; }
    uint4 o;
    o.x = __funnelshift_r(w0, w1, sh);
    o.y = __funnelshift_r(w1, w2, sh);
    o.z = __funnelshift_r(w2, w3, sh);
    o.w = __funnelshift_r(w3, w4, sh);
    return o;
}
// (four vectors per lane per round with the neighbour's words taken by shuffle instead of the second load: cfg3's copy
// 1.08 -> 1.57 ms)
template <int W>
__device__ __forceinline__ void copy_vecs_w(uint4* __restrict__ d4, const uint4* __restrict__ s4, int nvec, int sh, int lane) {
    int v = lane;
    for (; v + 32 < nvec; v += 64) {  // two rounds (four loads) in flight before the first store waits
        const uint4 a0 = s4[v], b0 = s4[v + 1], a1 = s4[v + 32], b1 = s4[v + 33];
        d4[v] = realign_w<W>(a0, b0, sh);
        d4[v + 32] = realign_w<W>(a1, b1, sh);
    }
    if (v < nvec) {
        const uint4 a0 = s4[v], b0 = s4[v + 1];
        d4[v] = realign_w<W>(a0, b0, sh);
    }
}
// The bytes in front of the first aligned 16-byte store and behind the last one are loaded before the vectors and stored
// after them, so that a run costs one memory round trip, not three (a store waits for its load, and nothing issues past
// a waiting store: with ~400-byte runs the three dependent steps were most of cfg5's copy time).
__device__ __forceinline__ void warp_copy_fast(uint8_t* dst, const uint8_t* src, int n, const uint8_t* src_hi, int lane) {
    if (n <= 0) return;
    int head = (int)((16u - (uint32_t)reinterpret_cast<uintptr_t>(dst)) & 15u);
    if (head > n) head = n;
    uint8_t hb = 0, tb = 0;
    if (lane < head) hb = src[lane];
    uint8_t* const dst0 = dst;
    dst += head;
    src += head;
    n -= head;
    const int shift = (int)(reinterpret_cast<uintptr_t>(src) & 15);
    const uint4* s4 = reinterpret_cast<const uint4*>(src - shift);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    int nvec = n >> 4;
    if (shift) {
        // the realigning read touches the next 16-byte vector: keep that inside [.., src_hi)
        const int64_t room = (src_hi - reinterpret_cast<const uint8_t*>(s4)) >> 4;  // whole vectors readable
        if ((int64_t)nvec + 1 > room) nvec = room > 0 ? (int)(room - 1) : 0;
    }
    const int t0 = (nvec << 4) + lane;  // this lane's first byte behind the vectors
    if (t0 < n) tb = src[t0];
    if (shift == 0) {
        int v = lane;
        for (; v + 32 < nvec; v += 64) {
            const uint4 a0 = s4[v], a1 = s4[v + 32];
            d4[v] = a0;
            d4[v + 32] = a1;
        }
        if (v < nvec) d4[v] = s4[v];
    } else {
        const int sh = (shift & 3) * 8;
        switch (shift >> 2) {
            case 0: copy_vecs_w<0>(d4, s4, nvec, sh, lane); break;
            case 1: copy_vecs_w<1>(d4, s4, nvec, sh, lane); break;
            case 2: copy_vecs_w<2>(d4, s4, nvec, sh, lane); break;
            default: copy_vecs_w<3>(d4, s4, nvec, sh, lane); break;
        }
    }
    if (lane < head) dst0[lane] = hb;
    if (t0 < n) dst[t0] = tb;
    for (int i = t0 + 32; i < n; i += 32) dst[i] = src[i];  // only when the vectors were cut short at the end of the buffer
}

}  // namespace qb
