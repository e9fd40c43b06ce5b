// SSE output framing: every filtered delta becomes
//     b'data: ' + json.dumps(event) + b'\n\n'
// (src/quorum/oai_proxy.py:629-652 per-backend events, :847-862 the final event).  Everything in the event except
// `delta.content` is fixed per backend, so an event is  prefix[p] + json_escape(content) + suffix  with prefix/suffix
// rendered once on the host.  json_escape is CPython's ensure_ascii string encoder: '"' '\\' -> backslash forms,
// \n \r \t \b \f short forms, other C0 controls and 0x7F -> \u00xx, every non-ASCII code point -> \uXXXX (lower-case
// hex; code points above U+FFFF as a surrogate pair).  Expansion, not removal: sizes per character, warp prefix sums,
// scatter of the escaped bytes.
#pragma once
#include "qb_common.cuh"

namespace qb {

// escaped size of the character whose first byte is b (UTF-8 continuation bytes belong to their lead byte: 0)
__device__ __forceinline__ int esc_size(uint8_t b) {
    if (b < 0x20) return (b == 8 || b == 9 || b == 10 || b == 12 || b == 13) ? 2 : 6;
    if (b == '"' || b == '\\') return 2;
    if (b < 0x7F) return 1;
    if (b == 0x7F) return 6;
    if (b < 0xC0) return 0;
    return b < 0xF0 ? 6 : 12;
}

__device__ __forceinline__ uint8_t hex_digit(uint32_t v) { return (uint8_t)(v < 10 ? '0' + v : 'a' + (v - 10)); }

__device__ __forceinline__ void put_u4(uint8_t* o, uint32_t cp) {  // \uXXXX
    o[0] = '\\';
    o[1] = 'u';
    o[2] = hex_digit((cp >> 12) & 15);
    o[3] = hex_digit((cp >> 8) & 15);
    o[4] = hex_digit((cp >> 4) & 15);
    o[5] = hex_digit(cp & 15);
}

// writes the escaped form of the character at src[x] (x < n) to o.  The continuation bytes are fetched and the code point
// assembled before the size is looked at, so that lanes escaping different kinds of character share most of the code.
__device__ __forceinline__ void esc_write(const uint8_t* src, int64_t x, int64_t n, uint8_t* o) {
    const uint32_t b = src[x];
    const uint32_t c1 = (b >= 0xC0 && x + 1 < n) ? (uint32_t)(src[x + 1] & 0x3F) : 0u;
    const uint32_t c2 = (b >= 0xE0 && x + 2 < n) ? (uint32_t)(src[x + 2] & 0x3F) : 0u;
    const uint32_t c3 = (b >= 0xF0 && x + 3 < n) ? (uint32_t)(src[x + 3] & 0x3F) : 0u;
    const int sz = esc_size((uint8_t)b);
    if (sz == 1) {
        o[0] = (uint8_t)b;
    } else if (sz == 2) {
        o[0] = '\\';
        o[1] = b == '"' ? '"' : b == '\\' ? '\\' : b == 8 ? 'b' : b == 9 ? 't' : b == 10 ? 'n' : b == 12 ? 'f' : 'r';
    } else if (sz == 6) {
        const uint32_t cp = b < 0x80 ? b : b < 0xE0 ? (((b & 0x1F) << 6) | c1) : (((b & 0x0F) << 12) | (c1 << 6) | c2);
        put_u4(o, cp);
    } else if (sz == 12) {
        const uint32_t cp = (((b & 0x07) << 18) | (c1 << 12) | (c2 << 6) | c3) - 0x10000u;
        put_u4(o, 0xD800u + ((cp >> 10) & 0x3FF));
        put_u4(o + 6, 0xDC00u + (cp & 0x3FF));
    }
}

constexpr uint32_t kSkipEmpty = 16u;  // QB200_SKIP_EMPTY

// Deltas are short (tens of bytes): a warp per delta spends its time on the loads in front of the work, and 2 Mi deltas
// are 2 Mi warps.  A warp takes 32 consecutive items instead: short items one per lane (aligned 4-byte loads, a whole word
// of plain ASCII counted at once), the rare long one by the whole warp afterwards.
constexpr int kLaneItem = 512;  // bytes: longer items are taken by the whole warp

// esc_size as a 256-byte table in shared memory: the lanes of a warp look at different items, and the branches of
// esc_size run one after the other for every kind of character any of the 32 lanes holds
__device__ __forceinline__ void fill_esc_table(uint8_t* tab) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) tab[i] = (uint8_t)esc_size((uint8_t)i);
    __syncthreads();
}

// true when no byte of v needs an escape: none below 0x20, none at or above 0x7F, no '"', no backslash
__device__ __forceinline__ bool plain_ascii_word(uint32_t v) {
    const uint32_t lo = (v - 0x20202020u) & ~v;  // bit 7 of a byte set when a byte below 0x20 exists (bytes with bit 7 are caught by v itself)
    const uint32_t q = v ^ 0x22222222u, bs = v ^ 0x5C5C5C5Cu, del = v ^ 0x7F7F7F7Fu;
    const uint32_t z = ((q - 0x01010101u) & ~q) | ((bs - 0x01010101u) & ~bs) | ((del - 0x01010101u) & ~del);
    return ((v | lo | z) & 0x80808080u) == 0;
}

// escaped size of the bytes [a, b) of `in`, by one lane; `in` 4-byte aligned, `total` = bytes readable
__device__ __forceinline__ int64_t esc_size_lane(const uint8_t* __restrict__ in, int64_t a, int64_t b, int64_t total, const uint8_t* tab) {
    int64_t sum = 0;
    const uint32_t* in32 = reinterpret_cast<const uint32_t*>(in);
    for (int64_t w = a >> 2; (w << 2) < b; w++) {
        const int64_t x0 = w << 2;
        if (x0 + 4 > total) {  // the buffer's last, partial word
            for (int64_t x = max(x0, a); x < b; x++) sum += tab[in[x]];
            break;
        }
        const uint32_t v = in32[w];
        if (x0 >= a && x0 + 4 <= b && plain_ascii_word(v)) {
            sum += 4;
            continue;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int64_t x = x0 + k;
            if (x >= a && x < b) sum += tab[(v >> (8 * k)) & 0xFFu];
        }
    }
    return sum;
}

// warp per 32 items: framed length (0 for a skipped empty item)
__global__ void __launch_bounds__(256) sse_count_kernel(const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off,
                                                        int64_t n_items, const int32_t* __restrict__ item_prefix,
                                                        const int64_t* __restrict__ prefix_off, int32_t suffix_len,
                                                        uint32_t flags, int64_t* __restrict__ lens) {
    __shared__ uint8_t s_tab[256];
    fill_esc_table(s_tab);
    const int lane = threadIdx.x & 31;
    const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t ngroups = (n_items + 31) >> 5;
    const int64_t total = in_off[n_items];
    const bool aligned = (reinterpret_cast<uintptr_t>(in) & 3) == 0;
    for (int64_t g = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); g < ngroups; g += nw) {
        const int64_t it = (g << 5) + lane;
        const bool have = it < n_items;
        int64_t a = 0, b = 0, sum = 0;
        if (have) {
            a = in_off[it];
            b = in_off[it + 1];
        }
        const bool mine = have && aligned && b - a <= kLaneItem;
        if (mine) sum = esc_size_lane(in, a, b, total, s_tab);
        unsigned todo = __ballot_sync(kFull, have && !mine);
        while (todo) {
            const int j = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t aj = __shfl_sync(kFull, a, j), bj = __shfl_sync(kFull, b, j);
            int64_t t = 0;
            for (int64_t x = aj + lane; x < bj; x += 32) t += esc_size(in[x]);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(kFull, t, d);
            if (lane == j) sum = t;
        }
        if (have) {
            const int p = item_prefix ? item_prefix[it] : 0;
            const bool skip = (flags & kSkipEmpty) && b == a;
            lens[it] = skip ? 0 : (prefix_off[p + 1] - prefix_off[p]) + sum + suffix_len;
        }
    }
}

// one item by the whole warp: prefix, escaped content, suffix at o
__device__ __forceinline__ void sse_write_item(const uint8_t* __restrict__ in, int64_t a, int64_t b, const uint8_t* __restrict__ prefix,
                                               int64_t plen, const uint8_t* __restrict__ suffix, int32_t suffix_len, uint8_t* o, int lane) {
    for (int64_t i = lane; i < plen; i += 32) o[i] = prefix[i];
    o += plen;
    int64_t run = 0;
    for (int64_t x0 = a; x0 < b; x0 += 32) {
        const int64_t x = x0 + lane;
        const int sz = x < b ? esc_size(in[x]) : 0;
        int incl = sz;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += t;
        }
        if (sz) esc_write(in + a, x - a, b - a, o + run + incl - sz);
        run += __shfl_sync(kFull, incl, 31);
    }
    o += run;
    for (int i = lane; i < suffix_len; i += 32) o[i] = suffix[i];
}

// Warp per 32 consecutive items.  Their events are consecutive in the output, so the warp builds them in shared memory -
// prefixes and suffixes copied by all lanes item after item, the contents escaped one item per lane - and then moves the
// block to its place with aligned 16-byte stores (the staging offset has the output address's low four bits).  A group
// whose events do not fit the staging area is written item by item straight to memory (sse_write_item).
// The characters that need an escape are not written where they are met: a lane notes (where it is, where its escape
// goes) in its own short list, and after the walk every lane works its list off, entry j of all lanes side by side - met
// in place, every escape in any of the 32 items stalls the other 31 lanes for its ~100 instructions (ncu: 500 M warp
// instructions for 2 Mi deltas, 1.1 lanes active in the escape code; a warp-wide list filled by ballot: 377 M).
constexpr int kSseStage = 9984;             // bytes of shared memory per warp
constexpr int kSseWarps = 4;                // warps per block: 48 KiB of shared memory, four blocks per SM
constexpr int kSsePref = 2560, kSseSuf = 256;  // the block's copy of the prefixes / the suffix, when they fit
constexpr int kSseList = 12;                // deferred escapes per lane (a lane whose list is full works it off at once)
struct SseEsc {
    uint16_t src;  // bits 0-13: byte inside the item; 14-15: bytes of its item behind it (3 = three or more)
    uint16_t dst;  // inside the staging area
};
__global__ void __launch_bounds__(kSseWarps * 32) sse_write_kernel(const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off,
                                                        int64_t n_items, const int32_t* __restrict__ item_prefix,
                                                        const uint8_t* __restrict__ prefix_bytes,
                                                        const int64_t* __restrict__ prefix_off, int32_t n_prefix,
                                                        const uint8_t* __restrict__ suffix, int32_t suffix_len,
                                                        const int64_t* __restrict__ out_off, uint8_t* __restrict__ out) {
    __shared__ __align__(16) uint8_t s_stage[kSseWarps][kSseStage];
    __shared__ uint8_t s_pref[kSsePref], s_suf[kSseSuf], s_tab[256];
    __shared__ SseEsc s_list[kSseWarps][kSseList][32];  // [entry][lane]
    fill_esc_table(s_tab);
    // the fixed text around the contents, once per block (a lane then copies its event's prefix and suffix from shared
    // memory; read item after item from global memory by all lanes, those copies were most of the kernel's time)
    const int64_t pref_total = prefix_off[n_prefix];
    const bool cached = pref_total <= kSsePref && suffix_len <= kSseSuf;
    if (cached) {
        for (int i = threadIdx.x; i < (int)pref_total; i += blockDim.x) s_pref[i] = prefix_bytes[i];
        for (int i = threadIdx.x; i < suffix_len; i += blockDim.x) s_suf[i] = suffix[i];
    }
    __syncthreads();
    const uint8_t* const psrc = cached ? s_pref : prefix_bytes;
    const uint8_t* const ssrc = cached ? s_suf : suffix;
    const int lane = threadIdx.x & 31;
    uint8_t* const stage = s_stage[threadIdx.x >> 5];
    const int64_t nw = (int64_t)gridDim.x * kSseWarps;
    const int64_t ngroups = (n_items + 31) >> 5;
    const int64_t total = in_off[n_items];
    const bool aligned = (reinterpret_cast<uintptr_t>(in) & 3) == 0;
    for (int64_t g = (int64_t)blockIdx.x * kSseWarps + (threadIdx.x >> 5); g < ngroups; g += nw) {
        const int64_t it0 = g << 5, it = it0 + lane;
        const int m = (int)min((int64_t)32, n_items - it0);
        const bool have = lane < m;
        const int64_t g0 = out_off[it0], g1 = out_off[it0 + m];
        if (g1 == g0) continue;
        int64_t a = 0, b = 0, o0 = g0, o1 = g0;
        int p = 0;
        if (have) {
            a = in_off[it];
            b = in_off[it + 1];
            o0 = out_off[it];
            o1 = out_off[it + 1];
            p = item_prefix ? item_prefix[it] : 0;
        }
        const int shift = (int)(reinterpret_cast<uintptr_t>(out + g0) & 15);
        const bool fits = aligned && (g1 - g0) + shift <= kSseStage && __all_sync(kFull, b - a <= kLaneItem);
        if (!fits) {
            for (int j = 0; j < m; j++) {
                const int64_t aj = __shfl_sync(kFull, a, j), bj = __shfl_sync(kFull, b, j), oj = __shfl_sync(kFull, o0, j),
                              ej = __shfl_sync(kFull, o1, j);
                const int pj = __shfl_sync(kFull, p, j);
                if (ej == oj) continue;  // skipped
                sse_write_item(in, aj, bj, prefix_bytes + prefix_off[pj], prefix_off[pj + 1] - prefix_off[pj], suffix, suffix_len, out + oj, lane);
            }
            continue;
        }
        const int my = shift + (int)(o0 - g0);  // where this lane's event starts in the staging area
        const int64_t pa = prefix_off[p];
        const int plen = (int)(prefix_off[p + 1] - pa);
        if (have && o1 > o0) {  // this lane's prefix and suffix
            uint8_t* o = stage + my;
            const uint8_t* ps = psrc + pa;
#pragma unroll 4
            for (int i = 0; i < plen; i++) o[i] = ps[i];
            o = stage + my + (int)(o1 - o0) - suffix_len;
#pragma unroll 4
            for (int i = 0; i < suffix_len; i++) o[i] = ssrc[i];
        }
        // contents, one item per lane: plain bytes straight to their place, escapes to the lane's list
        {
            SseEsc (*const list)[32] = s_list[threadIdx.x >> 5];
            int nlist = 0;
            if (have && o1 > o0) {
                const uint8_t* const item = in + a;
                const int n = (int)(b - a);
                auto flush = [&]() {
                    for (int e = 0; e < nlist; e++) {
                        const SseEsc q = list[e][lane];
                        const int x = q.src & 0x3FFF;
                        esc_write(item, x, x + 1 + (q.src >> 14), stage + q.dst);  // continuation bytes only inside the item
                    }
                    nlist = 0;
                };
                int o = my + plen;  // staging offset of this lane's next output byte
                const uint32_t* in32 = reinterpret_cast<const uint32_t*>(in);
                for (int64_t w = a >> 2; (w << 2) < b; w++) {
                    const int xr0 = (int)((w << 2) - a);  // item position of the word's first byte (negative in the first word)
                    uint32_t v;
                    if ((w << 2) + 4 <= total) {
                        v = in32[w];
                    } else {  // the buffer's last, partial word
                        v = 0;
                        for (int k = 0; k < 4 && (w << 2) + k < total; k++) v |= (uint32_t)in[(w << 2) + k] << (8 * k);
                    }
                    // (a fast path for a word of four plain bytes made the kernel slower, 0.63 -> 0.68 ms: the lanes of a warp
                    // are in different items and both paths run whenever the lanes disagree)
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const int xr = xr0 + k;
                        if (xr < 0 || xr >= n) continue;
                        const uint8_t c = (uint8_t)(v >> (8 * k));
                        const int sz = s_tab[c];
                        if (sz == 1) {
                            stage[o] = c;
                        } else if (sz > 1) {
                            SseEsc q;
                            q.src = (uint16_t)(xr | (min(3, n - 1 - xr) << 14));
                            q.dst = (uint16_t)o;
                            list[nlist++][lane] = q;
                            if (nlist == kSseList) flush();
                        }
                        o += sz;
                    }
                }
            }
            // entry j of every lane's list side by side
            const SseEsc (*const clist)[32] = list;
            int nmax = nlist;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) nmax = max(nmax, __shfl_xor_sync(kFull, nmax, d));
            for (int e = 0; e < nmax; e++) {
                if (e < nlist) {
                    const SseEsc q = clist[e][lane];
                    const int x = q.src & 0x3FFF;
                    esc_write(in + a, x, x + 1 + (q.src >> 14), stage + q.dst);
                }
            }
        }
        __syncwarp();
        // the block to its place: bytes up to the first aligned address, 16-byte vectors, the bytes behind them
        {
            const int nb = (int)(g1 - g0);
            uint8_t* dst = out + g0;
            const uint8_t* src = stage + shift;
            int head = (16 - shift) & 15;
            if (head > nb) head = nb;
            if (lane < head) dst[lane] = src[lane];
            const int nvec = (nb - head) >> 4;
            const uint4* s4 = reinterpret_cast<const uint4*>(src + head);
            uint4* d4 = reinterpret_cast<uint4*>(dst + head);
            for (int v = lane; v < nvec; v += 32) d4[v] = s4[v];
            for (int i = head + (nvec << 4) + lane; i < nb; i += 32) dst[i] = src[i];
        }
        __syncwarp();
    }
}

}  // namespace qb
