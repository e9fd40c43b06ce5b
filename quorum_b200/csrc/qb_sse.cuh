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

// writes the escaped form of the character at src[x] (x < n) to o
__device__ __forceinline__ void esc_write(const uint8_t* src, int64_t x, int64_t n, uint8_t* o) {
    const uint8_t b = src[x];
    if (b < 0x80) {
        const int sz = esc_size(b);
        if (sz == 1) {
            o[0] = b;
        } else if (sz == 2) {
            o[0] = '\\';
            o[1] = b == '"' ? '"' : b == '\\' ? '\\' : b == 8 ? 'b' : b == 9 ? 't' : b == 10 ? 'n' : b == 12 ? 'f' : 'r';
        } else {
            put_u4(o, b);
        }
        return;
    }
    auto cont = [&](int64_t i) -> uint32_t { return i < n ? (uint32_t)(src[i] & 0x3F) : 0u; };
    if (b < 0xE0) {
        put_u4(o, ((uint32_t)(b & 0x1F) << 6) | cont(x + 1));
    } else if (b < 0xF0) {
        put_u4(o, ((uint32_t)(b & 0x0F) << 12) | (cont(x + 1) << 6) | cont(x + 2));
    } else {
        const uint32_t cp = (((uint32_t)(b & 0x07) << 18) | (cont(x + 1) << 12) | (cont(x + 2) << 6) | cont(x + 3)) - 0x10000u;
        put_u4(o, 0xD800u + ((cp >> 10) & 0x3FF));
        put_u4(o + 6, 0xDC00u + (cp & 0x3FF));
    }
}

constexpr uint32_t kSkipEmpty = 16u;  // QB200_SKIP_EMPTY

// warp per item: framed length (0 for a skipped empty item)
__global__ void __launch_bounds__(256) sse_count_kernel(const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off,
                                                        int64_t n_items, const int32_t* __restrict__ item_prefix,
                                                        const int64_t* __restrict__ prefix_off, int32_t suffix_len,
                                                        uint32_t flags, int64_t* __restrict__ lens) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t it = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); it < n_items; it += nw) {
        const int64_t a = in_off[it], b = in_off[it + 1];
        int64_t sum = 0;
        for (int64_t x = a + lane; x < b; x += 32) sum += esc_size(in[x]);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(kFull, sum, d);
        if (lane == 0) {
            const int p = item_prefix ? item_prefix[it] : 0;
            const bool skip = (flags & kSkipEmpty) && b == a;
            lens[it] = skip ? 0 : (prefix_off[p + 1] - prefix_off[p]) + sum + suffix_len;
        }
    }
}

// warp per item: prefix, escaped content, suffix at out_off[it]
__global__ void __launch_bounds__(256) sse_write_kernel(const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off,
                                                        int64_t n_items, const int32_t* __restrict__ item_prefix,
                                                        const uint8_t* __restrict__ prefix_bytes,
                                                        const int64_t* __restrict__ prefix_off, const uint8_t* __restrict__ suffix,
                                                        int32_t suffix_len, const int64_t* __restrict__ out_off,
                                                        uint8_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t it = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); it < n_items; it += nw) {
        const int64_t o0 = out_off[it], o1 = out_off[it + 1];
        if (o1 == o0) continue;  // skipped
        const int64_t a = in_off[it], b = in_off[it + 1];
        const int p = item_prefix ? item_prefix[it] : 0;
        const int64_t pa = prefix_off[p], plen = prefix_off[p + 1] - pa;
        uint8_t* o = out + o0;
        for (int64_t i = lane; i < plen; i += 32) o[i] = prefix_bytes[pa + i];
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
}

}  // namespace qb
