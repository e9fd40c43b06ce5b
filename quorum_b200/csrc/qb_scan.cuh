// Single-pass prefix sums across thread blocks: decoupled look-back over tiles taken by ticket.
#pragma once
#include "qb_common.cuh"

namespace qb {

__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// Decoupled look-back over tiles (single pass prefix sum).  state word = status (2 bits) | value (62 bits).
// Called by one full warp.  Returns the exclusive prefix of `aggregate` over all earlier tiles.
// ---------------------------------------------------------------------------------------------------------------
constexpr unsigned long long kStAgg = 1ull << 62, kStPre = 2ull << 62, kStMask = (1ull << 62) - 1;

__device__ inline int64_t lookback_exclusive(unsigned long long* tile_state, int64_t tile, int64_t aggregate, int lane) {
    if (tile == 0) {
        if (lane == 0) atomicExch(&tile_state[0], kStPre | (unsigned long long)aggregate);
        return 0;
    }
    if (lane == 0) atomicExch(&tile_state[tile], kStAgg | (unsigned long long)aggregate);
    int64_t excl = 0;
    int64_t j = tile - 1;
    for (;;) {
        const int64_t idx = j - lane;
        unsigned long long v = idx >= 0 ? ld_volatile_u64(&tile_state[idx]) : kStPre;
        const unsigned st = (unsigned)(v >> 62);
        const unsigned has_p = __ballot_sync(kFull, st == 2);
        const unsigned inval = __ballot_sync(kFull, st == 0);
        const int firstp = has_p ? __ffs(has_p) - 1 : 32;
        const unsigned need = firstp >= 31 ? 0xffffffffu : ((2u << firstp) - 1u);
        if (inval & need) {  // a predecessor has not published yet: poll again, without hogging the issue slots
            __nanosleep(40);
            continue;
        }
        int64_t val = (lane <= firstp) ? (int64_t)(v & kStMask) : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(kFull, val, d);
        excl += val;
        if (has_p) break;
        j -= 32;
    }
    if (lane == 0) atomicExch(&tile_state[tile], kStPre | (unsigned long long)(excl + aggregate));
    return excl;
}

}  // namespace qb
