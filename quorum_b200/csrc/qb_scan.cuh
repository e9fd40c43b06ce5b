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

// ---------------------------------------------------------------------------------------------------------------
// Two-level look-back.  The one-level form above assumes that a tile finds a published PREFIX a few tiles back.  With a
// warp per tile there are ~4,700 tiles in flight, every one of them in the "aggregate only" state for most of its
// life, so a look-back would walk ~150 dependent 32-wide steps.  Here tiles are grouped by 32: a tile sums the
// aggregates of the tiles before it in its own group (one 32-wide load) and looks back over GROUP words, one per 32
// tiles, so the walk is ~32x shorter.  The tile with the last index of a group has just summed all the others: it
// publishes the group's aggregate, looks back over the groups before it and publishes the group's inclusive prefix.
// No atomics: every word has one writer, readers poll it (volatile).
// ---------------------------------------------------------------------------------------------------------------
struct Lookback2 {
    unsigned long long* state;   // per tile:  status | aggregate
    unsigned long long* gstate;  // per group: status | aggregate or inclusive prefix
};

// exclusive prefix over the groups before `g` (read only); called by one full warp
__device__ inline int64_t group_lookback(const unsigned long long* gstate, int64_t g, int lane) {
    int64_t excl = 0;
    int64_t j = g - 1;
    while (j >= 0) {
        const int64_t idx = j - lane;
        unsigned long long v = idx >= 0 ? ld_volatile_u64(&gstate[idx]) : kStPre;
        const unsigned st = (unsigned)(v >> 62);
        const unsigned has_p = __ballot_sync(kFull, st == 2);
        const unsigned inval = __ballot_sync(kFull, st == 0);
        const int firstp = has_p ? __ffs(has_p) - 1 : 32;
        const unsigned need = firstp >= 31 ? 0xffffffffu : ((2u << firstp) - 1u);
        if (inval & need) {  // a group before this one is not complete yet
            __nanosleep(60);
            continue;
        }
        int64_t val = (lane <= firstp) ? (int64_t)(v & kStMask) : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(kFull, val, d);
        excl += val;
        if (has_p) break;
        j -= 32;
    }
    return excl;
}

// Called by one full warp.  Returns the exclusive prefix of `aggregate` over all earlier tiles.  ntiles = number of tiles
// of the call (the last group may be short).
__device__ inline int64_t lookback2_exclusive(const Lookback2& L, int64_t tile, int64_t aggregate, int64_t ntiles, int lane) {
    const int64_t g = tile >> 5;
    const int k = (int)(tile & 31);
    int64_t gsize = ntiles - (g << 5);
    if (gsize > 32) gsize = 32;
    const bool closes = k == gsize - 1;  // the tile with the group's last index speaks for the group
    if (lane == 0) st_volatile_u64(&L.state[tile], kStAgg | (unsigned long long)aggregate);
    // the tiles before this one in its own group
    unsigned long long v = 0;
    if (lane < k) {
        const unsigned long long* p = &L.state[(g << 5) + lane];
        while (((v = ld_volatile_u64(p)) >> 62) == 0) __nanosleep(40);
        v &= kStMask;
    }
    int64_t within = (int64_t)v;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) within += __shfl_xor_sync(kFull, within, d);
    if (closes && g > 0 && lane == 0) st_volatile_u64(&L.gstate[g], kStAgg | (unsigned long long)(within + aggregate));
    const int64_t excl_g = group_lookback(L.gstate, g, lane);
    if (closes && lane == 0) st_volatile_u64(&L.gstate[g], kStPre | (unsigned long long)(excl_g + within + aggregate));
    return excl_g + within;
}

}  // namespace qb
