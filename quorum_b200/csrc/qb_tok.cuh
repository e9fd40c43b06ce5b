// Tokenizer pass: the one kernel that looks at every input byte.  Fully data parallel, no dependence between steps, so it
// runs at HBM speed: a warp takes a 1 KiB step of the flat buffer, lane j loads bytes [32j, 32j+32) with two 16-byte
// loads, finds '<' with a SWAR zero-byte test, matches the tag table at the (rare) candidates and writes one 32-bit
// record per '<' in position order (warp-scan ranks).  Everything downstream (the per-unit automata, the hold-back test,
// the copy of kept runs) works from these sparse records and never rescans the text.
#pragma once
#include "qb_walk.cuh"

namespace qb {

// Exact tag test at x for a flat global buffer: word-compare fast path when the longest possible tag (69 bytes) is readable, else byte-wise.
__device__ __forceinline__ int tag_at(const DevTags& tg, const uint8_t* buf, int64_t total, int64_t x, int& is_close, int& cls,
                                      int& exotic) {
    exotic = 0;
    if (x + 72 <= total) {
        int exrel;
        // fast_match_tag indexes with int: rebase so that the candidate sits at a small offset
        const int64_t base = x & ~(int64_t)3;
        const int len = fast_match_tag(tg, buf + base, (int)(x - base), 1 << 30, is_close, cls, exrel);
        exotic = exrel != 0xFF;
        return len;
    }
    const PtrAcc A{buf};
    int64_t ex;
    is_close = 0;
    int64_t tend = match_tag(tg, A, x, total, false, cls, ex);
    if (tend < 0) {
        is_close = 1;
        tend = match_tag(tg, A, x, total, true, cls, ex);
    }
    if (tend < 0) return 0;
    exotic = ex >= 0;
    return (int)(tend - x);
}

// The byte range to tokenize is [off[idx ? idx[0] : 0], off[idx ? idx[n] : n)): the units of this call (a chained call
// covers a slice of the flat buffer, see QB200_CHAIN).  Bytes at and beyond the end are never read.
__global__ void __launch_bounds__(256) tokenize_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ bytes,
                                                       const int64_t* __restrict__ off, const int64_t* __restrict__ idx, int64_t n,
                                                       uint8_t* __restrict__ cnt, uint32_t* __restrict__ rec, int64_t max_steps) {
    const int lane = threadIdx.x & 31;
    const int64_t begin = idx ? off[idx[0]] : off[0];
    const int64_t total = idx ? off[idx[n]] : off[n];
    int64_t nsteps = (total + 1023) >> 10;
    if (nsteps > max_steps) nsteps = max_steps;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t step = (begin >> 10) + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); step < nsteps; step += nwarps) {
        const int64_t a = (step << 10) + lane * 32;
        {   // this warp's next step: get its eight lines into L2 now, one iteration of work ahead of their use
            const int64_t an = ((step + nwarps) << 10) + lane * 128;
            if (lane < 8 && an + 128 <= total) asm volatile("prefetch.global.L2 [%0];" ::"l"(bytes + an));
        }
        uint32_t lt = 0;
        if (a + 32 <= total) {
            const uint4* p = reinterpret_cast<const uint4*>(bytes + a);
            const uint4 lo = __ldcs(p), hi = __ldcs(p + 1);  // streamed: every byte is read exactly once by this pass
            if (lt_any32(lo, hi)) lt = lt_mask16(lo) | (lt_mask16(hi) << 16);  // exact positions only where needed
        } else {
            for (int b = 0; b < 32 && a + b < total; b++)
                if (bytes[a + b] == '<') lt |= 1u << b;
        }
        const unsigned any = __ballot_sync(kFull, lt != 0);
        if (any == 0) {  // the common case: no '<' in this KiB
            if (lane == 0) cnt[step] = 0;
            continue;
        }
        // records of this lane's word (a 32-byte word holds at most 10 three-byte tags; more than 8 '<' is an overflow)
        uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0, r5 = 0, r6 = 0, r7 = 0;
        int n = 0;
        bool many = false;
        while (lt) {
            const int b = __ffs(lt) - 1;
            lt &= lt - 1;
            int is_close, cls, exotic;
            const int len = tag_at(tg, bytes, total, a + b, is_close, cls, exotic);
            const uint32_t r = (uint32_t)(lane * 32 + b) | ((uint32_t)len << 10) | ((len ? (uint32_t)is_close : 2u) << 17) |
                               ((uint32_t)cls << 19) | ((uint32_t)exotic << 26);
            if (n == 0) r0 = r;
            else if (n == 1) r1 = r;
            else if (n == 2) r2 = r;
            else if (n == 3) r3 = r;
            else if (n == 4) r4 = r;
            else if (n == 5) r5 = r;
            else if (n == 6) r6 = r;
            else if (n == 7) r7 = r;
            else many = true;
            n++;
        }
        int incl, tot;
        if (__all_sync(kFull, n <= 1)) {  // one record per lane at most: ranks straight from the ballot
            incl = __popc(any & (0xffffffffu >> (31 - lane)));
            tot = __popc(any);
        } else {
            incl = n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int t = __shfl_up_sync(kFull, incl, d);
                if (lane >= d) incl += t;
            }
            tot = __shfl_sync(kFull, incl, 31);
        }
        const bool bad = __any_sync(kFull, many) || tot > kRecPerStep;
        if (!bad && n > 0) {
            uint32_t* out = rec + step * kRecPerStep + (incl - n);
            out[0] = r0;
            if (n > 1) out[1] = r1;
            if (n > 2) out[2] = r2;
            if (n > 3) out[3] = r3;
            if (n > 4) out[4] = r4;
            if (n > 5) out[5] = r5;
            if (n > 6) out[6] = r6;
            if (n > 7) out[7] = r7;
        }
        if (lane == 0) cnt[step] = (uint8_t)(bad ? kRecOverflow : tot);
    }
}

}  // namespace qb
