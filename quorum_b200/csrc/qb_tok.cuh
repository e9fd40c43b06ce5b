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

__device__ __forceinline__ uint32_t lt_any16(const uint4& v) { return lt_any(v.x) | lt_any(v.y) | lt_any(v.z) | lt_any(v.w); }

// Per-warp scratch of the tokenizer (shared memory): the candidate list of a dense pair.
constexpr int kTokList = 2 * kRecPerStep;
struct TokWarpScratch {
    uint16_t list[kTokList];   // positions (inside the 2 KiB pair) of the '<' candidates, in position order
};

// One 2 KiB pair of steps `pr` of the flat buffer, by one warp: lane j loads bytes [64j, 64j+64) with four 16-byte loads
// (lanes 0-15 read the first step, lanes 16-31 the second), so one ballot and one divergent "look at the candidates"
// branch serve two steps, and each lane has 64 bytes in flight.
//   sparse pair (no lane has more than one '<', the usual case): every lane with a candidate matches it, ranks straight
//     from the ballot;
//   dense pair: the candidates are written, in position order, to a list in shared memory (ranks from a warp scan of the
//     per-lane counts) and matched ONE PER LANE, so the expensive part (tag_at) runs once per 32 candidates instead of
//     once per candidate of the busiest lane (adversarial text has 40 '<' per KiB in places: cfg5's tokenizer 1.16 ->
//     0.65 ms); a candidate's record goes straight to its slot: its rank inside its step is its list index minus the
//     step's first index.
// cnt / rec are indexed by absolute step (the caller may pass pointers shifted so that the steps of a tile land in shared
// memory); steps >= nsteps are not written, bytes at and beyond `total` are never read.  A step with more than
// kRecPerStep candidates gets cnt = kRecOverflow and no records.
__device__ __forceinline__ void tokenize_pair(const DevTags& tg, const uint8_t* __restrict__ bytes, int64_t total, int64_t nsteps,
                                              int64_t pr, int lane, uint8_t* cnt, uint32_t* rec, TokWarpScratch& W) {
    const int half = lane >> 4;
    const unsigned hm = half ? 0xFFFF0000u : 0x0000FFFFu;  // the lanes of this lane's step
    const int64_t a = (pr << 11) + lane * 64;
    uint32_t lt0 = 0, lt1 = 0;  // '<' positions in bytes [0, 32) and [32, 64) of the lane
    if (a + 64 <= total) {
        const uint4* p = reinterpret_cast<const uint4*>(bytes + a);
        // plain loads, not ld.cs: the candidates' tag match re-reads these lines from L1 (measured: cfg2 0.549 -> 0.541 ms,
        // cfg3 2.32 -> 2.27, cfg4 1.31 -> 1.29)
        const uint4 v0 = p[0], v1 = p[1], v2 = p[2], v3 = p[3];
        const uint32_t f0 = lt_any16(v0), f1 = lt_any16(v1), f2 = lt_any16(v2), f3 = lt_any16(v3);
        if (f0 | f1 | f2 | f3) {  // exact positions only where needed
            if (f0) lt0 = lt_mask16(v0);
            if (f1) lt0 |= lt_mask16(v1) << 16;
            if (f2) lt1 = lt_mask16(v2);
            if (f3) lt1 |= lt_mask16(v3) << 16;
        }
    } else {
        for (int b = 0; b < 32 && a + b < total; b++)
            if (bytes[a + b] == '<') lt0 |= 1u << b;
        for (int b = 0; b < 32 && a + 32 + b < total; b++)
            if (bytes[a + 32 + b] == '<') lt1 |= 1u << b;
    }
    const unsigned any = __ballot_sync(kFull, (lt0 | lt1) != 0);
    const int64_t step = (pr << 1) + half;
    if (any == 0) {  // the common case: no '<' in these 2 KiB
        if ((lane & 15) == 0 && step < nsteps) cnt[step] = 0;
        return;
    }
    const int nr = __popc(lt0) + __popc(lt1);
    if (__all_sync(kFull, nr <= 1)) {  // sparse: one record per lane at most, ranks straight from the ballot
        const int tot = __popc(any & hm);
        if (nr) {
            const int b = lt0 ? __ffs(lt0) - 1 : 32 + __ffs(lt1) - 1;
            int is_close, cls, exotic;
            const int len = tag_at(tg, bytes, total, a + b, is_close, cls, exotic);
            const uint32_t r = (uint32_t)((lane & 15) * 64 + b) | ((uint32_t)len << 10) | ((len ? (uint32_t)is_close : 2u) << 17) |
                               ((uint32_t)cls << 19) | ((uint32_t)exotic << 26);
            if (step < nsteps) rec[step * kRecPerStep + __popc(any & hm & (0xffffffffu >> (31 - lane))) - 1] = r;
        }
        if ((lane & 15) == 0 && step < nsteps) cnt[step] = (uint8_t)tot;  // at most 16
        return;
    }
    // dense: ranks from a warp scan, candidates compacted into the list, one per lane per round
    int incl = nr;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += t;
    }
    const int c0 = __shfl_sync(kFull, incl, 15), c1 = __shfl_sync(kFull, incl, 31) - c0;  // candidates of the two steps
    const bool bad0 = c0 > kRecPerStep, bad1 = c1 > kRecPerStep;
    const int first1 = bad0 ? 0 : c0;                 // list index of the second step's first candidate
    const int n_list = first1 + (bad1 ? 0 : c1);
    {
        const bool bad = half ? bad1 : bad0;
        int idx = half ? first1 + (incl - nr - c0) : incl - nr;
        if (!bad) {
            unsigned long long m = (unsigned long long)lt0 | ((unsigned long long)lt1 << 32);
            while (m) {
                const int b = __ffsll((long long)m) - 1;
                m &= m - 1;
                W.list[idx++] = (uint16_t)(lane * 64 + b);
            }
        }
    }
    __syncwarp();
    for (int i = lane; i < n_list; i += 32) {
        const int pos = W.list[i];
        const int sl = pos >> 10;
        int is_close, cls, exotic;
        const int len = tag_at(tg, bytes, total, (pr << 11) + pos, is_close, cls, exotic);
        const uint32_t r = (uint32_t)(pos & 1023) | ((uint32_t)len << 10) | ((len ? (uint32_t)is_close : 2u) << 17) |
                           ((uint32_t)cls << 19) | ((uint32_t)exotic << 26);
        const int64_t st = (pr << 1) + sl;
        if (st < nsteps) rec[st * kRecPerStep + (i - (sl ? first1 : 0))] = r;
    }
    if ((lane & 15) == 0 && step < nsteps) cnt[step] = (uint8_t)((half ? bad1 : bad0) ? kRecOverflow : (half ? c1 : c0));
    __syncwarp();
}

// The byte range to tokenize is [off[idx ? idx[0] : 0], off[idx ? idx[n] : n)): the units of this call (a chained call
// covers a slice of the flat buffer, see QB200_CHAIN).  A warp takes 2 KiB (two steps) per iteration.
__global__ void __launch_bounds__(256, 5) tokenize_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ bytes,
                                                       const int64_t* __restrict__ off, const int64_t* __restrict__ idx, int64_t n,
                                                       uint8_t* __restrict__ cnt, uint32_t* __restrict__ rec, int64_t max_steps,
                                                       unsigned long long* __restrict__ zero, int64_t zero_words) {
    __shared__ TokWarpScratch s_tok[8];
    // housekeeping for the kernel that follows on the stream: its look-back words and ticket start at zero
    if (blockIdx.x == 0)
        for (int64_t i = threadIdx.x; i < zero_words; i += blockDim.x) zero[i] = 0;
    const int lane = threadIdx.x & 31;
    const int64_t begin = idx ? off[idx[0]] : off[0];
    const int64_t total = idx ? off[idx[n]] : off[n];
    int64_t nsteps = (total + 1023) >> 10;
    if (nsteps > max_steps) nsteps = max_steps;
    const int64_t npairs = (nsteps + 1) >> 1;
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t pr = (begin >> 11) + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); pr < npairs; pr += nwarps) {
        {   // this warp's next 2 KiB: get its sixteen lines into L2 now, one iteration of work ahead of their use
            const int64_t an = ((pr + nwarps) << 11) + lane * 128;
            if (lane < 16 && an + 128 <= total) asm volatile("prefetch.global.L2 [%0];" ::"l"(bytes + an));
        }
        tokenize_pair(tg, bytes, total, nsteps, pr, lane, cnt, rec, s_tok[threadIdx.x >> 5]);
    }
}

}  // namespace qb
