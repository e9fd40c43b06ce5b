// Tokenizer pass: the one kernel that looks at every input byte.  Fully data parallel, no dependence between steps, so it
// runs at HBM speed: a warp takes a 1 KiB step of the flat buffer, lane j loads bytes [32j, 32j+32) with two 16-byte
// loads, finds '<' with a SWAR zero-byte test, matches the tag table at the (rare) candidates and writes one 32-bit
// record per '<' in position order (warp-scan ranks).  Everything downstream (the per-unit automata, the hold-back test,
// the copy of kept runs) works from these sparse records and never rescans the text.
#pragma once
#include "qb_walk.cuh"

namespace qb {

// Exact tag test at x for a flat global buffer: word-compare fast path when the longest possible tag (69 bytes) is readable, else byte-wise.
// Returns -1 when the window holds a lead byte of a fold form (U+0130/0131/017F/212A): the caller defers it (TokDefer).
template <bool kDefer>
__device__ __forceinline__ int tag_at(const DevTags& tg, const uint8_t* buf, int64_t total, int64_t x, int& is_close, int& cls,
                                      int& exotic) {
    exotic = 0;
    if (x + 72 <= total) {
        int exrel;
        // fast_match_tag indexes with int: rebase so that the candidate sits at a small offset
        const int64_t base = x & ~(int64_t)3;
        const int len = fast_match_tag<kDefer>(tg, buf + base, (int)(x - base), 1 << 30, is_close, cls, exrel);  // -1: deferred
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
constexpr int kTokDefer = 32 + 2 * kRecPerStep;
struct __align__(8) TokWarpScratch {
    int64_t dpos[kTokDefer];   // deferred candidates (need the exact byte matcher): absolute position ...
    uint8_t drank[kTokDefer];  // ... and rank inside its step (= record slot)
    int ndef;                  // entries waiting in dpos / drank
    uint16_t list[kTokList];   // positions (inside the 2 KiB pair) of the '<' candidates, in position order
};

// index of record `rank` of step s in the tokenizer's record buffer: [steps][kRecInline] first, then [steps][the rest] from
// word r2 on (TokSrc::rec / rec2)
__device__ __forceinline__ int64_t rec_index(int64_t s, int rank, int64_t r2) {
    return rank < kRecInline ? s * kRecInline + rank : r2 + s * (kRecPerStep - kRecInline) + (rank - kRecInline);
}
__device__ __forceinline__ uint32_t pack_rec(int pos, int len, int is_close, int cls, int exotic) {
    return (uint32_t)pos | ((uint32_t)len << 10) | ((len ? (uint32_t)is_close : 2u) << 17) | ((uint32_t)cls << 19) | ((uint32_t)exotic << 26);
}

// The tag test for deferred candidates [first, first + n) of the warp's list, one per lane.  Matching a candidate where
// it is found runs ~80 instructions (the byte-wise matcher for fold spellings ~10x that) with only the lanes that hold
// one active - two or three of 32 in ordinary text - and that was half of the tokenizer's issue slots; collected
// across pairs and matched 32 at a time the same work is a few instructions per candidate.
__device__ __forceinline__ void tok_drain(const DevTags& tg, const uint8_t* __restrict__ bytes, int64_t total, uint32_t* __restrict__ rec,
                                       int64_t r2, const TokWarpScratch& W, int first, int n, int lane) {
    if (lane < n) {
        const int64_t x = W.dpos[first + lane];
        int is_close, cls, exotic;
        const int len = tag_at<false>(tg, bytes, total, x, is_close, cls, exotic);
        rec[rec_index(x >> 10, W.drank[first + lane], r2)] = pack_rec((int)(x & 1023), len, is_close, cls, exotic);
    }
}

// Warp-collective: lanes with `need` append (x, rank) to the warp's deferred list (counter in shared memory: the pairs of
// ordinary text never come here and pay no register for it).  The list is drained between pairs (tokenize_kernel), so
// it holds at most 31 + 2 * kRecPerStep entries.
__device__ __forceinline__ bool tok_defer(TokWarpScratch& W, bool need, int64_t x, int rank, int lane) {
    const unsigned m = __ballot_sync(kFull, need);
    if (m) {
        const int nd = W.ndef;
        __syncwarp();
        if (need) {
            const int k = nd + __popc(m & ((1u << lane) - 1u));
            W.dpos[k] = x;
            W.drank[k] = (uint8_t)rank;
        }
        if (lane == 0) W.ndef = nd + __popc(m);
        __syncwarp();
    }
    return m != 0;
}

// One 2 KiB pair of steps `pr` of the flat buffer, by one warp: lane j loads bytes [64j, 64j+64) with four 16-byte loads
// (lanes 0-15 read the first step, lanes 16-31 the second), so one ballot and one divergent "look at the candidates"
// branch serve two steps, and each lane has 64 bytes in flight.
//   sparse pair (no lane has more than one '<', the usual case): every lane with a candidate matches it, ranks straight
//     from the ballot;
//   dense pair: the candidates are written, in position order, to a list in shared memory (ranks from a warp scan of the
//     per-lane counts) and matched ONE PER LANE, so the expensive part (tag_at) runs once per 32 candidates instead of
//     once per candidate of the busiest lane (adversarial text has 40 '<' per KiB in places: cfg5's tokenizer 1.16 ->
//     0.65 ms); a candidate's record goes straight to its slot: its rank inside its step is its list index minus the
//     step's first index.  Candidates whose window holds a fold-form lead byte need the exact byte-wise matcher: they are
//     deferred to the warp's list (tok_defer) and matched 32 at a time between pairs (tok_drain): cfg5 0.69 -> 0.50 ms.
//   The sparse path defers EVERY candidate (position and rank are known, the record slot waits): cfg3's tokenizer
//     0.74 -> 0.69 ms (6.0 TB/s), cfg2 0.201 -> 0.192, cfg4 0.70 -> 0.68, cfg5 0.52 -> 0.48.  (Deferring only the
//     candidates that need the exact matcher there cost more in ballots than it saved.)
// Returns true when something was deferred.
// cnt / rec are indexed by absolute step (rec through rec_index); steps >= nsteps are not written, bytes at and beyond
// `total` are never read.  A step with more than
// kRecPerStep candidates gets cnt = kRecOverflow and no records.
__device__ __forceinline__ bool tokenize_pair(const DevTags& tg, const uint8_t* __restrict__ bytes, int64_t total, int64_t nsteps,
                                              int64_t pr, int lane, uint8_t* cnt, uint32_t* rec, int64_t r2, TokWarpScratch& W) {
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
        return false;
    }
    const int nr = __popc(lt0) + __popc(lt1);
    if (__all_sync(kFull, nr <= 1)) {  // sparse: one record per lane at most, ranks straight from the ballot
        const int tot = __popc(any & hm);
        bool need = false;
        int b = 0;
        const int rank = __popc(any & hm & (0xffffffffu >> (31 - lane))) - 1;
        if (nr && step < nsteps) {
            b = lt0 ? __ffs(lt0) - 1 : 32 + __ffs(lt1) - 1;
            need = true;
        }
        if ((lane & 15) == 0 && step < nsteps) cnt[step] = (uint8_t)tot;  // at most 16
        return tok_defer(W, need, a + b, rank, lane);
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
    bool deferred = false;
    for (int i0 = 0; i0 < n_list; i0 += 32) {
        const int i = i0 + lane;
        bool need = false;
        int64_t x = 0;
        int rank = 0;
        if (i < n_list) {
            const int pos = W.list[i];
            const int sl = pos >> 10;
            const int64_t st = (pr << 1) + sl;
            if (st < nsteps) {
                int is_close, cls, exotic;
                x = (pr << 11) + pos;
                const int len = tag_at<true>(tg, bytes, total, x, is_close, cls, exotic);
                rank = i - (sl ? first1 : 0);
                need = len < 0;
                if (!need) rec[rec_index(st, rank, r2)] = pack_rec(pos & 1023, len, is_close, cls, exotic);
            }
        }
        deferred |= tok_defer(W, need, x, rank, lane);
    }
    if ((lane & 15) == 0 && step < nsteps) cnt[step] = (uint8_t)((half ? bad1 : bad0) ? kRecOverflow : (half ? c1 : c0));
    __syncwarp();
    return deferred;
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
    const int64_t r2 = max_steps * kRecInline;  // where the second record array starts (words)
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    TokWarpScratch& W = s_tok[threadIdx.x >> 5];
    if (lane == 0) W.ndef = 0;
    __syncwarp();
    for (int64_t pr = (begin >> 11) + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); pr < npairs; pr += nwarps) {
        {   // this warp's next 2 KiB: get its sixteen lines into L2 now, one iteration of work ahead of their use
            const int64_t an = ((pr + nwarps) << 11) + lane * 128;
            if (lane < 16 && an + 128 <= total) asm volatile("prefetch.global.L2 [%0];" ::"l"(bytes + an));
        }
        if (!tokenize_pair(tg, bytes, total, nsteps, pr, lane, cnt, rec, r2, W)) continue;
        int nd = W.ndef;
        if (nd >= 32) {  // full warps' worth of deferred candidates
            do {
                nd -= 32;
                tok_drain(tg, bytes, total, rec, r2, W, nd, 32, lane);
            } while (nd >= 32);
            __syncwarp();
            if (lane == 0) W.ndef = nd;
            __syncwarp();
        }
    }
    if (W.ndef) tok_drain(tg, bytes, total, rec, r2, W, 0, W.ndef, lane);
}

}  // namespace qb
