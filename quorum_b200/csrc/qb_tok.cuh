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

// Per-warp scratch of the tokenizer (shared memory): candidates waiting for their tag test.  Two lists in one array:
// fresh candidates grow from index 0 up (every '<' found, position + rank inside its step), candidates that turned out to
// need the exact byte-wise matcher grow from the last index down.  Between pairs at most 31 of each wait; a pair adds at
// most 2 * kRecPerStep fresh ones.
constexpr int kTokDefer = 64 + 2 * kRecPerStep;
struct __align__(8) TokWarpScratch {
    int64_t dpos[kTokDefer];   // absolute position of the '<' ...
    uint8_t drank[kTokDefer];  // ... and rank inside its step (= record slot)
    int ndef;                  // fresh entries: [0, ndef)
    int nex;                   // entries for the exact matcher: [kTokDefer - nex, kTokDefer)
};

// index of record `rank` of step s in the tokenizer's record buffer: [steps][kRecInline] first, then [steps][the rest] from
// word r2 on (TokSrc::rec / rec2)
__device__ __forceinline__ int64_t rec_index(int64_t s, int rank, int64_t r2) {
    return rank < kRecInline ? s * kRecInline + rank : r2 + s * (kRecPerStep - kRecInline) + (rank - kRecInline);
}
__device__ __forceinline__ uint32_t pack_rec(int pos, int len, int is_close, int cls, int exotic) {
    return (uint32_t)pos | ((uint32_t)len << 10) | ((len ? (uint32_t)is_close : 2u) << 17) | ((uint32_t)cls << 19) | ((uint32_t)exotic << 26);
}

// The tag test for deferred candidates, one per lane.  Matching a candidate where it is found runs ~80 instructions (the
// byte-wise matcher for fold spellings ~10x that) with only the lanes that hold one active - two or three of 32 in
// ordinary text - and that was half of the tokenizer's issue slots; collected across pairs and matched 32 at a time the
// same work is a few instructions per candidate.  Two stages, each with all lanes on the same code: the word-compare
// matcher for fresh candidates [first, first + n); those whose window holds a fold-form lead byte move to the second list
// (returns how many), which tok_drain_exact works off with the byte-wise matcher.
__device__ __forceinline__ int tok_drain_fast(const DevTags& tg, const uint8_t* __restrict__ bytes, int64_t total, uint32_t* __restrict__ rec,
                                              int64_t r2, TokWarpScratch& W, int first, int n, int nex, int lane) {
    bool need = false;
    int64_t x = 0;
    int rank = 0;
    if (lane < n) {
        x = W.dpos[first + lane];
        rank = W.drank[first + lane];
        int is_close, cls, exotic;
        const int len = tag_at<true>(tg, bytes, total, x, is_close, cls, exotic);
        need = len < 0;
        if (!need) rec[rec_index(x >> 10, rank, r2)] = pack_rec((int)(x & 1023), len, is_close, cls, exotic);
    }
    const unsigned m = __ballot_sync(kFull, need);  // (every lane has read its entry: the slots may be reused now)
    if (need) {
        const int k = kTokDefer - 1 - (nex + __popc(m & ((1u << lane) - 1u)));
        W.dpos[k] = x;
        W.drank[k] = (uint8_t)rank;
    }
    __syncwarp();
    return __popc(m);
}
__device__ __forceinline__ void tok_drain_exact(const DevTags& tg, const uint8_t* __restrict__ bytes, int64_t total, uint32_t* __restrict__ rec,
                                                int64_t r2, const TokWarpScratch& W, int first, int n, int lane) {
    if (lane < n) {
        const int64_t x = W.dpos[first + lane];
        const PtrAcc A{bytes};
        const int is_close = bytes[x + 1] == '/';
        int cls = 0;
        int64_t ex;
        const int64_t tend = match_tag(tg, A, x, total, is_close != 0, cls, ex);
        const int len = tend >= 0 ? (int)(tend - x) : 0;
        rec[rec_index(x >> 10, W.drank[first + lane], r2)] = pack_rec((int)(x & 1023), len, is_close, len ? cls : 0, len && ex >= 0);
    }
}

// Warp-collective: lanes with `need` append (x, rank) to the warp's list of fresh candidates (counter in shared memory:
// the pairs without a '<' never come here and pay no register for it).  The list is drained between pairs
// (tokenize_kernel), so it holds at most 31 + 2 * kRecPerStep entries.
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
//   Every '<' found goes to the warp's list of fresh candidates with its position and its rank inside its step (the
//   record slot it will fill) and is matched later, 32 at a time (tok_drain_fast / tok_drain_exact):
//   sparse pair (no lane has more than one '<', the usual case): ranks straight from the ballot;
//   dense pair: ranks from a warp scan of the per-lane counts, every lane writes its candidates in position order.
//   History of this part: candidates matched where they are found by the lane that holds them; dense pairs compacted into
//   a per-pair list and matched one per lane (cfg5 1.16 -> 0.65 ms); candidates that need the byte-wise matcher deferred
//   across pairs (0.69 -> 0.50); every candidate of a sparse pair deferred (cfg3 0.74 -> 0.69 = 6.0 TB/s); every
//   candidate of a dense pair too, and the byte-wise matcher as a second stage of the drain (cfg5 0.475 -> 0.396).
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
    // every candidate straight to the warp's list of fresh candidates, in position order, with its rank inside its step
    const bool ok0 = !bad0 && (pr << 1) < nsteps, ok1 = !bad1 && (pr << 1) + 1 < nsteps;
    const int first1 = ok0 ? c0 : 0;
    const int n_list = first1 + (ok1 ? c1 : 0);
    const int nd0 = W.ndef;
    __syncwarp();
    if (half ? ok1 : ok0) {
        int rk = half ? incl - nr - c0 : incl - nr;  // rank of this lane's first candidate inside its step
        int k = nd0 + (half ? first1 : 0) + rk;
        unsigned long long m = (unsigned long long)lt0 | ((unsigned long long)lt1 << 32);
        while (m) {
            const int b = __ffsll((long long)m) - 1;
            m &= m - 1;
            W.dpos[k] = a + b;
            W.drank[k] = (uint8_t)rk;
            k++;
            rk++;
        }
    }
    if (lane == 0) W.ndef = nd0 + n_list;
    const bool deferred = n_list > 0;
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
    grid_dep_wait();
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
    if (lane == 0) W.ndef = W.nex = 0;
    __syncwarp();
    for (int64_t pr = (begin >> 11) + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); pr < npairs; pr += nwarps) {
        {   // this warp's next 2 KiB: get its sixteen lines into L2 now, one iteration of work ahead of their use
            const int64_t an = ((pr + nwarps) << 11) + lane * 128;
            if (lane < 16 && an + 128 <= total) asm volatile("prefetch.global.L2 [%0];" ::"l"(bytes + an));
        }
        if (!tokenize_pair(tg, bytes, total, nsteps, pr, lane, cnt, rec, r2, W)) continue;
        int nd = W.ndef;
        if (nd >= 32) {  // full warps' worth of fresh candidates, and of candidates for the exact matcher
            int nx = W.nex;
            do {
                nd -= 32;
                nx += tok_drain_fast(tg, bytes, total, rec, r2, W, nd, 32, nx, lane);
                if (nx >= 32) {
                    nx -= 32;
                    tok_drain_exact(tg, bytes, total, rec, r2, W, kTokDefer - nx - 32, 32, lane);
                    __syncwarp();
                }
            } while (nd >= 32);
            if (lane == 0) {
                W.ndef = nd;
                W.nex = nx;
            }
            __syncwarp();
        }
    }
    {   // what is left at the end
        int nx = W.nex;
        const int nd = W.ndef;
        if (nd) nx += tok_drain_fast(tg, bytes, total, rec, r2, W, 0, nd, nx, lane);
        for (; nx > 0; nx -= 32) {
            const int n = nx < 32 ? nx : 32;
            tok_drain_exact(tg, bytes, total, rec, r2, W, kTokDefer - nx, n, lane);
        }
    }
}

}  // namespace qb
