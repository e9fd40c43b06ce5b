// Warp-cooperative strip_thinking_tags over the token records, with the automaton's state carried across the unit's steps
// by a SCAN OF FINITE-STATE TRANSITION VECTORS (BASELINE.json north star, step 2; semantics: SURVEY.md A.1,
// oai_proxy.py:137-139).  This is the walker for long responses: the lane walker of qb_lane.cuh visits a unit's records
// one after the other, which is what bounds the adversarial configuration (one lane walking the ~500 records of a 64 KB
// response while the other 31 are done); here lane j owns step j of the unit (1 KiB), 32 steps at a time:
//   A  last closer position of every backreference class (lanes in parallel, atomicMax in shared memory)
//   B  every lane folds the records of its step into a transition vector v[s] = state after the step when it is entered
//      in state s (states: OUT, IN(c) for each class c; a live open tag sends OUT to IN(c), a close tag of class c sends
//      IN(c) to OUT, everything else is the identity), five packed words, byte-parallel updates
//   C  scan of the vectors over the lanes: applied in lane order to the state carried in from the previous 32 steps, which
//      gives every lane its entry state (the composition of whole vectors, (v2 . v1)[s] = v2[v1[s]], in log2(32) rounds was
//      built and measured slower, see below)
//   D  every lane walks its records again from its entry state and emits the run ends (live open met in OUT) and run
//      starts (matching close met in IN(c)); warp prefix sums rank them, the k-th start pairs with the k-th end
//   E  empty runs dropped (compaction), `.strip()` at the two ends of the list, kept length; up to kDescRuns runs go to
//      the descriptor, more stay in the run pool where they were written
// Returns false (nothing usable written) when the unit is left to the run-list warp walker: fewer than one tag per eight
// steps (that walker only visits steps that hold records), more classes than kLaneCls, an overflowed step, 2 GiB and
// more, pool exhausted.
#pragma once
#include "qb_lane.cuh"

namespace qb {

struct TVec {
    uint32_t w[5];  // next state for entry states 0..19, one byte each (0 = OUT, c + 1 = IN(class c))
};
__device__ __forceinline__ TVec tvec_identity() {
    TVec v;
    v.w[0] = 0x03020100u;
    v.w[1] = 0x07060504u;
    v.w[2] = 0x0B0A0908u;
    v.w[3] = 0x0F0E0D0Cu;
    v.w[4] = 0x13121110u;
    return v;
}
__device__ __forceinline__ int tvec_get(const TVec& v, int s) {
    const uint32_t w = s < 4 ? v.w[0] : s < 8 ? v.w[1] : s < 12 ? v.w[2] : s < 16 ? v.w[3] : v.w[4];
    return (int)((w >> ((s & 3) * 8)) & 0xFFu);
}
// bytes of w equal to `val` become `to` (exact zero-byte test, no borrow smear)
__device__ __forceinline__ uint32_t bytes_replace(uint32_t w, uint32_t val, uint32_t to) {
    const uint32_t v = w ^ (val * 0x01010101u);
    uint32_t t = (v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    t = ~(t | v | 0x7F7F7F7Fu);            // 0x80 in every byte of v that is zero
    const uint32_t m = (t >> 7) * 0xFFu;   // 0xFF in those bytes
    return (w & ~m) | ((to * 0x01010101u) & m);
}
__device__ __forceinline__ void tvec_apply(TVec& v, int from, int to, int nwords) {  // then: state `from` -> `to`
#pragma unroll
    for (int i = 0; i < 5; i++)
        if (i < nwords) v.w[i] = bytes_replace(v.w[i], (uint32_t)from, (uint32_t)to);
}
// first / last non-white byte of [s, e) of the unit (warp-cooperative), -1 when all white
__device__ __forceinline__ int scan_first_nonwhite(const PtrAcc& T, int s, int e, int n, int lane) {
    for (int base = s; base < e; base += 32) {
        const int x = base + lane;
        const unsigned b = __ballot_sync(kFull, x < e && !is_white_byte(T, x, n));
        if (b) return base + __ffs(b) - 1;
    }
    return -1;
}
__device__ __forceinline__ int scan_last_nonwhite(const PtrAcc& T, int s, int e, int n, int lane) {
    for (int top = e; top > s; top -= 32) {
        const int x = top - 32 + lane;
        const unsigned b = __ballot_sync(kFull, x >= s && !is_white_byte(T, x, n));
        if (b) return top - 32 + (31 - __clz(b));
    }
    return -1;
}

// `lc`: kMaxCls words of shared memory of this warp; `stage`: kDescRuns + 1 int2 of shared memory of this warp
__device__ __noinline__ bool strip_scan_walk(const DevTags& tg, const uint8_t* buf, int64_t ustart, int64_t n64, const TokSrc& ts,
                                             UnitDesc* dsc, int64_t* kept_len, unsigned long long* lc, int2* stage,
                                             const RunPool& pool, int lane) {
    if (n64 <= 0 || n64 >= (int64_t)1 << 31 || tg.ncls > kLaneCls) return false;
    const int n = (int)n64;
    const int nwords = (tg.ncls + 1 + 3) >> 2;
    const PtrAcc T{buf + ustart};
    const int64_t s_first = ustart >> 10, s_last = (ustart + n - 1) >> 10;
    for (int c = lane; c < tg.ncls; c += 32) lc[c] = 0;
    __syncwarp();
    // ---- A: last closer per class, number of open tags ----
    int n_open = 0, n_close = 0;
    bool bad = false;
    for (int64_t sj = s_first + lane; sj <= s_last; sj += 32) {
        const int cnt = ts.count(sj);
        if (cnt == kRecOverflow) {
            bad = true;
            continue;
        }
        const uint32_t* rp = ts.head(sj);
        const int sb = (int)((sj << 10) - ustart);
        for (int q = 0; q < cnt; q++) {
            if (q == kRecInline) rp = ts.tail(sj);
            const uint32_t r = rp[q];
            const int kind = (r >> 17) & 3;
            const int x = sb + (int)(r & 1023);
            if (kind == 2 || x < 0 || x + (int)((r >> 10) & 127) > n) continue;
            if (kind == 1) {
                atomicMax(&lc[(r >> 19) & 127], (unsigned long long)(x + 1));
                n_close++;
            } else {
                n_open++;
            }
        }
    }
    if (__any_sync(kFull, bad)) return false;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        n_open += __shfl_xor_sync(kFull, n_open, d);
        n_close += __shfl_xor_sync(kFull, n_close, d);
    }
    __syncwarp();
    // Few tags over many steps: the run-list walker, which only visits the steps that hold records, is the faster one
    // (4,096 responses x 256 KiB on one B200: 6 blocks each 0.030 vs 0.105 ms, 128 blocks each 0.210 vs 0.175 ms and the
    // copy kernel 0.59 vs 0.31 ms, because this walker leaves its runs in the pool).
    if ((int64_t)(n_open + n_close) * 8 < s_last - s_first + 1) return false;
    // where the runs go: the stage (shared memory) when they surely fit a descriptor, else pool slots taken up front
    const bool pooled = n_open + 1 > kDescRuns;
    int2* runs = stage;
    unsigned long long at = 0;
    if (pooled) {
        if (!pool.runs) return false;
        if (lane == 0) at = atomicAdd(pool.top, (unsigned long long)(n_open + 1));
        at = __shfl_sync(kFull, at, 0);
        if (at + (unsigned long long)(n_open + 1) > (unsigned long long)pool.cap) return false;
        runs = pool.runs + at;
    }
    // ---- B-D, 32 steps at a time ----
    int carry = 0;       // state at the start of the group (0 = OUT)
    int kS = 1, kE = 0;  // run starts / run ends ranked so far (start 0 is the start of the unit)
    if (lane == 0) runs[0].x = 0;
    for (int64_t S0 = s_first; S0 <= s_last; S0 += 32) {
        const int64_t sj = S0 + lane;
        const int cnt = sj <= s_last ? ts.count(sj) : 0;
        if (!__any_sync(kFull, cnt > 0)) continue;  // 32 steps without a '<': every vector is the identity, the state carries
        const uint32_t *rh = ts.head(sj), *rt = ts.tail(sj);
        auto recq = [&](int q) { return q < kRecInline ? rh[q] : rt[q]; };
        const int sb = (int)((sj << 10) - ustart);
        // B: this step's transition vector
        TVec v = tvec_identity();
        for (int q = 0; q < cnt; q++) {
            const uint32_t r = recq(q);
            const int kind = (r >> 17) & 3;
            const int x = sb + (int)(r & 1023);
            if (kind == 2 || x < 0 || x + (int)((r >> 10) & 127) > n) continue;
            const int cls = (r >> 19) & 127;
            if (kind == 0) {
                if (lc[cls] > (unsigned long long)x) tvec_apply(v, 0, cls + 1, nwords);  // live open tag: OUT -> IN(cls)
            } else {
                tvec_apply(v, cls + 1, 0, nwords);                                       // close tag: IN(cls) -> OUT
            }
        }
        // C: scan of the vectors over the lanes: applied in lane order to the state carried into the group, every lane
        // keeps the state it is entered in.  32 dependent shuffle + lookup steps; the logarithmic form - five rounds of
        // (v2 . v1)[s] = v2[v1[s]] over whole vectors - was measured slower (count of 4,096 x 256 KiB dense responses 0.174 vs
        // 0.116 ms, cfg5 1.82 vs 1.74 ms): composing two vectors costs ~190 instructions, the lookup ten.
        int state = carry, run_state = carry;
        for (int l = 0; l < 32; l++) {
            if (lane == l) state = run_state;
            run_state = __shfl_sync(kFull, tvec_get(v, run_state), l);
        }
        const int carry_next = run_state;
        // D: events of this step from its entry state.  Two passes over the records: count, then (with the ranks) write.
        int nS = 0, nE = 0;
        {
            int st = state;
            for (int q = 0; q < cnt; q++) {
                const uint32_t r = recq(q);
                const int kind = (r >> 17) & 3;
                const int x = sb + (int)(r & 1023);
                if (kind == 2 || x < 0 || x + (int)((r >> 10) & 127) > n) continue;
                const int cls = (r >> 19) & 127;
                if (st == 0) {
                    if (kind == 0 && lc[cls] > (unsigned long long)x) {
                        nE++;
                        st = cls + 1;
                    }
                } else if (kind == 1 && cls + 1 == st) {
                    nS++;
                    st = 0;
                }
            }
        }
        int iS = nS, iE = nE;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int a = __shfl_up_sync(kFull, iS, d), b = __shfl_up_sync(kFull, iE, d);
            if (lane >= d) {
                iS += a;
                iE += b;
            }
        }
        int oS = kS + iS - nS, oE = kE + iE - nE;
        {
            int st = state;
            for (int q = 0; q < cnt; q++) {
                const uint32_t r = recq(q);
                const int kind = (r >> 17) & 3;
                const int x = sb + (int)(r & 1023);
                const int tend = x + (int)((r >> 10) & 127);
                if (kind == 2 || x < 0 || tend > n) continue;
                const int cls = (r >> 19) & 127;
                if (st == 0) {
                    if (kind == 0 && lc[cls] > (unsigned long long)x) {
                        runs[oE++].y = x;
                        st = cls + 1;
                    }
                } else if (kind == 1 && cls + 1 == st) {
                    runs[oS++].x = tend;
                    st = 0;
                }
            }
        }
        kS += __shfl_sync(kFull, iS, 31);
        kE += __shfl_sync(kFull, iE, 31);
        carry = carry_next;
    }
    // a live open tag always has its closer inside the unit, so the unit ends in OUT and the last run ends at n
    if (lane == 0) runs[kE].y = n;
    kE++;
    __syncwarp();
    if (carry != 0 || kE != kS) return false;  // cannot happen; never trust a state machine that says otherwise
    // ---- E: drop the empty runs (32 at a time, in place: entries only move down) ----
    int nruns = 0;
    for (int base = 0; base < kS; base += 32) {
        const int i = base + lane;
        int2 run = make_int2(0, 0);
        if (i < kS) run = runs[i];
        const bool keep = i < kS && run.y > run.x;
        const unsigned km = __ballot_sync(kFull, keep);
        __syncwarp();
        if (keep) runs[nruns + __popc(km & ((1u << lane) - 1u))] = run;
        nruns += __popc(km);
        __syncwarp();
    }
    // .strip(): white space goes from the front of the first run that keeps anything and from the back of the last
    int lead = -1;
    for (int i = 0; i < nruns && lead < 0; i++) {
        const int2 run = runs[i];
        const int f = scan_first_nonwhite(T, run.x, run.y, n, lane);
        __syncwarp();
        if (f >= 0) {
            if (lane == 0) runs[i] = make_int2(f, run.y);
            lead = i;
        } else if (lane == 0) {
            runs[i] = make_int2(run.x, run.x);  // all white
        }
        __syncwarp();
    }
    if (lead >= 0) {
        for (int i = nruns - 1; i >= lead; i--) {
            const int2 run = runs[i];
            const int l = scan_last_nonwhite(T, run.x, run.y, n, lane);
            __syncwarp();
            if (l >= 0) {
                if (lane == 0) runs[i] = make_int2(run.x, l + 1);
                __syncwarp();
                break;
            }
            if (lane == 0) runs[i] = make_int2(run.x, run.x);
            __syncwarp();
        }
    }
    int64_t kl = 0;
    for (int i = lane; i < nruns; i += 32) {
        const int2 run = runs[i];
        kl += run.y - run.x;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) kl += __shfl_xor_sync(kFull, kl, d);
    *kept_len = kl;
    UnitDesc& d = *dsc;
    if (lane == 0) {
        d.nruns = nruns;
        d.total = (int32_t)kl;
        d.held_kept = 0;
        d.depth = 0;
        d.last_tok_end = -1;
        d.rlast = n;
        d.held_len = 0;
        d.flags = pooled ? kDescPooled : 0;
    }
    __syncwarp();
    if (lane < kDescRuns) {
        int2 run = make_int2(0, 0);
        if (!pooled && lane < nruns) run = runs[lane];
        d.run_s[lane] = run.x;
        d.run_e[lane] = run.y;
    }
    __syncwarp();
    if (pooled && lane == 0) {
        d.run_s[0] = (int32_t)(uint32_t)(at & 0xffffffffull);
        d.run_s[1] = (int32_t)(uint32_t)(at >> 32);
    }
    __syncwarp();
    return true;
}

}  // namespace qb
