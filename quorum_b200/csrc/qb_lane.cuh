// Lane-per-stream walk.  The tag automaton of one stream is scalar work over a handful of records; run by a whole warp
// (StreamRunWalker) it costs ~1000 warp instructions per stream with 31 lanes repeating lane 0.  Here every lane walks
// its own stream, 32 streams per warp, straight from the token records: same semantics, same descriptor, same carried
// state as StreamRunWalker::run (qb_walk.cuh), which stays for streams too long for one lane and for units that do not
// fit the descriptor.
#pragma once
#include "qb_walk.cuh"

namespace qb {

// Returns false (nothing written) when the stream must be walked by a warp: 2 GiB and more, a step whose records
// overflowed, more kept runs or hold intervals than the descriptor holds.
__device__ __forceinline__ bool stream_walk_lane(const DevTags& tg, const uint8_t* buf, const int64_t* co, int64_t m, uint32_t flags,
                                                 const TokSrc& ts, const StreamState* st_in, StreamState* st_out, StreamDesc* dsc,
                                                 int64_t& total_out) {
    total_out = 0;
    int depth = 0, h = 0;
    if (st_in) {
        depth = st_in->depth;
        h = st_in->held_len;
    }
    if (m <= 0) {  // nothing fed: state passes through, nothing to copy
        if (st_out) {
            StreamState so;
            so.depth = depth;
            so.tail_len = -1;
            so.held_len = (flags & 4u) ? 0 : (uint8_t)h;
            for (int i = 0; i < kHeldMax; i++) so.held[i] = i < so.held_len ? st_in->held[i] : 0;
            *st_out = so;
        }
        dsc->nruns = 0;
        dsc->held_kept = 0;
        return true;
    }
    const int64_t ustart = co[0];
    const int64_t n64 = co[m] - ustart;
    if (n64 >= (int64_t)1 << 31) return false;
    const int n = (int)n64;
    const uint8_t* heldp = st_in ? st_in->held : buf;
    const HeldAcc A{buf + ustart, heldp, h};
    const PtrAcc T{buf + ustart};

    int last_tok_end = -1, seg_start = 0, held_kept = 0, nruns = 0, cum = 0, nh = 0;
    bool held_unconsumed = false;
    int rs[kDescRuns], re[kDescRuns], hx[kDescHolds], he[kDescHolds];

    auto first_boundary_after = [&](int64_t pos) -> int64_t {
        int64_t lo = 1, hi = m + 1;
        {   // evenly cut chunks (the usual shape): the answer is where proportion says; two loads confirm it and replace the
            // eight dependent ones of the search
            int64_t g = pos >= 0 ? (pos * m) / (n > 0 ? n : 1) + 1 : 1;
            if (g < 1) g = 1;
            if (g > m) g = m;
            const int64_t cg = co[g] - ustart, cp = co[g - 1] - ustart;
            if (cg > pos && cp <= pos) return g;
            if (cg > pos) hi = g; else lo = g + 1;
        }
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (co[mid] - ustart > pos) hi = mid;
            else lo = mid + 1;
        }
        return lo;
    };
    auto recognised = [&](bool is_close, int x, int tend, int ex) -> bool {
        if (is_close) return depth > 0;
        if (depth > 0) return true;
        if (tg.hold_enabled && ex < 0) return true;
        const int64_t lo = tg.hold_enabled ? ex : x;
        int64_t k = first_boundary_after(lo);
        return !(k <= m && co[k] - ustart < tend);
    };
    auto close_run = [&](int e) {
        if (e > seg_start) {
            if (nruns < kDescRuns) {
                rs[nruns] = seg_start;
                re[nruns] = e;
            }
            nruns++;
            cum += e - seg_start;
        }
    };
    auto on_tag = [&](int x, int tend, bool is_close, int ex) -> bool {
        if (!recognised(is_close, x, tend, ex)) return false;
        if (depth == 0) {
            close_run(x);
            depth = 1;
        } else {
            depth += is_close ? -1 : 1;
            if (depth == 0) seg_start = tend;
        }
        last_tok_end = tend;
        return true;
    };
    auto hold_candidate = [&](const auto& acc, int x) {  // see StreamRunWalker: a '<' that stays text at depth 0
        if (!tg.hold_enabled) return;
        const int c = (int)(co[first_boundary_after(x)] - ustart);
        if (c - x > kHeldMax) return;
        const int e = hold_interval_end(tg, acc, x, c, n);
        if (e < c) return;
        if (nh < kDescHolds) {
            hx[nh] = x;
            he[nh] = e;
        }
        nh++;
    };

    // ---- the carried partial tag (positions -h .. -1) ----
    if (h > 0) {
        int cls;
        int64_t ex;
        int64_t tend = match_tag(tg, A, -h, n, false, cls, ex);
        bool is_close = false;
        if (tend < 0 && depth > 0) {
            tend = match_tag(tg, A, -h, n, true, cls, ex);
            is_close = true;
        }
        if (tend >= 0 && recognised(is_close, -h, (int)tend, (int)ex)) {
            depth += is_close ? -1 : 1;
            last_tok_end = seg_start = (int)tend;
        } else {
            held_unconsumed = true;
            if (depth == 0) {
                held_kept = h;
                hold_candidate(A, -h);
            }
        }
    }
    cum = held_kept;

    // ---- the stream's token records, step by step ----
    if (n > 0) {
        const int64_t s_first = ustart >> 10, s_last = (ustart + n - 1) >> 10;
        // eight steps at a time: their record counts are fetched together and the record lines of the steps that have
        // any are pulled towards the SM before the first one is walked (a lane has nothing else to hide a miss behind)
        for (int64_t s0 = s_first; s0 <= s_last; s0 += 8) {
          unsigned long long cpack = 0;
#pragma unroll
          for (int i = 0; i < 8; i++)
              if (s0 + i <= s_last) cpack |= (unsigned long long)ts.count(s0 + i) << (8 * i);
          if (cpack == 0) continue;
#pragma unroll
          for (int i = 0; i < 8; i++)
              if ((cpack >> (8 * i)) & 0xFF) asm volatile("prefetch.global.L1 [%0];" ::"l"(ts.head(s0 + i)));
          // One loop over this lane's records of the eight steps, whatever step each is in: the lanes of a warp then
          // iterate max-over-lanes(records) times, not sum-over-steps(max-over-lanes(records in the step)).
          unsigned long long rest = cpack;
          const uint32_t *rp = nullptr, *rp2 = nullptr;
          int cnt = 0, q = 0, sb = 0;
          for (;;) {
                if (q == cnt) {  // next step of the eight that has records
                    if (rest == 0) break;
                    const int i = (__ffsll((long long)rest) - 1) >> 3;
                    cnt = (int)((rest >> (8 * i)) & 0xFF);
                    rest &= ~(0xFFull << (8 * i));
                    if (cnt == kRecOverflow) return false;
                    const int64_t sj = s0 + i;
                    rp = ts.head(sj);
                    rp2 = ts.tail(sj);
                    sb = (int)((sj << 10) - ustart);  // unit position of the step's first byte
                    q = 0;
                }
                if (q == kRecInline) rp = rp2;  // the step's records beyond the first eight live in the second array
                const uint32_t r = rp[q++];
                const int kind = (r >> 17) & 3;
                const int x = sb + (int)(r & 1023);
                const int tend = x + (int)((r >> 10) & 127);
                if (x < 0 || x >= n || x < last_tok_end) continue;  // another unit's / inside a consumed tag
                bool took = false;
                if (kind != 2 && tend <= n) {  // a complete tag inside the unit
                    int ex = -1;
                    if (((r >> 26) & 1) && depth == 0 && kind == 0) {  // rare: which letter uses a non-ASCII form?
                        int cls;
                        int64_t e64;
                        match_tag(tg, T, x, n, false, cls, e64);
                        ex = (int)e64;
                    }
                    took = on_tag(x, tend, kind == 1, ex);
                }
                if (!took && depth == 0) hold_candidate(T, x);
          }
        }
    }
    const int nclosed = nruns;     // runs ended by an open tag
    if (depth == 0) close_run(n);  // the final run, clipped later at the last release boundary
    const bool has_final = nruns > nclosed;
    if (nruns > kDescRuns || nh > kDescHolds) return false;

    // ---- the last boundary (c = n): inside a hold interval? ----
    int R_last = n, P_last = cum;
    for (int i = 0; i < nh; i++)
        if (hx[i] < n && n <= he[i]) {
            R_last = hx[i];
            P_last = cum - (n - R_last);
        }
    total_out = P_last;

    // ---- descriptor ----
    {
        StreamDesc& d = *dsc;
        d.nruns = nruns;
        d.total = P_last;
        d.held_kept = (held_kept && R_last > -h) ? held_kept : 0;  // carried literal bytes that are released
        d.depth = depth;
        d.last_tok_end = last_tok_end;
        d.rlast = R_last;
        d.flags = (held_unconsumed ? 1 : 0) | (has_final ? 2 : 0) | (held_kept ? 4 : 0);
        d.held_len = h;
        d.ustart = ustart;
#pragma unroll
        for (int i = 0; i < kDescRuns; i++) {  // unclipped; the copy clips the open-ended final run at rlast
            d.run_s[i] = i < nruns ? rs[i] : 0;
            d.run_e[i] = i < nruns ? re[i] : 0;
        }
#pragma unroll
        for (int i = 0; i < kDescHolds; i++) {
            d.hold_x[i] = i < nh ? hx[i] : INT32_MAX;
            d.hold_e[i] = i < nh ? he[i] : INT32_MIN;
        }
        if (held_kept && R_last > -h)
            for (int i = 0; i < h; i++) d.held[i] = heldp[i];
    }

    // ---- carried state (state_in may alias state_out: everything is read before the one write) ----
    if (st_out) {
        int hp = INT32_MIN;  // start of the carried tail (logical coords)
        if (!(flags & 4u)) {
            if (depth == 0) {
                if (R_last < n) hp = R_last;
            } else {
                int lo = n - kHeldMax;
                if (lo < 0) lo = 0;
                if (last_tok_end > lo) lo = last_tok_end;
                int p = INT32_MIN;
                for (int x = n - 1; x >= lo; x--)
                    if (T(x) == '<') {
                        p = x;
                        break;
                    }
                if (p == INT32_MIN && held_unconsumed && last_tok_end < 0 && n + h <= kHeldMax) p = -h;
                if (p != INT32_MIN && is_tag_prefix(tg, A, p, n, false, true)) hp = p;
            }
        }
        StreamState so;
        so.depth = depth;
        so.tail_len = last_tok_end >= 0 ? (n - last_tok_end) : -1;
        so.held_len = 0;
        for (int i = 0; i < kHeldMax; i++) so.held[i] = 0;
        if (hp != INT32_MIN) {
            so.held_len = (uint8_t)(n - hp);
            for (int i = 0; i < so.held_len; i++) so.held[i] = A(hp + i);
        }
        *st_out = so;
    }
    return true;
}

// Kept runs of the units that have more than a descriptor holds (kDescRuns): a bump-allocated pool in the call's scratch.
// The count kernel writes a unit's runs once, the copy kernel reads them back - no second walk over the records.  `top`
// is zeroed by the tokenizer kernel of the call; a unit that does not get its slots (pool exhausted) falls back to the
// warp walkers, which settle any number of runs 32 at a time.
struct RunPool {
    int2* runs = nullptr;
    unsigned long long* top = nullptr;
    int64_t cap = 0;  // slots
};
constexpr int kLaneCls = 16;       // backreference classes a lane can track (the default 5-tag table has 9)
constexpr uint32_t kDescPooled = 8;  // UnitDesc.flags: runs live in the pool at index run_s[0] | run_s[1] << 32

// One lane visits the records of the steps [s_first, s_last] in position order: f(r, x) with x = unit position of the
// '<' (records of other units in a shared step have x < 0 or end beyond n: the callee filters).  Eight steps at a time:
// their record counts are fetched together and the record lines of the steps that have any are pulled towards the SM
// before the first one is read (a lane has nothing else to hide a miss behind).  Returns false on an overflowed step.
template <class F>
__device__ __forceinline__ bool for_each_record(const TokSrc& ts, int64_t s_first, int64_t s_last, int64_t ustart, F&& f) {
    // (measured on cfg5: a two-batch-deep software pipeline of the count loads plus one flattened loop over the records of
    // a batch, as the stream walker has, made this slower - 1.09 vs 0.61 ms - so the loops stay nested)
    for (int64_t s0 = s_first; s0 <= s_last; s0 += 8) {
        unsigned long long cpack = 0;
#pragma unroll
        for (int i = 0; i < 8; i++)
            if (s0 + i <= s_last) cpack |= (unsigned long long)ts.count(s0 + i) << (8 * i);
        if (cpack == 0) continue;
#pragma unroll
        for (int i = 0; i < 8; i++)
            if ((cpack >> (8 * i)) & 0xFF) asm volatile("prefetch.global.L1 [%0];" ::"l"(ts.head(s0 + i)));
        for (int i = 0; i < 8; i++) {
            const int cnt = (int)((cpack >> (8 * i)) & 0xFF);
            if (cnt == 0) continue;
            if (cnt == kRecOverflow) return false;
            const int64_t sj = s0 + i;
            const uint32_t* rp = ts.head(sj);
            const int sb = (int)((sj << 10) - ustart);
            for (int q = 0; q < cnt; q++) {
                if (q == kRecInline) rp = ts.tail(sj);
                const uint32_t r = rp[q];
                f(r, sb + (int)(r & 1023));
            }
        }
    }
    return true;
}

// Lane-per-response strip_thinking_tags (same semantics as StripRunWalker<false>): liveness of an open tag ("a closer of
// its class follows", oai_proxy.py:137-139 as restated in SURVEY.md A.1) comes from one pass over the response's records
// that keeps the LAST closer position of every class (lc[], one int per class and thread in shared memory, strided by the
// block size) and counts the open tags; a second pass is the two-state automaton; `.strip()` looks at the two ends of
// the run list.  Any number of records; a response that can have more than kDescRuns kept runs (runs <= open tags + 1)
// takes that many slots of the run pool before the second pass and writes its runs there, otherwise they stay in
// registers and go to the descriptor.  Returns false (nothing written) when the response needs the warp walker: 2 GiB
// and more, overflowed record slots, more classes than kLaneCls, pool exhausted.
__device__ __forceinline__ bool strip_walk_lane(const DevTags& tg, const uint8_t* buf, int64_t ustart, int64_t n64, const TokSrc& ts,
                                                UnitDesc* dsc, int64_t& kept_len, int* lc, int lc_stride, const RunPool& pool) {
    kept_len = 0;
    if (n64 <= 0) {  // empty response: nothing to walk, nothing to copy
        dsc->nruns = 0;
        dsc->flags = 0;
        return true;
    }
    if (n64 >= (int64_t)1 << 31 || tg.ncls > kLaneCls) return false;
    const int n = (int)n64;
    const PtrAcc T{buf + ustart};
    const int64_t s_first = ustart >> 10, s_last = (ustart + n - 1) >> 10;
    // (prefetching the two ends of the response for .strip() here made the kernel slower: cfg3 0.217 -> 0.238 ms)
    // pass A: last closer of every class, number of open tags
    for (int c = 0; c < tg.ncls; c++) lc[c * lc_stride] = 0;
    int n_open = 0;
    if (!for_each_record(ts, s_first, s_last, ustart, [&](uint32_t r, int x) {
            const int kind = (r >> 17) & 3;
            if (kind == 2 || x < 0 || x + (int)((r >> 10) & 127) > n) return;
            if (kind == 1) lc[((r >> 19) & 127) * lc_stride] = x + 1;
            else n_open++;
        }))
        return false;
    const bool pooled = n_open + 1 > kDescRuns;
    int2* pr = nullptr;
    unsigned long long at = 0;
    if (pooled) {
        if (!pool.runs) return false;
        at = atomicAdd(pool.top, (unsigned long long)(n_open + 1));
        if (at + (unsigned long long)(n_open + 1) > (unsigned long long)pool.cap) return false;
        pr = pool.runs + at;
    }
    // pass B: the automaton over the tags in position order
    int rs[kDescRuns], re[kDescRuns];
    int nruns = 0, state = -1, seg_start = 0;
    auto emit = [&](int s, int e) {
        if (pooled) {
            pr[nruns] = make_int2(s, e);
        } else {
            rs[nruns] = s;
            re[nruns] = e;
        }
        nruns++;
    };
    if (n_open > 0)
        for_each_record(ts, s_first, s_last, ustart, [&](uint32_t r, int x) {
            const int kind = (r >> 17) & 3;
            const int tend = x + (int)((r >> 10) & 127);
            if (kind == 2 || x < 0 || tend > n) return;
            const int cls = (r >> 19) & 127;
            if (state < 0) {
                if (kind == 0 && lc[cls * lc_stride] > x) {  // live: a closer of its class starts later
                    if (x > seg_start) emit(seg_start, x);
                    state = cls;
                }
            } else if (kind == 1 && cls == state) {
                seg_start = tend;
                state = -1;
            }
        });
    if (n > seg_start) emit(seg_start, n);
    // .strip(): white space goes from the front of the first run that keeps anything and from the back of the last
    auto get = [&](int i) -> int2 {
        return pooled ? pr[i] : make_int2(rs[i], re[i]);
    };
    auto put = [&](int i, int s, int e) {
        if (pooled) {
            pr[i] = make_int2(s, e);
        } else {
            rs[i] = s;
            re[i] = e;
        }
    };
    int lead = -1;
    for (int i = 0; i < nruns && lead < 0; i++) {
        const int2 run = get(i);
        int f = -1;
        for (int x = run.x; x < run.y; x++)
            if (!is_white_byte(T, x, n)) {
                f = x;
                break;
            }
        if (f >= 0) {
            put(i, f, run.y);
            lead = i;
        } else {
            put(i, run.x, run.x);  // all white
        }
    }
    if (lead >= 0) {
        for (int i = nruns - 1; i >= lead; i--) {
            const int2 run = get(i);
            int l = -1;
            for (int x = run.y - 1; x >= run.x; x--)
                if (!is_white_byte(T, x, n)) {
                    l = x;
                    break;
                }
            if (l >= 0) {
                put(i, run.x, l + 1);
                break;
            }
            put(i, run.x, run.x);
        }
    }
    int64_t kl = 0;
    if (pooled) {
        for (int i = 0; i < nruns; i++) {
            const int2 run = pr[i];
            kl += run.y - run.x;
        }
    } else {
        for (int i = 0; i < nruns; i++) kl += re[i] - rs[i];
    }
    kept_len = kl;
    UnitDesc& d = *dsc;
    d.nruns = nruns;
    d.total = (int32_t)kl;
    d.held_kept = 0;
    d.depth = 0;
    d.last_tok_end = -1;
    d.rlast = n;
    d.held_len = 0;
    d.flags = pooled ? kDescPooled : 0;
#pragma unroll
    for (int i = 0; i < kDescRuns; i++) {
        d.run_s[i] = (!pooled && i < nruns) ? rs[i] : 0;
        d.run_e[i] = (!pooled && i < nruns) ? re[i] : 0;
    }
    if (pooled) {
        d.run_s[0] = (int32_t)(uint32_t)(at & 0xffffffffull);
        d.run_s[1] = (int32_t)(uint32_t)(at >> 32);
    }
    return true;
}

}  // namespace qb
