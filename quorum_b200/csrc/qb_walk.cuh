// Warp-cooperative walkers: one warp runs one unit (a response, or one stream's data of this call).
//
// Shared structure: the unit is visited in steps of 32 words x 32 bytes (1 KiB); lane j owns word j of the step and a
// 32-bit bitmap of the '<' bytes in it.  Only '<' positions can start a tag, so everything sequential (the tag
// automaton) runs over the sparse candidate list in position order, while the dense work (finding candidates,
// counting and copying kept bytes) is lane-parallel.
#pragma once
#include "qb_common.cuh"

namespace qb {

// Token records written by the tokenizer pass (qb_tok.cuh), one group per absolute 1 KiB step of the flat buffer:
//   cnt[step]  number of records (0..kRecPerStep), kRecOverflow when the step had more '<' than fit
//   record k of a step, in position order: the first kRecInline of every step sit side by side in `rec` (32 bytes per
//   step: a walker reads one sector per step, four steps per line - with all 48 slots of a step in one place every
//   step cost a 128-byte line of DRAM traffic for its one or two records), the rest in `rec2` (a walker keeps the
//   first kRecInline of a step in registers and fetches the rest, lane per record, only for the rare step that has more):
//     bits 0-9 position inside the step | 10-16 tag length in bytes (0 for a literal '<') | 17-18 kind
//     (0 open tag, 1 close tag, 2 '<' that starts no complete tag) | 19-25 backreference class | 26 the tag uses a
//     U+0130/U+0131/U+017F spelling (the walker re-reads the bytes when that matters)
constexpr int kRecPerStep = 48;
constexpr int kRecInline = 8;
constexpr int kRecOverflow = 255;
struct TokSrc {
    const uint8_t* cnt = nullptr;
    const uint32_t* rec = nullptr;   // [step][kRecInline]
    const uint32_t* rec2 = nullptr;  // [step][kRecPerStep - kRecInline]
    int64_t nsteps = 0;  // steps the record arrays were sized for (from the caller's byte bound, out_capacity)
    // Record count of step s.  A step beyond the arrays reads as "overflowed": when a caller states a byte bound below the
    // real input size the units out there are walked from their bytes instead of from records that were never written.
    __device__ __forceinline__ int count(int64_t s) const { return s < nsteps ? (int)cnt[s] : kRecOverflow; }
    __device__ __forceinline__ const uint32_t* head(int64_t s) const { return rec + s * kRecInline; }  // records 0..kRecInline-1
    // p[k] is record k of step s for k >= kRecInline
    __device__ __forceinline__ const uint32_t* tail(int64_t s) const { return rec2 + s * (kRecPerStep - kRecInline) - kRecInline; }
    __device__ __forceinline__ uint32_t get(int64_t s, int k) const { return k < kRecInline ? head(s)[k] : tail(s)[k]; }
};
__device__ __forceinline__ uint32_t rec_select(const uint4& a, const uint4& b, int k) {
    switch (k) {
        case 0: return a.x;
        case 1: return a.y;
        case 2: return a.z;
        case 3: return a.w;
        case 4: return b.x;
        case 5: return b.y;
        case 6: return b.z;
        default: return b.w;
    }
}
// What the walk of one unit leaves behind for the kernels that follow (per-chunk offsets, gather copy): the kept runs of
// the unit in unit coordinates plus the few scalars of the automaton's final state.  nruns < 0: the unit did not fit
// (more than kDescRuns runs, overflowed record slots, 2 GiB and more) and is redone by the byte-rescanning walkers.
constexpr int kDescRuns = 8;
constexpr int kDescHolds = 4;
struct UnitDesc {
    int32_t nruns, total, held_kept, depth, last_tok_end, rlast, flags, held_len;  // flags: 1 held '<' unconsumed, 2 final run open-ended, 4 carried bytes are literal text
    int32_t run_s[kDescRuns], run_e[kDescRuns];
};
static_assert(sizeof(UnitDesc) == 96, "descriptor layout");
// What a stream adds for the per-chunk offsets (stream_offsets_kernel reads only descriptors, never text or records)
struct StreamDesc : UnitDesc {
    int64_t ustart;  // flat position of the stream's first byte in this call
    uint8_t held[24];  // the carried bytes when they are released as text (held_kept > 0): the finish kernel writes them
    // '<' that stay text at depth 0 and can be held back by a feed boundary: a boundary c with hold_x < c <= hold_e
    // releases only the bytes before hold_x (unused slots never match)
    int32_t hold_x[kDescHolds], hold_e[kDescHolds];
};
static_assert(sizeof(StreamDesc) == 160, "descriptor layout");

// true when any step touched by the unit [ustart, ustart+n) overflowed its record slots (then the unit is walked
// the slow way, straight from its bytes)
__device__ inline bool unit_has_overflow(const TokSrc& ts, int64_t ustart, int64_t n, int lane) {
    if (!ts.cnt || n <= 0) return false;
    const int64_t s0 = ustart >> 10, s1 = (ustart + n - 1) >> 10;
    bool bad = false;
    for (int64_t s = s0 + lane; s <= s1; s += 32) bad |= ts.count(s) == kRecOverflow;
    return __any_sync(kFull, bad);
}

// ------------------------------------------------------------------------------------------------------------------
// strip_thinking_tags  (oai_proxy.py:120-139), SURVEY.md Appendix A.1
//
// States: OUT, or IN(c) for a backreference class c.  An open tag of class c met in OUT is LIVE iff a closer of class c
// starts later in the response (regex `<(..)>.*?</\1>` must find its closer, else the open tag is literal text); a live
// open tag deletes through the FIRST closer of its class; everything else is inert.  Pass A records the last closer
// position per class, pass B walks the tokens.  Kept text = the gaps between deleted regions, then .strip().
// ------------------------------------------------------------------------------------------------------------------
struct StripResult {
    int64_t first_pos;  // unit-relative position of the first kept non-white byte (valid when kept_len > 0)
    int64_t kept_len;   // output length after filter + strip
};

template <bool COPY>
struct StripWalker {
    const DevTags& tg;
    const uint8_t* buf;   // flat buffer base (16-byte aligned)
    int64_t total;        // readable bytes in buf
    int64_t ustart, n;    // unit = buf[ustart, ustart+n)
    unsigned long long* lastclose;  // per-warp shared array [kMaxCls]: 1 + position of the last closer of the class
    int lane;
    PtrAcc T;
    // count pass
    bool lead_done = false;
    int64_t cum = 0;
    StripResult res{0, 0};
    // copy pass
    uint8_t* dst = nullptr;

    __device__ StripWalker(const DevTags& tg_, const uint8_t* buf_, int64_t total_, int64_t ustart_, int64_t n_,
                           unsigned long long* lc, int lane_)
        : tg(tg_), buf(buf_), total(total_), ustart(ustart_), n(n_), lastclose(lc), lane(lane_), T{buf_ + ustart_} {}

    // '<' bitmap of this lane's word in step `s`, restricted to the unit; wb = unit-relative position of bit 0
    __device__ __forceinline__ uint32_t step_mask(int64_t w_first, int64_t w_last, int64_t s, int64_t& wb) const {
        int64_t w = w_first + s * 32 + lane;
        wb = (w << 5) - ustart;
        if (w > w_last) return 0;
        uint32_t m = load_lt_word(buf, total, w);
        m &= mask_ge(wb, 0);
        m &= ~mask_ge(wb, n);
        return m;
    }

    __device__ int64_t first_nonwhite(int64_t s, int64_t e) const {
        for (int64_t base = s; base < e; base += 32) {
            int64_t x = base + lane;
            bool nw = x < e && !is_white_byte(T, x, n);
            unsigned b = __ballot_sync(kFull, nw);
            if (b) return base + __ffs(b) - 1;
        }
        return -1;
    }
    __device__ int64_t last_nonwhite(int64_t s, int64_t e) const {
        for (int64_t top = e; top > s; top -= 32) {
            int64_t x = top - 32 + lane;
            bool nw = x >= s && !is_white_byte(T, x, n);
            unsigned b = __ballot_sync(kFull, nw);
            if (b) return top - 32 + (31 - __clz(b));
        }
        return -1;
    }

    // kept segment [s, e) in position order
    __device__ void emit(int64_t s, int64_t e) {
        if (s >= e) return;
        if (COPY) {
            if (e <= res.first_pos) return;
            if (s < res.first_pos) s = res.first_pos;
            int64_t take = e - s;
            if (take > res.kept_len - cum) take = res.kept_len - cum;
            if (take > 0) warp_copy<true>(dst + cum, T.p + s, take, buf + total, lane);
            cum += e - s;
        } else {
            if (!lead_done) {
                int64_t f = first_nonwhite(s, e);
                if (f < 0) return;  // all white: still leading
                lead_done = true;
                res.first_pos = f;
                s = f;
            }
            int64_t l = last_nonwhite(s, e);
            if (l >= 0) res.kept_len = cum + (l + 1 - s);
            cum += e - s;
        }
    }

    // walker state of pass B
    int state = -1;
    int64_t seg_start = 0;
    __device__ __forceinline__ void event(int64_t x, int64_t tend, int cls, bool is_close) {
        if (state < 0) {
            if (!is_close && lastclose[cls] > (unsigned long long)x) {
                emit(seg_start, x);
                state = cls;
            }
        } else if (is_close && cls == state) {
            seg_start = tend;
            state = -1;
        }
    }

    __device__ void run() {
        if (n <= 0) return;
        const int64_t w_first = ustart >> 5, w_last = (ustart + n - 1) >> 5;
        const int64_t nsteps = (w_last - w_first + 32) >> 5;
        for (int c = lane; c < tg.ncls; c += 32) lastclose[c] = 0;
        __syncwarp();
        // pass A: last closer per class
        for (int64_t s = 0; s < nsteps; s++) {
            int64_t wb;
            uint32_t m = step_mask(w_first, w_last, s, wb);
            while (m) {
                int b = __ffs(m) - 1;
                m &= m - 1;
                int cls;
                int64_t ex;
                if (match_tag(tg, T, wb + b, n, true, cls, ex) >= 0)
                    atomicMax(&lastclose[cls], (unsigned long long)(wb + b + 1));
            }
        }
        __syncwarp();
        // pass B: walk tokens in order
        for (int64_t s = 0; s < nsteps; s++) {
            int64_t wb;
            uint32_t m = step_mask(w_first, w_last, s, wb);
            uint32_t tok = 0;
            while (m) {  // keep only real tags
                int b = __ffs(m) - 1;
                m &= m - 1;
                int cls;
                int64_t ex;
                if (match_tag(tg, T, wb + b, n, false, cls, ex) >= 0 || match_tag(tg, T, wb + b, n, true, cls, ex) >= 0)
                    tok |= 1u << b;
            }
            unsigned act = __ballot_sync(kFull, tok != 0);
            while (act) {
                int src = __ffs(act) - 1;
                int64_t x = 0, tend = 0;
                int kc = 0;
                if (lane == src) {
                    int b = __ffs(tok) - 1;
                    tok &= tok - 1;
                    x = wb + b;
                    int cls;
                    int64_t ex;
                    tend = match_tag(tg, T, x, n, false, cls, ex);
                    kc = cls;
                    if (tend < 0) {
                        tend = match_tag(tg, T, x, n, true, cls, ex);
                        kc = cls | 0x100;
                    }
                }
                x = __shfl_sync(kFull, x, src);
                tend = __shfl_sync(kFull, tend, src);
                kc = __shfl_sync(kFull, kc, src);
                event(x, tend, kc & 0xff, (kc & 0x100) != 0);
                act = __ballot_sync(kFull, tok != 0);
            }
        }
        emit(seg_start, n);
    }
};

// ------------------------------------------------------------------------------------------------------------------
// ThinkingTagFilter.feed / flush  (oai_proxy.py:278-371), SURVEY.md Appendix A.2, flat formulation
//
// The stream's text of this call is its chunks back to back, preceded logically by the carried partial tag.  Tags are
// recognised on the flat text: an open tag bumps depth at any depth, a close tag lowers it only at depth > 0; a byte is
// kept iff depth is 0 and it is not part of a consumed tag.  Chunking matters in exactly two places:
//   * at depth 0 a tag split by a feed boundary is only recognised if the part before the boundary passed the
//     hold-back test (str.lower() fold), i.e. contains no U+0130/U+0131/U+017F form before the boundary;
//   * the return value of feed k is the kept bytes in [R_{k-1}, R_k) where R_k is the chunk end c_k, or the position of
//     the last '<' before c_k when that '<' is kept text and [p, c_k) is a holdable partial open tag.
// ------------------------------------------------------------------------------------------------------------------
struct StreamResult {
    int64_t total_out;  // kept bytes released by this call
};

template <bool COPY>
struct StreamWalker {
    const DevTags& tg;
    const uint8_t* buf;
    int64_t total;
    const int64_t* co;  // absolute chunk offsets of this stream: co[0..m]
    int64_t m;
    int lane;
    uint32_t flags;
    // copy pass
    uint8_t* dst = nullptr;          // out_bytes + this stream's output base
    int64_t out_base = 0;            // absolute output offset of this stream
    int64_t* out_chunk_off = nullptr;  // global array, indexed by absolute chunk index
    int64_t chunk0 = 0;              // absolute index of this stream's first chunk

    __device__ StreamWalker(const DevTags& tg_, const uint8_t* buf_, int64_t total_, const int64_t* co_, int64_t m_,
                            int lane_, uint32_t flags_)
        : tg(tg_), buf(buf_), total(total_), co(co_), m(m_), lane(lane_), flags(flags_) {}

    // smallest k in [1, m] with co[k] - ustart > pos, or m+1
    __device__ int64_t first_boundary_after(int64_t ustart, int64_t pos) const {
        int64_t lo = 1, hi = m + 1;
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (co[mid] - ustart > pos) hi = mid;
            else lo = mid + 1;
        }
        return lo;
    }

    __device__ StreamResult run(const StreamState* st_in, StreamState* st_out) {
        StreamState si;
        if (st_in) {
            si = *st_in;
        } else {
            si.depth = 0;
            si.tail_len = -1;
            si.held_len = 0;
            for (int i = 0; i < kHeldMax; i++) si.held[i] = 0;
        }
        StreamResult res{0};
        if (m <= 0) {  // nothing fed: state passes through
            if (COPY && st_out && lane == 0) {
                if (flags & 4u) si.held_len = 0;
                for (int i = si.held_len; i < kHeldMax; i++) si.held[i] = 0;
                si.tail_len = -1;
                *st_out = si;
            }
            return res;
        }
        const int64_t ustart = co[0], n = co[m] - co[0];
        const int h = si.held_len;
        const uint8_t* heldp = st_in ? st_in->held : buf;
        const HeldAcc A{buf + ustart, heldp, h};
        const PtrAcc T{buf + ustart};

        int32_t depth = si.depth;
        int64_t last_tok_end = -1;   // end of the last consumed tag (unit coords), -1 if none
        int64_t skip_until = 0;      // bytes before this position belong to a consumed tag
        int64_t seg_start = 0;       // start of the current kept run (valid while depth == 0)
        int held_kept = 0;           // carried bytes that turned out to be literal kept text
        bool held_unconsumed = false;  // the carried '<' was not consumed as a tag (still a partial-tag candidate)

        // x = position of the tag's '<'.  At depth 0 an open tag cut by a feed boundary c is only seen by the reference
        // if [x, c) passed the hold-back test: never when holding is off (empty tag list), and not when a form the
        // str.lower() fold rejects (first one at ex) lies before c.
        auto recognised = [&](bool is_close, int64_t x, int64_t tend, int64_t ex) -> bool {
            if (is_close) return depth > 0;
            if (depth > 0) return true;
            if (tg.hold_enabled && ex < 0) return true;
            const int64_t lo = tg.hold_enabled ? ex : x;
            int64_t k = first_boundary_after(ustart, lo);
            return !(k <= m && co[k] - ustart < tend);
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
            if (tend >= 0 && recognised(is_close, -h, tend, ex)) {
                depth += is_close ? -1 : 1;
                last_tok_end = skip_until = seg_start = tend;
            } else {
                held_unconsumed = true;
                if (depth == 0) held_kept = h;
            }
        }

        int64_t cum = held_kept;   // kept bytes before the current step (carried literal included)
        int64_t out_pos = held_kept;  // copy pass: next output byte (carried literal bytes come first)
        int64_t kb = COPY ? 1 : m;  // next feed boundary to account (co[kb] is the end of chunk kb-1)
        int64_t R_last = 0;        // release boundary of the last feed
        int64_t P_last = 0;        // kept bytes before R_last
        bool have_last = false;    // this lane accounted the last boundary
        uint32_t prev_lt31 = 0, prev_kept31 = 0;
        if (COPY && lane == 0) out_chunk_off[chunk0] = out_base;

        const int64_t w_first = (ustart >> 10) << 5;  // steps sit on the absolute 1 KiB grid (the tokenizer's grid)
        const int64_t w_last = n > 0 ? ((ustart + n - 1) >> 5) : (w_first - 1);
        const int64_t nsteps = n > 0 ? ((w_last - w_first + 32) >> 5) : 1;  // one (empty) step accounts empty chunks
        // feed boundaries are fetched one step ahead: cpre = co[pre_base + lane]
        int64_t pre_base = kb;
        int64_t cpre = (kb + lane <= m) ? co[kb + lane] - ustart : INT64_MAX;
        for (int64_t s = 0; s < nsteps; s++) {
            const int64_t w = w_first + s * 32 + lane;
            const int64_t wb = (w << 5) - ustart;
            const uint32_t range = (w <= w_last) ? (mask_ge(wb, 0) & ~mask_ge(wb, n)) : 0u;
            uint32_t lt = 0;
            uint32_t kept = depth == 0 ? mask_ge(wb, skip_until) : 0u;
            // one recognised-or-not tag at unit position x
            auto on_tag = [&](int64_t x, int64_t tend, bool is_close, int64_t ex) {
                if (!recognised(is_close, x, tend, ex)) return;
                if (depth == 0) {  // open tag at depth 0: the kept run ends at x
                    kept &= ~mask_ge(wb, x);
                    if (COPY) {
                        int64_t len = x - seg_start;
                        if (len > 0) warp_copy<true>(dst + out_pos, T.p + seg_start, len, buf + total, lane);
                        out_pos += len > 0 ? len : 0;
                    }
                    depth = 1;
                } else {
                    depth += is_close ? -1 : 1;
                    if (depth == 0) {
                        kept |= mask_ge(wb, tend);
                        seg_start = tend;
                    }
                }
                last_tok_end = skip_until = tend;
            };
            {
                lt = (w <= w_last) ? (load_lt_word(buf, total, w) & range) : 0u;
                uint32_t tok = 0;
                {
                    uint32_t mm = lt & mask_ge(wb, skip_until);
                    while (mm) {
                        int b = __ffs(mm) - 1;
                        mm &= mm - 1;
                        int cls;
                        int64_t ex;
                        if (match_tag(tg, T, wb + b, n, false, cls, ex) >= 0 || match_tag(tg, T, wb + b, n, true, cls, ex) >= 0)
                            tok |= 1u << b;
                    }
                }
                unsigned act = __ballot_sync(kFull, tok != 0);
                while (act) {
                    int src = __ffs(act) - 1;
                    int64_t x = 0, tend = 0, ex = -1;
                    int is_close = 0;
                    if (lane == src) {
                        int b = __ffs(tok) - 1;
                        tok &= tok - 1;
                        x = wb + b;
                        int cls;
                        tend = match_tag(tg, T, x, n, false, cls, ex);
                        if (tend < 0) {
                            tend = match_tag(tg, T, x, n, true, cls, ex);
                            is_close = 1;
                        }
                    }
                    x = __shfl_sync(kFull, x, src);
                    tend = __shfl_sync(kFull, tend, src);
                    ex = __shfl_sync(kFull, ex, src);
                    is_close = __shfl_sync(kFull, is_close, src);
                    on_tag(x, tend, is_close != 0, ex);
                    act = __ballot_sync(kFull, tok != 0);
                }
            }
            kept &= range;
            // exclusive prefix of kept counts over the words of this step (the count pass only needs it where the
            // unit's last feed boundary lies: the released total depends on nothing else)
            const int pc = __popc(kept);
            int incl = pc, excl = 0, step_total;
            const bool need_scan = COPY || s == nsteps - 1;
            if (need_scan) {
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    int v = __shfl_up_sync(kFull, incl, d);
                    if (lane >= d) incl += v;
                }
                excl = incl - pc;
                step_total = __shfl_sync(kFull, incl, 31);
            } else {
                step_total = (int)__reduce_add_sync(kFull, (unsigned)pc);
            }
            const int64_t step_end = wb - (int64_t)lane * 32 + 1024 < n ? wb - (int64_t)lane * 32 + 1024 : n;
            const int64_t step_wb0 = wb - (int64_t)lane * 32;

            // ---- feed boundaries whose last byte lies in this step (c == 0: before any byte) ----
            for (; need_scan;) {
                const int64_t k = kb + lane;
                int64_t c;
                if (kb == pre_base) c = cpre;
                else c = k <= m ? co[k] - ustart : INT64_MAX;
                const bool in = k <= m && c <= step_end;
                // word (lane index inside this step) that holds byte c-1
                int lw = (in && c > 0) ? (int)((c - 1 - step_wb0) >> 5) : 0;
                const uint32_t kept_w = __shfl_sync(kFull, kept, lw);
                const uint32_t lt_w = __shfl_sync(kFull, lt, lw);
                const int excl_w = __shfl_sync(kFull, excl, lw);
                uint32_t kept_p = __shfl_sync(kFull, kept, lw > 0 ? lw - 1 : 0);
                uint32_t lt_p = __shfl_sync(kFull, lt, lw > 0 ? lw - 1 : 0);
                if (lw == 0) { kept_p = prev_kept31; lt_p = prev_lt31; }
                if (in) {
                    int64_t P = cum, R = c;
                    bool found_lt = false;
                    if (c > 0) {
                        const int64_t wbw = step_wb0 + (int64_t)lw * 32;
                        const int nb = (int)(c - wbw);  // 1..32 bytes of word lw lie before c
                        const uint32_t low = nb >= 32 ? 0xffffffffu : ((1u << nb) - 1u);
                        P = cum + excl_w + __popc(kept_w & low);
                        // last '<' before c, within the two words that can hold a partial tag
                        int64_t p = 0;
                        bool p_kept = false;
                        if (lt_w & low) {
                            int b = 31 - __clz(lt_w & low);
                            p = wbw + b;
                            p_kept = (kept_w >> b) & 1u;
                            found_lt = true;
                        } else if (lt_p) {
                            int b = 31 - __clz(lt_p);
                            p = wbw - 32 + b;
                            p_kept = (kept_p >> b) & 1u;
                            found_lt = true;
                        }
                        if (found_lt && p_kept && tg.hold_enabled && c - p <= kHeldMax &&
                            is_tag_prefix(tg, T, p, c, true, false)) {
                            R = p;
                            P -= (c - p);
                        }
                    }
                    // the carried partial is still the last '<' and still a holdable prefix
                    if (!found_lt && held_kept && tg.hold_enabled && c + h <= kHeldMax &&
                        is_tag_prefix(tg, A, -h, c, true, false)) {
                        R = -h;
                        P = 0;
                    }
                    if (COPY) out_chunk_off[chunk0 + k] = out_base + P;
                    if (k == m) { R_last = R; P_last = P; have_last = true; }
                }
                const unsigned inb = __ballot_sync(kFull, in);
                kb += __popc(inb);
                if (inb != kFull) break;
            }
            if (need_scan && kb != pre_base) {  // issue the next step's boundary loads now, use them a step later
                pre_base = kb;
                cpre = (kb + lane <= m) ? co[kb + lane] - ustart : INT64_MAX;
            }
            cum += step_total;
            prev_lt31 = __shfl_sync(kFull, lt, 31);
            prev_kept31 = __shfl_sync(kFull, kept, 31);
        }

        {
            const unsigned who = __ballot_sync(kFull, have_last);
            const int srcl = who ? __ffs(who) - 1 : 0;
            R_last = __shfl_sync(kFull, R_last, srcl);
            P_last = __shfl_sync(kFull, P_last, srcl);
        }
        res.total_out = P_last;
        if (COPY) {
            // carried literal bytes, then the kept runs, all clipped at the last release boundary
            if (held_kept && R_last > -h) {
                if (lane < h) dst[lane] = heldp[lane];
            }
            if (depth == 0) {
                int64_t e = R_last < n ? R_last : n;
                int64_t len = e - seg_start;
                if (len > 0) warp_copy<true>(dst + out_pos, T.p + seg_start, len, buf + total, lane);
            }
            if (st_out) {
                // what stays carried
                int64_t hp = INT64_MIN;  // start of the carried tail (logical coords)
                if (!(flags & 4u)) {
                    if (depth == 0) {
                        if (R_last < n) hp = R_last;
                    } else {
                        // last '<' after the last consumed tag, if it can still become a tag (oai_proxy.py:332-334)
                        int64_t lo = n - kHeldMax;
                        if (lo < 0) lo = 0;
                        if (last_tok_end > lo) lo = last_tok_end;
                        int64_t x = lo + lane;
                        bool islt = x < n && T(x) == '<';
                        unsigned bl = __ballot_sync(kFull, islt);
                        int64_t p = INT64_MIN;
                        if (bl) p = lo + (31 - __clz(bl));
                        else if (held_unconsumed && last_tok_end < 0 && n + h <= kHeldMax) p = -h;
                        if (p != INT64_MIN && is_tag_prefix(tg, A, p, n, false, true)) hp = p;
                    }
                }
                if (lane == 0) {
                    StreamState so;
                    so.depth = depth;
                    so.tail_len = last_tok_end >= 0 ? (int32_t)(n - last_tok_end) : -1;
                    so.held_len = 0;
                    for (int i = 0; i < kHeldMax; i++) so.held[i] = 0;  // unused bytes are defined (zero), not stale
                    if (hp != INT64_MIN) {
                        so.held_len = (uint8_t)(n - hp);
                        for (int i = 0; i < so.held_len; i++) so.held[i] = A(hp + i);
                    }
                    *st_out = so;
                }
            }
        }
        return res;
    }
};

// ------------------------------------------------------------------------------------------------------------------
// Record-driven stream walker, one warp per stream (the count kernel uses it for streams too long for one lane, see
// qb_lane.cuh for the lane-per-stream version of the same walk).
//
// Same semantics as StreamWalker, different shape of work: nothing here loops over the text.
//   1  the warp walks the unit's token records (only steps that have any): the kept runs of the unit, one run per lane,
//      and the hold intervals of the '<' that stay text at depth 0 (more than the descriptor holds, or a step whose
//      records overflowed: `overflow`, the caller falls back to StreamWalker);
//   2  the last feed boundary (c = n) decides the released total and what is carried;
//   3  descriptor (for the offsets and finish kernels) and carried state are written.
// Positions are unit-relative int32 (units of 2 GiB and more take the fallback).
// ------------------------------------------------------------------------------------------------------------------
// The largest e in [c, min(x + kHeldMax, n)] such that text[x, e) is a proper prefix of an open tag (hold-back test,
// oai_proxy.py:311-323), or c - 1 when text[x, c) is none.  The test is monotone in e.  Out of line: rare.
template <class A>
__device__ __noinline__ int hold_interval_end(const DevTags& tg, const A acc, int x, int c, int n) {
    if (!is_tag_prefix(tg, acc, x, c, true, false)) return c - 1;
    int lo = c, hi = min(x + kHeldMax, n);
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (is_tag_prefix(tg, acc, x, mid, true, false)) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

// One pass per stream (count kernel): walk, released total, descriptor for the offsets / finish kernels, carried state.
struct StreamRunWalker {
    StreamDesc* desc = nullptr;
    const DevTags& tg;
    const uint8_t* buf;
    int64_t total;
    const int64_t* co;
    int64_t m;
    int lane;
    uint32_t flags;
    TokSrc ts;
    bool overflow = false;

    __device__ StreamRunWalker(const DevTags& tg_, const uint8_t* buf_, int64_t total_, const int64_t* co_, int64_t m_, int lane_,
                               uint32_t flags_, TokSrc ts_)
        : tg(tg_), buf(buf_), total(total_), co(co_), m(m_), lane(lane_), flags(flags_), ts(ts_) {}

    __device__ StreamResult run(const StreamState* st_in, StreamState* st_out) {
        StreamState si;
        if (st_in) {
            si = *st_in;
        } else {
            si.depth = 0;
            si.tail_len = -1;
            si.held_len = 0;
            for (int i = 0; i < kHeldMax; i++) si.held[i] = 0;
        }
        StreamResult res{0};
        if (m <= 0) {  // nothing fed: state passes through
            if (st_out && lane == 0) {
                if (flags & 4u) si.held_len = 0;
                for (int i = si.held_len; i < kHeldMax; i++) si.held[i] = 0;
                si.tail_len = -1;
                *st_out = si;
            }
            return res;
        }
        const int64_t ustart = co[0];
        const int64_t n64 = co[m] - ustart;
        if (n64 >= (int64_t)1 << 31) {
            overflow = true;
            return res;
        }
        const int n = (int)n64;
        const int h = si.held_len;
        const uint8_t* heldp = st_in ? st_in->held : buf;
        const HeldAcc A{buf + ustart, heldp, h};
        const PtrAcc T{buf + ustart};

        int depth = si.depth;
        int last_tok_end = -1, seg_start = 0, held_kept = 0;
        bool held_unconsumed = false;
        // runs: lane i keeps run i = [run_s, run_e) and run_cum = released-candidate bytes before it
        int run_s = 0, run_e = 0, run_cum = 0, nruns = 0, cum = 0;

        auto first_boundary_after = [&](int64_t pos) -> int64_t {
            int64_t lo = 1, hi = m + 1;
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
                if (lane == nruns) {
                    run_s = seg_start;
                    run_e = e;
                    run_cum = cum;
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
        // A '<' at x that stays text at depth 0.  The reference holds it back at a feed boundary c when text[x, c) could
        // still grow into an open tag (oai_proxy.py:311-323): such a boundary releases only the bytes before x.  The
        // prefix test is monotone in c, so the boundaries concerned are the ones in (x, e] for one e; candidates with no
        // boundary in their interval are dropped here, the others go to the descriptor (lane i keeps candidate i).
        int nh = 0, hold_x = INT32_MAX, hold_e = INT32_MIN;
        auto hold_candidate = [&](const auto& acc, int x) {
            if (!tg.hold_enabled) return;
            const int c = (int)(co[first_boundary_after(x)] - ustart);
            if (c - x > kHeldMax) return;
            const int e = hold_interval_end(tg, acc, x, c, n);
            if (e < c) return;
            if (lane == nh) {
                hold_x = x;
                hold_e = e;
            }
            nh++;
        };

        bool has_final = false;
        {
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

            // ---- 1: token records of the unit, steps with records only ----
            if (n > 0) {
                const int64_t s_first = ustart >> 10, s_last = (ustart + n - 1) >> 10;
                for (int64_t S0 = s_first; S0 <= s_last; S0 += 32) {
                    const int64_t sj = S0 + lane;
                    int cnt = 0;
                    uint4 ra = make_uint4(0, 0, 0, 0), rb = ra;
                    if (sj <= s_last) {
                        cnt = ts.count(sj);
                        if (cnt == kRecOverflow) cnt = -1;
                        if (cnt > 0) {
                            const uint4* rp = reinterpret_cast<const uint4*>(ts.head(sj));
                            ra = rp[0];
                            if (cnt > 4) rb = rp[1];
                        }
                    }
                    if (__any_sync(kFull, cnt < 0)) {
                        overflow = true;
                        return res;
                    }
                    unsigned live = __ballot_sync(kFull, cnt > 0);
                    while (live) {
                        const int j = __ffs(live) - 1;
                        live &= live - 1;
                        const int c = __shfl_sync(kFull, cnt, j);
                        const int sb = (int)(((S0 + j) << 10) - ustart);  // unit position of the step's first byte
                        auto handle = [&](uint32_t r) {
                            const int kind = (r >> 17) & 3;
                            const int x = sb + (int)(r & 1023);
                            const int tend = x + (int)((r >> 10) & 127);
                            if (x < 0 || x >= n || x < last_tok_end) return;  // another unit's / inside a consumed tag
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
                        };
                        if (c <= kRecInline) {
#pragma unroll
                            for (int k = 0; k < kRecInline; k++) {
                                if (k >= c) break;
                                handle(__shfl_sync(kFull, rec_select(ra, rb, k), j));
                            }
                        } else {  // a dense step: one record per lane straight from memory, 32 at a time
                            for (int k0 = 0; k0 < c; k0 += 32) {
                                const uint32_t rd = k0 + lane < c ? ts.get(S0 + j, k0 + lane) : 0u;
                                for (int k = 0; k < 32 && k0 + k < c; k++) handle(__shfl_sync(kFull, rd, k));
                            }
                        }
                    }
                }
            }
            const int nclosed = nruns;       // runs ended by an open tag
            if (depth == 0) close_run(n);    // the final run, clipped later at the last release boundary
            has_final = nruns > nclosed;
            if (nruns > kDescRuns || nh > kDescHolds) {
                overflow = true;
                return res;
            }
        }

        // ---- 2: feed boundaries, one per lane ----
        int R_last = 0, P_last = 0;
        {  // only the last boundary (c = n) matters here: is it inside a hold interval?
            const unsigned who = __ballot_sync(kFull, lane < nh && hold_x < n && n <= hold_e);
            R_last = n;
            P_last = cum;
            if (who) {
                R_last = __shfl_sync(kFull, hold_x, __ffs(who) - 1);
                P_last = cum - (n - R_last);
            }
        }
        res.total_out = P_last;
        {
            if (desc) {
                StreamDesc& d = *desc;
                if (lane == 0) {
                    d.nruns = nruns;
                    d.total = P_last;
                    d.held_kept = (held_kept && R_last > -h) ? held_kept : 0;  // carried literal bytes that are released
                    d.depth = depth;
                    d.last_tok_end = last_tok_end;
                    d.rlast = R_last;
                    d.flags = (held_unconsumed ? 1 : 0) | (has_final ? 2 : 0) | (held_kept ? 4 : 0);
                    d.held_len = h;
                    d.ustart = ustart;
                }
                if (held_kept && R_last > -h && lane < h) d.held[lane] = heldp[lane];
                if (lane < kDescRuns) {  // unclipped; the copy clips the open-ended final run at rlast
                    d.run_s[lane] = lane < nruns ? run_s : 0;
                    d.run_e[lane] = lane < nruns ? run_e : 0;
                }
                if (lane < kDescHolds) {
                    d.hold_x[lane] = lane < nh ? hold_x : INT32_MAX;
                    d.hold_e[lane] = lane < nh ? hold_e : INT32_MIN;
                }
            }
        }

        // ---- 3: carried state (the kept runs and the carried literal bytes are written by the finish kernel) ----
        if (st_out) {
            int hp = INT32_MIN;  // start of the carried tail (logical coords)
            if (!(flags & 4u)) {
                if (depth == 0) {
                    if (R_last < n) hp = R_last;
                } else {
                    int lo = n - kHeldMax;
                    if (lo < 0) lo = 0;
                    if (last_tok_end > lo) lo = last_tok_end;
                    const int x = lo + lane;
                    const bool islt = x < n && T(x) == '<';
                    const unsigned bl = __ballot_sync(kFull, islt);
                    int p = INT32_MIN;
                    if (bl) p = lo + (31 - __clz(bl));
                    else if (held_unconsumed && last_tok_end < 0 && n + h <= kHeldMax) p = -h;
                    if (p != INT32_MIN && is_tag_prefix(tg, A, p, n, false, true)) hp = p;
                }
            }
            __syncwarp();  // state_in may alias state_out: every lane is done reading the carried bytes
            if (lane == 0) {
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
        }
        return res;
    }
};

// ------------------------------------------------------------------------------------------------------------------
// Record-driven strip walker ("run list"): the fast path of strip_thinking_tags.
//
// Same semantics as StripWalker.  The warp reads the unit's token records once (one step per lane), records the last
// closer of every class (pass A, lanes in parallel), walks the tags in position order (pass B) and keeps the kept runs
// of the response one per lane.  `.strip()` then only has to look at the two ends of that list.  More than 32 runs, a
// step whose records overflowed, or a response of 2 GiB and more set `overflow` (the caller falls back to StripWalker).
// ------------------------------------------------------------------------------------------------------------------
template <bool COPY, bool BATCH = false>  // BATCH: more than 32 kept runs are settled 32 at a time (out-of-line callers only)
struct StripRunWalker {
    const DevTags& tg;
    const uint8_t* buf;
    int64_t total, ustart, n64;
    unsigned long long* lastclose;  // per-warp shared array [kMaxCls]
    int lane;
    TokSrc ts;
    bool overflow = false;
    bool many = false;  // overflow because of more than 32 kept runs (and nothing else)
    bool wide = false;
    int64_t kept_len = 0;
    uint8_t* dst = nullptr;
    UnitDesc* desc = nullptr;  // count pass: where the trimmed runs go (copied later by strip_copy_kernel)

    __device__ StripRunWalker(const DevTags& tg_, const uint8_t* buf_, int64_t total_, int64_t ustart_, int64_t n_,
                              unsigned long long* lc, int lane_, TokSrc ts_)
        : tg(tg_), buf(buf_), total(total_), ustart(ustart_), n64(n_), lastclose(lc), lane(lane_), ts(ts_) {}

    __device__ void run() {
        if (n64 <= 0) return;
        if (n64 >= (int64_t)1 << 31) {
            overflow = true;
            return;
        }
        const int n = (int)n64;
        const PtrAcc T{buf + ustart};
        const int64_t s_first = ustart >> 10, s_last = (ustart + n - 1) >> 10;
        const bool one_group = s_last - s_first < 32;  // the usual case: all records stay in registers for both passes
        for (int c = lane; c < tg.ncls; c += 32) lastclose[c] = 0;
        __syncwarp();
        int run_s = 0, run_e = 0, nruns = 0;  // lane i keeps run i
        int state = -1, seg_start = 0;
        bool dead = false;  // a batch could not be settled (overflow is set): ignore the rest
        int cnt = 0;
        uint4 ra = make_uint4(0, 0, 0, 0), rb = ra;
        for (int pass = 0; pass < 2; pass++) {
            for (int64_t S0 = s_first; S0 <= s_last; S0 += 32) {
                const int64_t sj = S0 + lane;
                if (pass == 0 || !one_group) {
                    cnt = 0;
                    if (sj <= s_last) {
                        cnt = ts.count(sj);
                        if (cnt == kRecOverflow) cnt = -1;
                        if (cnt > 0) {
                            const uint4* rp = reinterpret_cast<const uint4*>(ts.head(sj));
                            ra = rp[0];
                            if (cnt > 4) rb = rp[1];
                        }
                    }
                    if (__any_sync(kFull, cnt < 0)) {
                        overflow = true;
                        return;
                    }
                }
                const int sb0 = (int)((S0 << 10) - ustart);
                if (pass == 0) {  // pass A: last closer per class, every lane over its own step
                    for (int k = 0; k < cnt; k++) {
                        const uint32_t r = k < kRecInline ? rec_select(ra, rb, k) : ts.tail(sj)[k];
                        if (((r >> 17) & 3) != 1) continue;
                        const int x = sb0 + lane * 1024 + (int)(r & 1023);
                        const int tend = x + (int)((r >> 10) & 127);
                        if (x >= 0 && tend <= n) atomicMax(&lastclose[(r >> 19) & 127], (unsigned long long)(x + 1));
                    }
                } else {  // pass B: tags in position order
                    unsigned live = __ballot_sync(kFull, cnt > 0);
                    while (live) {
                        const int j = __ffs(live) - 1;
                        live &= live - 1;
                        const int c = __shfl_sync(kFull, cnt, j);
                        auto handle = [&](uint32_t r) {
                            const int kind = (r >> 17) & 3;
                            if (kind == 2 || dead) return;
                            const int x = sb0 + j * 1024 + (int)(r & 1023);
                            const int tend = x + (int)((r >> 10) & 127);
                            if (x < 0 || tend > n) return;
                            const int cls = (r >> 19) & 127;
                            if (state < 0) {
                                if (kind == 0 && lastclose[cls] > (unsigned long long)x) {  // live open tag
                                    if (x > seg_start) {
                                        if (lane == nruns) { run_s = seg_start; run_e = x; }
                                        nruns++;
                                        if (BATCH && nruns == 32) {  // the lanes are full: settle this batch, start the next
                                            if (!flush_batch(run_s, run_e, nruns, false, n, T)) dead = true;
                                            nruns = 0;
                                        }
                                    }
                                    state = cls;
                                }
                            } else if (kind == 1 && cls == state) {
                                seg_start = tend;
                                state = -1;
                            }
                        };
                        if (c <= kRecInline) {
#pragma unroll
                            for (int k = 0; k < kRecInline; k++) {
                                if (k >= c) break;
                                handle(__shfl_sync(kFull, rec_select(ra, rb, k), j));
                            }
                        } else {  // a dense step: one record per lane straight from memory, 32 at a time.
                            // Every lane decodes its own record and looks up liveness; the automaton then only visits the
                            // records that change its state (ballot + find-first-set), not every record.
                            for (int k0 = 0; k0 < c && !dead; k0 += 32) {
                                const uint32_t r = k0 + lane < c ? ts.get(S0 + j, k0 + lane) : (2u << 17);
                                const int kind = (r >> 17) & 3;
                                const int x = sb0 + j * 1024 + (int)(r & 1023);
                                const int tend = x + (int)((r >> 10) & 127);
                                const int cls = (r >> 19) & 127;
                                const bool valid = kind != 2 && x >= 0 && tend <= n;
                                const bool live = valid && kind == 0 && lastclose[cls] > (unsigned long long)x;
                                const unsigned open_m = __ballot_sync(kFull, live);
                                int pos = 0;
                                while (pos < 32) {
                                    const unsigned from = 0xffffffffu << pos;
                                    if (state < 0) {
                                        const unsigned mm = open_m & from;
                                        if (!mm) break;
                                        const int i = __ffs(mm) - 1;
                                        const int xi = __shfl_sync(kFull, x, i);
                                        if (xi > seg_start) {
                                            if (lane == nruns) { run_s = seg_start; run_e = xi; }
                                            nruns++;
                                            if (BATCH && nruns == 32) {
                                                if (!flush_batch(run_s, run_e, nruns, false, n, T)) dead = true;
                                                nruns = 0;
                                            }
                                        }
                                        state = __shfl_sync(kFull, cls, i);
                                        pos = i + 1;
                                    } else {
                                        const unsigned mm = __ballot_sync(kFull, valid && kind == 1 && cls == state) & from;
                                        if (!mm) break;
                                        const int i = __ffs(mm) - 1;
                                        seg_start = __shfl_sync(kFull, tend, i);
                                        state = -1;
                                        pos = i + 1;
                                    }
                                    if (dead) break;
                                }
                            }
                        }
                    }
                }
            }
            __syncwarp();
        }
        if (dead) return;
        if (n > seg_start) {
            if (lane == nruns) { run_s = seg_start; run_e = n; }
            nruns++;
            if (BATCH && nruns == 32) {
                if (!flush_batch(run_s, run_e, nruns, false, n, T)) return;
                nruns = 0;
            }
        }
        if (!BATCH && nruns > 32) {  // more runs than lanes: the caller retries with the batching variant
            overflow = true;
            many = true;
            return;
        }
        flush_batch(run_s, run_e, nruns, true, n, T);
    }

    // ---- batches of up to 32 runs (one per lane) --------------------------------------------------------------------
    // A response with more than 32 kept runs is handled in batches: the leading `.strip()` cut is made on the first
    // batch that has text, every batch remembers where its text ends (so a white-space-only tail of later batches is
    // cut correctly), and the count pass's kept_len bounds what the copy pass writes.  A whole non-final batch of white
    // space before any text sets `overflow` (byte-rescanning walker).
    int64_t flushed = 0;     // kept bytes of the batches before this one
    int64_t text_end = 0;    // kept length up to the last non-white byte seen in the batches before this one
    bool lead_done = false;
    int batches = 0;
    int64_t klimit = INT64_MAX;  // copy pass: kept_len from the count pass

    __device__ __forceinline__ bool flush_batch(int& run_s, int& run_e, int nruns, bool final, int n, const PtrAcc& T) {
        int lead = 0;
        if (!lead_done) {  // .strip(), front: drop white space from the first runs that keep anything
            lead = nruns;
            for (int i = 0; i < nruns; i++) {
                const int rs = __shfl_sync(kFull, run_s, i), re = __shfl_sync(kFull, run_e, i);
                int f = -1;
                for (int base = rs; base < re; base += 32) {
                    const int x = base + lane;
                    const bool nw = x < re && !is_white_byte(T, x, n);
                    const unsigned b = __ballot_sync(kFull, nw);
                    if (b) { f = base + __ffs(b) - 1; break; }
                }
                if (f >= 0) {
                    if (lane == i) run_s = f;
                    lead = i;
                    lead_done = true;
                    break;
                }
                if (lane == i) run_e = run_s;  // all white
            }
            if (!lead_done && !final && nruns > 0) {
                overflow = true;
                return false;
            }
        }
        // .strip(), back: where does the text of this batch end?  The final batch is cut there; for an earlier batch the
        // answer is remembered in case every later batch turns out to be white space only.
        bool found = false;
        int end_run = -1, end_pos = 0;
        if (lead_done) {
            for (int i = nruns - 1; i >= lead; i--) {
                const int rs = __shfl_sync(kFull, run_s, i), re = __shfl_sync(kFull, run_e, i);
                int l = -1;
                for (int top = re; top > rs; top -= 32) {
                    const int x = top - 32 + lane;
                    const bool nw = x >= rs && !is_white_byte(T, x, n);
                    const unsigned b = __ballot_sync(kFull, nw);
                    if (b) { l = top - 32 + (31 - __clz(b)); break; }
                }
                if (l >= 0) {
                    found = true;
                    end_run = i;
                    end_pos = l + 1;
                    break;
                }
            }
        }
        if (final) {
            if (found) {
                if (lane == end_run) run_e = end_pos;
                if (lane > end_run) run_e = run_s;
            } else {
                run_e = run_s;  // nothing but white space in this batch
            }
        }
        const int mylen = lane < nruns ? run_e - run_s : 0;
        int incl = mylen;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += t;
        }
        const int batch_len = __shfl_sync(kFull, incl, 31);
        if (found && !final) {  // kept length if the text ended inside this batch
            const int before = __shfl_sync(kFull, incl - mylen, end_run), rs_e = __shfl_sync(kFull, run_s, end_run);
            text_end = flushed + before + (end_pos - rs_e);
        }
        if (COPY) {
            const int my_off = incl - mylen;
            if (batch_len <= 64 * nruns) {  // short runs: one run per lane, byte by byte
                int64_t len = mylen;
                if (flushed + my_off + len > klimit) len = klimit - flushed - my_off;
                for (int64_t k = 0; k < len; k++) dst[flushed + my_off + k] = T(run_s + k);
            } else {
                for (int i = 0; i < nruns; i++) {
                    const int rs = __shfl_sync(kFull, run_s, i), ro = __shfl_sync(kFull, my_off, i);
                    int64_t len = __shfl_sync(kFull, mylen, i);
                    if (flushed + ro + len > klimit) len = klimit - flushed - ro;  // trailing white space of the whole text
                    if (len > 0) warp_copy<true>(dst + flushed + ro, T.p + rs, len, buf + total, lane);
                }
            }
        }
        if (final) {
            wide = batches > 0 || nruns > kDescRuns;  // does not fit a descriptor: the copy kernel walks this unit again
            kept_len = found ? flushed + batch_len : text_end;  // all white at the end: the text ended in an earlier batch
            if (desc) {
                UnitDesc& d = *desc;
                if (lane == 0) {
                    d.nruns = wide ? -2 : nruns;
                    d.total = (int32_t)kept_len;
                    d.held_kept = 0;
                    d.depth = 0;
                    d.last_tok_end = -1;
                    d.rlast = n;
                    d.flags = 0;
                    d.held_len = 0;
                }
                if (lane < kDescRuns) {
                    d.run_s[lane] = lane < nruns ? run_s : 0;
                    d.run_e[lane] = lane < nruns ? run_e : 0;
                }
            }
        }
        flushed += batch_len;
        batches++;
        return true;
    }
};

}  // namespace qb
