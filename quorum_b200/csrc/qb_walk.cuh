// Warp-cooperative walkers: one warp runs one unit (a response, or one stream's data of this call).
//
// Shared structure: the unit is visited in steps of 32 words x 32 bytes (1 KiB); lane j owns word j of the step and a
// 32-bit bitmap of the '<' bytes in it.  Only '<' positions can start a tag, so everything sequential (the tag
// automaton) runs over the sparse candidate list in position order, while the dense work (finding candidates,
// counting and copying kept bytes) is lane-parallel.
#pragma once
#include "qb_common.cuh"

namespace qb {

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
            if (take > 0) warp_copy(dst + cum, T.p + s, take, buf + total, lane);
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
        int state = -1;
        int64_t seg_start = 0;
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
                const int cls = kc & 0xff;
                const bool is_close = (kc & 0x100) != 0;
                if (state < 0) {
                    if (!is_close && lastclose[cls] > (unsigned long long)x) {
                        emit(seg_start, x);
                        state = cls;
                    }
                } else if (is_close && cls == state) {
                    seg_start = tend;
                    state = -1;
                }
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
        int64_t kb = 1;            // next feed boundary to account (co[kb] is the end of chunk kb-1)
        int64_t R_last = 0;        // release boundary of the last feed
        int64_t P_last = 0;        // kept bytes before R_last
        bool have_last = false;    // this lane accounted the last boundary
        uint32_t prev_lt31 = 0, prev_kept31 = 0;
        if (COPY && lane == 0) out_chunk_off[chunk0] = out_base;

        const int64_t w_first = ustart >> 5;
        const int64_t w_last = n > 0 ? ((ustart + n - 1) >> 5) : (w_first - 1);
        const int64_t nsteps = n > 0 ? ((w_last - w_first + 32) >> 5) : 1;  // one (empty) step accounts empty chunks
        for (int64_t s = 0; s < nsteps; s++) {
            const int64_t w = w_first + s * 32 + lane;
            const int64_t wb = (w << 5) - ustart;
            const uint32_t range = (w <= w_last) ? (mask_ge(wb, 0) & ~mask_ge(wb, n)) : 0u;
            const uint32_t lt = (w <= w_last) ? (load_lt_word(buf, total, w) & range) : 0u;
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
            uint32_t kept = depth == 0 ? mask_ge(wb, skip_until) : 0u;
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
                if (recognised(is_close != 0, x, tend, ex)) {
                    if (depth == 0) {  // open tag at depth 0: the kept run ends at x
                        kept &= ~mask_ge(wb, x);
                        if (COPY) {
                            int64_t len = x - seg_start;
                            if (len > 0) warp_copy(dst + out_pos, T.p + seg_start, len, buf + total, lane);
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
                }
                act = __ballot_sync(kFull, tok != 0);
            }
            kept &= range;
            // exclusive prefix of kept counts over the words of this step
            int pc = __popc(kept), incl = pc;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int v = __shfl_up_sync(kFull, incl, d);
                if (lane >= d) incl += v;
            }
            const int excl = incl - pc;
            const int step_total = __shfl_sync(kFull, incl, 31);
            const int64_t step_end = wb - (int64_t)lane * 32 + 1024 < n ? wb - (int64_t)lane * 32 + 1024 : n;
            const int64_t step_wb0 = wb - (int64_t)lane * 32;

            // ---- feed boundaries whose last byte lies in this step (c == 0: before any byte) ----
            for (;;) {
                const int64_t k = kb + lane;
                const int64_t c = k <= m ? co[k] - ustart : INT64_MAX;
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
                if (len > 0) warp_copy(dst + out_pos, T.p + seg_start, len, buf + total, lane);
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

}  // namespace qb
