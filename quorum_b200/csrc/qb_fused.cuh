// Fused single-pass kernels (the default path): one persistent kernel per call does what the token path does in four to
// seven launches, tile by tile, so that the text comes from HBM once.
//
// A tile is the run of consecutive units (requests with their responses / streams with their chunks) whose first byte
// lies in one window of 2^log2T bytes of the flat input, so units are never split and a tile's bytes are contiguous.  A
// CTA takes tiles by ticket (tiles start in index order, which the look-back needs for forward progress) and runs, per
// tile:
//   1  tokenise the tile's bytes (tokenize_pair, all warps): records into shared memory (global scratch when the tile
//      spans more steps than the shared arrays hold)
//   2  walk: lane per unit over the records (strip_walk_lane / stream_walk_lane; warp-per-unit and byte-rescanning
//      fallbacks exactly as in the token path) -> kept length, descriptor, carried state
//   3  block scan of the lengths (+ separators), tile aggregate, decoupled look-back over tiles -> output base
//   4  (streams) per-chunk output offsets, thread per feed boundary
//   5  copy: warp per unit, kept runs to their final place (gather_unit)
// Nothing is staged in shared memory: step 1 pulls the tile through L2, step 5 re-reads the kept runs a few microseconds
// later, while the tile is still L2-resident (tiles in flight: 4 CTAs x 148 SMs x 32-64 KiB, well inside the 126 MB L2).
// The serial chain of one tile (load -> records -> walk -> look-back -> copy) is hidden by the other CTAs of the SM,
// which are in other phases; the descriptors, lengths and records a tile needs live in shared memory, so the chain has
// no DRAM round trip besides the first touch of the text.
#pragma once
#include "qb_lane.cuh"
#include "qb_scan.cuh"
#include "qb_tok.cuh"

namespace qb {

constexpr int kFusedSteps = 80;      // steps of records a tile keeps in shared memory (a 64 KiB window + a 16 KiB overhang)
constexpr int kFusedUnits = 64;      // units whose descriptor / lengths a tile keeps in shared memory
constexpr int kFusedMinLog2T = 15;   // smallest window: 32 KiB

// Tile tables: first[t] / last[t] = 1 + index of the first / 1 + index of the last unit that starts in window t (0: none).
// Zeroed by the host (cudaMemsetAsync) together with the look-back words and the ticket.
struct TileTables {
    unsigned long long* state;
    unsigned int* ticket;
    int64_t* first;
    int64_t* last;
    int64_t max_tiles;
    int log2T;
};

// Block-wide exclusive scan of one int64 per thread (kThreads threads); returns the exclusive prefix, total in `all`.
// Two barriers; `s_w` is a shared array of kWarpsPerBlock entries that may be reused after the call returns.
__device__ __forceinline__ int64_t block_excl_scan(int64_t v, int64_t* s_w, int lane, int wib, int64_t& all) {
    int64_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int64_t t = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) s_w[wib] = incl;
    __syncthreads();
    int64_t before = 0;
    all = 0;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; w++) {
        const int64_t t = s_w[w];
        if (w < wib) before += t;
        all += t;
    }
    __syncthreads();
    return before + incl - v;
}

// Tokenise the steps that cover bytes [tb, te) of the flat buffer, all warps of the CTA.  Returns where the records are:
// shared memory when the range spans at most kFusedSteps + 2 steps, the call's global record arrays otherwise.
__device__ __forceinline__ TokSrc tokenize_tile(const DevTags& tg, const uint8_t* bytes, int64_t total, int64_t tb, int64_t te,
                                                uint8_t* s_cnt, uint32_t* s_rec, uint8_t* g_cnt, uint32_t* g_rec, int64_t g_max_steps) {
    TokSrc ts;
    if (te <= tb) return ts;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t pr0 = tb >> 11, pr1 = (te - 1) >> 11;
    const bool in_smem = (pr1 - pr0 + 1) * 2 <= kFusedSteps + 2;
    int64_t nsteps = (total + 1023) >> 10;
    uint8_t* cnt = g_cnt;
    uint32_t* rec = g_rec;
    if (in_smem) {  // index by absolute step: shift the shared arrays so that the tile's first step lands on entry 0
        cnt = s_cnt - (pr0 << 1);
        rec = s_rec - (pr0 << 1) * kRecPerStep;
    } else if (nsteps > g_max_steps) {
        nsteps = g_max_steps;
    }
    // the tile's lines start moving towards L2 before the first load waits on one
    for (int64_t o = (tb & ~(int64_t)127) + (int64_t)threadIdx.x * 128; o < te; o += (int64_t)kThreads * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(bytes + o));
    for (int64_t pr = pr0 + wib; pr <= pr1; pr += kWarpsPerBlock) tokenize_pair(tg, bytes, total, nsteps, pr, lane, cnt, rec);
    ts.cnt = cnt;
    ts.rec = rec;
    return ts;
}

// ---------------------------------------------------------------------------------------------------------------
// strip_thinking_tags + concatenate (oai_proxy.py:120-139, 1194-1199 + 1286, 759-764 + 841), fused
// ---------------------------------------------------------------------------------------------------------------
struct StripFusedParams {
    const uint8_t* in;
    const int64_t* in_off;
    int64_t n_resp;
    const int64_t* req_off;  // may be null: one response per request
    int64_t n_req;
    const uint8_t* sep;
    int32_t sep_len;
    int32_t hide, stream_join, chain;
    uint8_t* out;
    int64_t* out_off;
    int64_t* resp_out_len;  // may be null
    int64_t* klen;          // scratch, n_resp entries each
    int64_t* first;
    int64_t* dst;
    UnitDesc* desc;
    uint8_t* g_cnt;         // global record arrays (tiles too long for shared memory)
    uint32_t* g_rec;
    int64_t g_max_steps;
    TileTables tt;
    int32_t lane_steps;
};

// thread per request: the window its first byte lies in
__global__ void __launch_bounds__(256) strip_tile_map_kernel(const int64_t* __restrict__ in_off, const int64_t* __restrict__ req_off,
                                                             int64_t n_req, TileTables tt) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_req) return;
    const int64_t begin = in_off[req_off ? req_off[0] : 0];
    auto win = [&](int64_t qq) { return (in_off[req_off ? req_off[qq] : qq] - begin) >> tt.log2T; };
    const int64_t id = win(q);
    if (id >= tt.max_tiles) return;  // only when out_capacity is below the input size (contract violated): stay inside the tables
    if (q == 0 || win(q - 1) != id) tt.first[id] = q + 1;
    if (q == n_req - 1 || win(q + 1) != id) tt.last[id] = q + 1;
}

__global__ void __launch_bounds__(kThreads, 4)
strip_fused_kernel(const __grid_constant__ DevTags tg, const __grid_constant__ StripFusedParams P) {
    __shared__ unsigned long long lastclose[kWarpsPerBlock][kMaxCls];
    __shared__ __align__(16) uint32_t s_rec[(kFusedSteps + 2) * kRecPerStep];
    __shared__ __align__(16) UnitDesc s_desc[kFusedUnits];
    __shared__ int64_t s_klen[kFusedUnits], s_first[kFusedUnits], s_dst[kFusedUnits];
    __shared__ int64_t s_w[kWarpsPerBlock];
    __shared__ int64_t s_ticket, s_base;
    __shared__ uint8_t s_cnt[kFusedSteps + 2 + 6];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t* __restrict__ in_off = P.in_off;
    const int64_t* __restrict__ req_off = P.req_off;
    const int64_t begin = in_off[0], total = in_off[P.n_resp];
    int64_t ntiles = ((total - begin) >> P.tt.log2T) + 1;
    if (ntiles > P.tt.max_tiles) ntiles = P.tt.max_tiles;
    const int64_t init = P.chain ? P.out_off[0] : 0;  // chained call: this slice's output starts where the previous one ended
    const int hide = P.hide, sj = P.stream_join;
    const int32_t sep_len = P.sep_len;
    for (;;) {
        if (threadIdx.x == 0) s_ticket = (int64_t)atomicAdd(P.tt.ticket, 1u);
        __syncthreads();
        const int64_t t = s_ticket;
        if (t >= ntiles) break;
        const int64_t q0 = P.tt.first[t] - 1;
        if (q0 < 0) {  // no request starts in this window
            if (wib == 0) lookback_exclusive(P.tt.state, t, 0, lane);
            __syncthreads();
            continue;
        }
        const int64_t q1 = P.tt.last[t];
        const int64_t r0 = req_off ? req_off[q0] : q0, r1 = req_off ? req_off[q1] : q1;
        const int64_t tb = in_off[r0], te = in_off[r1];
        // small tile: descriptors, lengths and destinations stay in shared memory (indexed by response, shifted)
        const bool small = r1 - r0 <= kFusedUnits;
        UnitDesc* desc = small ? s_desc - r0 : P.desc;
        int64_t* klen = small ? s_klen - r0 : P.klen;
        int64_t* first = small ? s_first - r0 : P.first;
        int64_t* dst = small ? s_dst - r0 : P.dst;

        // ---- 1: records ----
        TokSrc ts;
        if (hide) ts = tokenize_tile(tg, P.in, total, tb, te, s_cnt, s_rec, P.g_cnt, P.g_rec, P.g_max_steps);
        __syncthreads();

        // ---- 2: walk (lane per response, then warp per response for the ones that need it) ----
        if (hide) {
            for (int64_t gb = r0; gb < r1; gb += kThreads) {
                const int64_t g0 = gb + (wib << 5);
                if (g0 >= r1) continue;
                const int64_t r = g0 + lane;
                const bool have = r < r1;
                int64_t a = 0, b = 0, kl = 0, fp = 0;
                bool done = false;
                if (have) {
                    a = in_off[r];
                    b = in_off[r + 1];
                    if (b <= a) {
                        desc[r].nruns = 0;
                        done = true;
                    } else if (((b - 1) >> 10) - (a >> 10) < P.lane_steps) {
                        done = strip_walk_lane(tg, P.in, a, b - a, ts, desc + r, kl);
                    }
                }
                __syncwarp();
                unsigned todo = __ballot_sync(kFull, have && !done);
                while (todo) {
                    const int j = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const int64_t rj = g0 + j;
                    const int64_t aj = __shfl_sync(kFull, a, j), bj = __shfl_sync(kFull, b, j);
                    StripResult res;
                    StripRunWalker<false> w(tg, P.in, total, aj, bj - aj, lastclose[wib], lane, ts);
                    w.desc = desc + rj;
                    w.run();
                    res.kept_len = w.kept_len;
                    res.first_pos = 0;
                    bool slow = w.overflow;
                    if (w.many) slow = !strip_count_wide(tg, P.in, total, aj, bj - aj, lastclose[wib], lane, ts, desc + rj, &res.kept_len);
                    if (slow) {
                        res = strip_count_slow(tg, P.in, total, aj, bj - aj, lastclose[wib], lane);
                        if (lane == 0) desc[rj].nruns = -1;  // redone from the bytes by the copy step
                    }
                    if (lane == j) {
                        kl = res.kept_len;
                        fp = res.first_pos;
                    }
                    __syncwarp();
                }
                if (have) {
                    klen[r] = kl;
                    first[r] = fp;
                }
            }
            __syncthreads();
        }

        // ---- 3: request lengths (separators included), scan, look-back ----
        auto request_len = [&](int64_t q, int64_t base2, bool write) -> int64_t {
            // length of request q; with write: dst[r] = 2 * (absolute offset of response r's text) | (separator before it)
            const int64_t ra = req_off ? req_off[q] : q, rb = req_off ? req_off[q + 1] : q + 1;
            int64_t o = 0;
            bool fst = true;
            for (int64_t r = ra; r < rb; r++) {
                const int64_t len = in_off[r + 1] - in_off[r];
                if (sj && len == 0) {  // oai_proxy.py:763 `if text`
                    if (write) dst[r] = -1;
                    continue;
                }
                if (!fst) o += sep_len;
                if (write) dst[r] = ((base2 + o) << 1) | (fst ? 0 : 1);
                fst = false;
                o += hide ? klen[r] : len;
            }
            return o;
        };
        int64_t run = 0, rel0 = 0;
        for (int64_t qb = q0; qb < q1; qb += kThreads) {
            const int64_t q = qb + threadIdx.x;
            const int64_t len = q < q1 ? request_len(q, 0, false) : 0;
            int64_t all;
            const int64_t excl = run + block_excl_scan(len, s_w, lane, wib, all);
            if (qb == q0) rel0 = excl;
            else if (q < q1) P.out_off[q] = excl;  // tile-relative until the base is known
            run += all;
        }
        if (wib == 0) {
            const int64_t excl = lookback_exclusive(P.tt.state, t, run, lane);
            if (lane == 0) s_base = excl;
        }
        __syncthreads();
        const int64_t base = init + s_base;
        for (int64_t qb = q0; qb < q1; qb += kThreads) {
            const int64_t q = qb + threadIdx.x;
            if (q < q1) {
                const int64_t o = base + (qb == q0 ? rel0 : P.out_off[q]);
                P.out_off[q] = o;
                request_len(q, o, true);
            }
        }
        if (q1 == P.n_req && threadIdx.x == 0) P.out_off[P.n_req] = base + run;
        __syncthreads();

        // ---- 5: copy (warp per response) ----
        for (int64_t r = r0 + wib; r < r1; r += kWarpsPerBlock) {
            const int64_t a = in_off[r], b = in_off[r + 1];
            const int64_t de = dst[r];
            if (de < 0) {  // excluded from the join
                if (P.resp_out_len && lane == 0) P.resp_out_len[r] = 0;
                continue;
            }
            const int64_t d = de >> 1;
            if (de & 1)  // the separator between this response and the previous included one of its request
                for (int i = lane; i < sep_len; i += 32) P.out[d - sep_len + i] = P.sep[i];
            const int64_t kl = hide ? klen[r] : (b - a);
            if (P.resp_out_len && lane == 0) P.resp_out_len[r] = kl;
            if (!hide) {
                warp_copy<true>(P.out + d, P.in + a, b - a, P.in + total, lane);
            } else if (kl > 0) {
                const int nr = desc[r].nruns;
                if (nr >= 0) {
                    gather_unit<0>(desc + r, P.in + a, P.out + d, P.in + total, lane);
                } else if (nr == -2) {  // 9..32 kept runs: walk the records again and copy from the lanes
                    strip_copy_wide(tg, P.in, total, a, b - a, lastclose[wib], lane, ts, kl, P.out + d);
                } else {
                    strip_copy_slow(tg, P.in, total, a, b - a, lastclose[wib], lane, first[r], kl, P.out + d);
                }
            }
            __syncwarp();
        }
        __syncthreads();  // the next tile overwrites the shared records, descriptors and the ticket
    }
}

// ---------------------------------------------------------------------------------------------------------------
// ThinkingTagFilter.feed / flush over many streams (oai_proxy.py:278-371), fused
// ---------------------------------------------------------------------------------------------------------------
struct StreamFusedParams {
    const uint8_t* bytes;
    const int64_t* chunk_off;
    int64_t n_chunks;
    const int64_t* stream_off;
    int64_t n_streams;
    const StreamState* st_in;  // may be null
    StreamState* st_out;       // may be null
    uint32_t flags;
    int32_t chain;
    uint8_t* out;
    int64_t* out_chunk_off;
    int64_t* bases;            // scratch, n_streams entries (tile-relative output offset of every stream)
    StreamDesc* desc;          // scratch, n_streams entries
    uint8_t* g_cnt;
    uint32_t* g_rec;
    int64_t g_max_steps;
    TileTables tt;
    int32_t lane_steps;
};

// thread per stream: the window its first byte lies in
__global__ void __launch_bounds__(256) stream_tile_map_kernel(const int64_t* __restrict__ chunk_off, const int64_t* __restrict__ stream_off,
                                                              int64_t n_streams, TileTables tt) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    const int64_t begin = chunk_off[stream_off[0]];
    auto win = [&](int64_t ss) { return (chunk_off[stream_off[ss]] - begin) >> tt.log2T; };
    const int64_t id = win(s);
    if (id >= tt.max_tiles) return;  // only when out_capacity is below the input size (contract violated)
    if (s == 0 || win(s - 1) != id) tt.first[id] = s + 1;
    if (s == n_streams - 1 || win(s + 1) != id) tt.last[id] = s + 1;
}

// bytes the stream of descriptor d has released by the feed boundary at unit position c (see stream_offsets_kernel)
__device__ __forceinline__ int released_before(const StreamDesc* d, int c) {
    int P = (d->flags & 4) ? d->held_len : 0;
#pragma unroll
    for (int i = 0; i < kDescRuns; i++) {  // unused run slots are [0, 0): they add nothing
        const int s = d->run_s[i];
        P += min(max(c - s, 0), d->run_e[i] - s);
    }
    // a boundary inside a hold interval releases only the bytes before the held '<' (everything in between is text)
    if (d->hold_x[0] != INT32_MAX) {
#pragma unroll
        for (int i = 0; i < kDescHolds; i++) {
            const int hx = d->hold_x[i];
            if (c > hx && c <= d->hold_e[i]) P -= c - hx;
        }
    }
    return P;
}

__global__ void __launch_bounds__(kThreads, 4)
stream_fused_kernel(const __grid_constant__ DevTags tg, const __grid_constant__ StreamFusedParams P) {
    __shared__ __align__(16) uint32_t s_rec[(kFusedSteps + 2) * kRecPerStep];
    __shared__ __align__(16) StreamDesc s_desc[kFusedUnits];
    __shared__ int64_t s_bases[kFusedUnits];
    __shared__ int64_t s_w[kWarpsPerBlock];
    __shared__ int64_t s_ticket, s_base;
    __shared__ uint8_t s_cnt[kFusedSteps + 2 + 6];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t* __restrict__ chunk_off = P.chunk_off;
    const int64_t* __restrict__ stream_off = P.stream_off;
    const int64_t total = chunk_off[P.n_chunks];
    const int64_t cb = stream_off[0];
    const int64_t begin = chunk_off[cb], end = chunk_off[stream_off[P.n_streams]];
    int64_t ntiles = ((end - begin) >> P.tt.log2T) + 1;
    if (ntiles > P.tt.max_tiles) ntiles = P.tt.max_tiles;
    const int64_t init = P.chain ? P.out_chunk_off[cb] : 0;  // chained call: continue where the previous slice ended
    const uint32_t flags = P.flags;
    for (;;) {
        if (threadIdx.x == 0) s_ticket = (int64_t)atomicAdd(P.tt.ticket, 1u);
        __syncthreads();
        const int64_t t = s_ticket;
        if (t >= ntiles) break;
        const int64_t s0 = P.tt.first[t] - 1;
        if (s0 < 0) {  // no stream starts in this window
            if (wib == 0) lookback_exclusive(P.tt.state, t, 0, lane);
            __syncthreads();
            continue;
        }
        const int64_t s1 = P.tt.last[t];
        const int64_t k0 = stream_off[s0], k1 = stream_off[s1];  // the tile's chunks
        const int64_t tb = chunk_off[k0], te = chunk_off[k1];
        const bool small = s1 - s0 <= kFusedUnits;
        StreamDesc* desc = small ? s_desc - s0 : P.desc;
        int64_t* bases = small ? s_bases - s0 : P.bases;

        // ---- 1: records ----
        const TokSrc ts = tokenize_tile(tg, P.bytes, total, tb, te, s_cnt, s_rec, P.g_cnt, P.g_rec, P.g_max_steps);
        __syncthreads();

        // ---- 2 + 3: walk (lane per stream, warp per stream for the ones that need it), scan of the released totals ----
        int64_t run = 0;
        for (int64_t gb = s0; gb < s1; gb += kThreads) {
            const int64_t g0 = gb + (wib << 5);
            const int64_t s = g0 + lane;
            const bool have = s < s1;
            int64_t c0 = 0, c1 = 0, tot = 0;
            bool done = false;
            if (have) {
                c0 = stream_off[s];
                c1 = stream_off[s + 1];
                bool lane_ok = true;
                if (c1 > c0) {
                    const int64_t a = chunk_off[c0], b = chunk_off[c1];
                    lane_ok = b <= a || (ts.cnt && ((b - 1) >> 10) - (a >> 10) < P.lane_steps);
                }
                if (lane_ok)
                    done = stream_walk_lane(tg, P.bytes, chunk_off + c0, c1 - c0, flags, ts, P.st_in ? P.st_in + s : nullptr,
                                            P.st_out ? P.st_out + s : nullptr, desc + s, tot);
            }
            __syncwarp();
            unsigned todo = __ballot_sync(kFull, have && !done);
            while (todo) {  // one warp per stream
                const int j = __ffs(todo) - 1;
                todo &= todo - 1;
                const int64_t sj = g0 + j;
                const int64_t d0 = __shfl_sync(kFull, c0, j), d1 = __shfl_sync(kFull, c1, j);
                StreamResult r;
                bool slow = ts.cnt == nullptr;
                if (!slow) {
                    StreamRunWalker w(tg, P.bytes, total, chunk_off + d0, d1 - d0, lane, flags, ts);
                    w.desc = desc + sj;
                    if (d1 <= d0 && lane == 0) desc[sj].nruns = desc[sj].held_kept = 0;  // nothing fed: nothing to copy
                    r = w.run(P.st_in ? P.st_in + sj : nullptr, P.st_out ? P.st_out + sj : nullptr);
                    slow = w.overflow;  // does not fit the descriptor: redo this unit from its bytes
                }
                if (slow) {
                    r = stream_count_slow(tg, P.bytes, total, chunk_off + d0, d1 - d0, lane, flags, P.st_in ? P.st_in + sj : nullptr);
                    if (lane == 0) desc[sj].nruns = -1;
                }
                if (lane == j) tot = r.total_out;
                __syncwarp();
            }
            int64_t all;
            const int64_t excl = run + block_excl_scan(tot, s_w, lane, wib, all);
            if (have) bases[s] = excl;  // tile-relative
            run += all;
        }
        if (wib == 0) {
            const int64_t excl = lookback_exclusive(P.tt.state, t, run, lane);
            if (lane == 0) s_base = excl;
        }
        __syncthreads();
        const int64_t base = init + s_base;

        // ---- 4: per-chunk output offsets, thread per feed boundary ----
        if (s0 == 0 && threadIdx.x == 0) P.out_chunk_off[cb] = base;
        for (int64_t k = k0 + threadIdx.x; k < k1; k += kThreads) {
            int64_t lo = s0, hi = s1 - 1;  // the last stream that starts at or before chunk k owns it
            while (lo < hi) {
                const int64_t mid = (lo + hi + 1) >> 1;
                if (stream_off[mid] <= k) lo = mid;
                else hi = mid - 1;
            }
            const StreamDesc* d = desc + lo;
            if (d->nruns < 0) continue;  // redone by the byte-rescanning walker, which writes its own offsets
            const int c = (int)(chunk_off[k + 1] - d->ustart);
            P.out_chunk_off[k + 1] = base + bases[lo] + released_before(d, c);
        }

        // ---- 5: finish (warp per stream): released carried bytes and kept runs to their final place ----
        for (int64_t s = s0 + wib; s < s1; s += kWarpsPerBlock) {
            const StreamDesc* d = desc + s;
            const int nruns = d->nruns;
            const int64_t ob = base + bases[s];
            if (nruns < 0) {
                const int64_t c0 = stream_off[s], c1 = stream_off[s + 1];
                StreamState si;  // state_in may alias state_out: the input state is read before the output state is written
                if (P.st_in) si = P.st_in[s];
                stream_copy_slow(tg, P.bytes, total, chunk_off + c0, c1 - c0, lane, flags, P.st_in ? &si : nullptr,
                                 P.st_out ? P.st_out + s : nullptr, ob, P.out, P.out_chunk_off, c0);
            } else {
                const int hk = d->held_kept;
                if (lane < hk) P.out[ob + lane] = d->held[lane];
                if (nruns > 0) gather_unit<0>(d, P.bytes + d->ustart, P.out + ob, P.bytes + total, lane);
            }
            __syncwarp();
        }
        __syncthreads();  // the next tile overwrites the shared records, descriptors and the ticket
    }
}

}  // namespace qb
