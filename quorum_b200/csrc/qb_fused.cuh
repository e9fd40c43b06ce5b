// Fused single-pass kernels (the default path): one persistent kernel per call does what the token path does in four to
// seven launches, tile by tile, so that the text comes from HBM once.
//
// A tile is the run of consecutive units (requests with their responses / streams with their chunks) whose first byte
// lies in one window of 2^log2T bytes of the flat input, so units are never split and a tile's bytes are contiguous.
// ONE WARP owns a tile from start to end; warps take tiles by ticket (tiles start in index order, which the look-back
// needs for forward progress) and run, per tile:
//   1  tokenise the tile's bytes (tokenize_pair): records into the warp's slice of shared memory (global scratch when
//      the tile spans more steps than the slice holds)
//   2  walk: lane per unit over the records (strip_walk_lane / stream_walk_lane; warp-per-unit and byte-rescanning
//      fallbacks exactly as in the token path) -> kept length, descriptor, carried state
//   3  warp scan of the lengths (+ separators), tile aggregate, decoupled look-back over tiles -> output base
//   4  (streams) per-chunk output offsets, lane per feed boundary
//   5  copy: the kept runs of every unit to their final place (gather_unit)
// There is no block-wide barrier anywhere: the first version of this kernel gave a tile to a CTA and spent its time at
// barriers (ncu: 13 barrier-stalled warps per issued instruction, 4 tile chains per SM, 34 % of the issue slots used)
// because the walk keeps one warp busy while seven wait.  With a warp per tile an SM runs 32 independent chains.
// Nothing is staged in shared memory: step 1 pulls the tile through L2, step 5 re-reads the kept runs a few microseconds
// later while the tile is still L2-resident (measured on cfg3: DRAM read 1.005x the input, written 0.95x the output).
// The descriptors, lengths and records a tile needs live in shared memory, so the chain has no DRAM round trip besides
// the first touch of the text.
#pragma once
#include "qb_lane.cuh"
#include "qb_scan.cuh"
#include "qb_tok.cuh"

namespace qb {

constexpr int kFusedSteps = 18;      // steps of records a warp keeps in shared memory (a 16 KiB window, pair-aligned)
constexpr int kFusedUnits = 8;       // units whose descriptor / lengths a warp keeps in shared memory
constexpr int kFusedMinLog2T = 13;   // smallest window: 8 KiB

// Written by the map kernel: the units / items / bytes a tile owns.  u0p1 == 0: no unit starts in this window.
//   strip:   u = requests [u0p1 - 1, u1), i = responses [i0, i1), b = bytes [b0, b1)
//   streams: u = streams  [u0p1 - 1, u1), i = chunks    [i0, i1), b = bytes [b0, b1)
struct __align__(16) TileRec {
    int64_t u0p1, u1, i0, i1, b0, b1, pad0, pad1;
};
static_assert(sizeof(TileRec) == 64, "tile record layout");

// Zeroed by the host (cudaMemsetAsync): look-back words (tiles and groups of 32 tiles), tile records.
struct TileTables {
    Lookback2 lb;
    TileRec* rec;
    int64_t max_tiles;
    int log2T;
};

// warp-wide inclusive scan
__device__ __forceinline__ int64_t warp_incl_scan(int64_t v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int64_t t = __shfl_up_sync(kFull, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// Tokenise the steps that cover bytes [tb, te) of the flat buffer, one warp.  Returns where the records are: the warp's
// shared slice when the range spans at most kFusedSteps steps (pair-aligned), the call's global record arrays otherwise.
__device__ __forceinline__ TokSrc tokenize_tile(const DevTags& tg, const uint8_t* bytes, int64_t total, int64_t tb, int64_t te,
                                                uint8_t* s_cnt, uint32_t* s_rec, TokWarpScratch& W, uint8_t* g_cnt, uint32_t* g_rec,
                                                int64_t g_max_steps, int lane) {
    TokSrc ts;
    if (te <= tb) return ts;
    const int64_t step_lo = (tb >> 11) << 1, step_hi = (((te - 1) >> 11) + 1) << 1;
    const bool in_smem = step_hi - step_lo <= kFusedSteps;
    int64_t nsteps = (total + 1023) >> 10;
    uint8_t* cnt = g_cnt;
    uint32_t* rec = g_rec;
    if (in_smem) {  // index by absolute step: shift the shared arrays so that the tile's first step lands on entry 0
        cnt = s_cnt - step_lo;
        rec = s_rec - step_lo * kRecPerStep;
    } else if (nsteps > g_max_steps) {
        nsteps = g_max_steps;
    }
    // (the head of the tile was prefetched one tile of work ago; a tile that overhangs its window gets the rest now)
    for (int64_t o = ((tb & ~(int64_t)127) + 16384) + (int64_t)lane * 128; o < te; o += 32 * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(bytes + o));
    for (int64_t pr = tb >> 11; pr <= ((te - 1) >> 11); pr++) tokenize_pair(tg, bytes, total, nsteps, pr, lane, cnt, rec, W);
    __syncwarp();
    ts.cnt = cnt;
    ts.rec = rec;
    return ts;
}

// Per-phase time of the tile chains (debug builds only, -DQB_PHASE_TIMING): lane 0 of every warp sums the nanoseconds it
// spends in each phase; tools/phase_times.py prints the split.  Not compiled into the product library.
#ifdef QB_PHASE_TIMING
__device__ unsigned long long g_phase_ns[16];
__device__ __forceinline__ unsigned long long qb_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define QB_PHASE_DECL unsigned long long ph_t = qb_now(), ph_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define QB_PHASE(i) do { const unsigned long long n_ = qb_now(); ph_acc[i] += n_ - ph_t; ph_t = n_; } while (0)
#define QB_PHASE_FLUSH do { if (lane == 0) for (int i_ = 0; i_ < 8; i_++) atomicAdd(&g_phase_ns[i_], ph_acc[i_]); } while (0)
#else
#define QB_PHASE_DECL
#define QB_PHASE(i)
#define QB_PHASE_FLUSH
#endif

// every lane gets the tile's record (lanes 0-5 load one field each)
__device__ __forceinline__ void load_tile(const TileTables& tt, int64_t t, int lane, TileRec& r) {
    int64_t v = 0;
    if (lane < 6) v = reinterpret_cast<const int64_t*>(tt.rec + t)[lane];
    r.u0p1 = __shfl_sync(kFull, v, 0);
    r.u1 = __shfl_sync(kFull, v, 1);
    r.i0 = __shfl_sync(kFull, v, 2);
    r.i1 = __shfl_sync(kFull, v, 3);
    r.b0 = __shfl_sync(kFull, v, 4);
    r.b1 = __shfl_sync(kFull, v, 5);
}

// the warp's next tile is known in advance (static assignment): get its record and the head of its text moving
__device__ __forceinline__ void prefetch_tile(const TileTables& tt, int64_t t, int64_t ntiles, const uint8_t* bytes, int64_t begin,
                                              int64_t total, int lane) {
    if (t >= ntiles) return;
    if (lane == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(tt.rec + t));
    // units start in window t, so the tile's bytes start inside [begin + t * T, begin + (t + 1) * T)
    const int64_t w0 = begin + (t << tt.log2T);
    for (int64_t o = (w0 & ~(int64_t)127) + (int64_t)lane * 128; o < w0 + ((int64_t)1 << tt.log2T) && o < total; o += 32 * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(bytes + o));
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

// thread per request: the window its first byte lies in; the first / last request of a window write the tile's record
__global__ void __launch_bounds__(256) strip_tile_map_kernel(const int64_t* __restrict__ in_off, const int64_t* __restrict__ req_off,
                                                             int64_t n_req, TileTables tt) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_req) return;
    const int64_t begin = in_off[req_off ? req_off[0] : 0];
    auto first_resp = [&](int64_t qq) { return req_off ? req_off[qq] : qq; };
    auto win = [&](int64_t qq) { return (in_off[first_resp(qq)] - begin) >> tt.log2T; };
    const int64_t id = win(q);
    if (id >= tt.max_tiles) return;  // only when out_capacity is below the input size (contract violated): stay inside the tables
    TileRec* r = tt.rec + id;
    if (q == 0 || win(q - 1) != id) {
        const int64_t r0 = first_resp(q);
        r->u0p1 = q + 1;
        r->i0 = r0;
        r->b0 = in_off[r0];
    }
    if (q == n_req - 1 || win(q + 1) != id) {
        const int64_t r1 = first_resp(q + 1);
        r->u1 = q + 1;
        r->i1 = r1;
        r->b1 = in_off[r1];
    }
}

struct __align__(16) WarpStripShared {
    uint32_t rec[kFusedSteps * kRecPerStep];
    UnitDesc desc[kFusedUnits];
    int64_t klen[kFusedUnits], first[kFusedUnits], dst[kFusedUnits];
    unsigned long long lastclose[kMaxCls];
    TokWarpScratch tok;
    uint8_t cnt[kFusedSteps + 6];
};

__global__ void __launch_bounds__(kThreads, 4)
strip_fused_kernel(const __grid_constant__ DevTags tg, const __grid_constant__ StripFusedParams P) {
    __shared__ __align__(16) WarpStripShared s_warp[kWarpsPerBlock];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    WarpStripShared& S = s_warp[wib];
    const int64_t* __restrict__ in_off = P.in_off;
    const int64_t* __restrict__ req_off = P.req_off;
    const int64_t begin = in_off[0], total = in_off[P.n_resp];
    int64_t ntiles = ((total - begin) >> P.tt.log2T) + 1;
    if (ntiles > P.tt.max_tiles) ntiles = P.tt.max_tiles;
    const int64_t init = P.chain ? P.out_off[0] : 0;  // chained call: this slice's output starts where the previous one ended
    const int hide = P.hide, sj = P.stream_join;
    const int32_t sep_len = P.sep_len;
    // Tiles are dealt round-robin to the warps of the (co-resident, cooperatively launched) grid: consecutive tiles run at
    // the same time on different warps, which is what the look-back wants, and there is no ticket counter for ~4,700
    // warps to contend on (measured: a ticket per tile costs ~25 ns of serialised same-address atomics, 7 ms on cfg3).
    const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;
    QB_PHASE_DECL;
    for (int64_t t = (int64_t)blockIdx.x * kWarpsPerBlock + wib; t < ntiles; t += nwarps) {
        TileRec tr;
        load_tile(P.tt, t, lane, tr);
        prefetch_tile(P.tt, t + nwarps, ntiles, P.in, begin, total, lane);
        QB_PHASE(0);
        if (tr.u0p1 == 0) {  // no request starts in this window
            lookback2_exclusive(P.tt.lb, t, 0, ntiles, lane);
            continue;
        }
        const int64_t q0 = tr.u0p1 - 1, q1 = tr.u1, r0 = tr.i0, r1 = tr.i1, tb = tr.b0, te = tr.b1;
        // small tile: descriptors, lengths and destinations stay in shared memory (indexed by response, shifted)
        const bool small = r1 - r0 <= kFusedUnits;
        UnitDesc* desc = small ? S.desc - r0 : P.desc;
        int64_t* klen = small ? S.klen - r0 : P.klen;
        int64_t* first = small ? S.first - r0 : P.first;
        int64_t* dst = small ? S.dst - r0 : P.dst;

        // ---- 1: records ----
        TokSrc ts;
        if (hide) ts = tokenize_tile(tg, P.in, total, tb, te, S.cnt, S.rec, S.tok, P.g_cnt, P.g_rec, P.g_max_steps, lane);
        QB_PHASE(1);

        // ---- 2: walk (lane per response, then the whole warp for the ones that need it) ----
        if (hide) {
            for (int64_t g0 = r0; g0 < r1; g0 += 32) {
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
                    StripRunWalker<false> w(tg, P.in, total, aj, bj - aj, S.lastclose, lane, ts);
                    w.desc = desc + rj;
                    w.run();
                    res.kept_len = w.kept_len;
                    res.first_pos = 0;
                    bool slow = w.overflow;
                    if (w.many) slow = !strip_count_wide(tg, P.in, total, aj, bj - aj, S.lastclose, lane, ts, desc + rj, &res.kept_len);
                    if (slow) {
                        res = strip_count_slow(tg, P.in, total, aj, bj - aj, S.lastclose, lane);
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
            __syncwarp();
        }

        QB_PHASE(2);
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
        for (int64_t qb = q0; qb < q1; qb += 32) {
            const int64_t q = qb + lane;
            const int64_t len = q < q1 ? request_len(q, 0, false) : 0;
            const int64_t incl = warp_incl_scan(len, lane);
            const int64_t excl = run + incl - len;
            if (qb == q0) rel0 = excl;
            else if (q < q1) P.out_off[q] = excl;  // tile-relative until the base is known
            run += __shfl_sync(kFull, incl, 31);
        }
        QB_PHASE(3);
        const int64_t base = init + lookback2_exclusive(P.tt.lb, t, run, ntiles, lane);
        QB_PHASE(4);
        for (int64_t qb = q0; qb < q1; qb += 32) {
            const int64_t q = qb + lane;
            if (q < q1) {
                const int64_t o = base + (qb == q0 ? rel0 : P.out_off[q]);
                P.out_off[q] = o;
                request_len(q, o, true);
            }
        }
        if (q1 == P.n_req && lane == 0) P.out_off[P.n_req] = base + run;
        __syncwarp();
        QB_PHASE(5);

        // ---- 5: copy ----
        for (int64_t r = r0; r < r1; r++) {
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
                    gather_unit<2>(desc + r, P.in + a, P.out + d, P.in + total, lane);
                } else if (nr == -2) {  // 9..32 kept runs: walk the records again and copy from the lanes
                    strip_copy_wide(tg, P.in, total, a, b - a, S.lastclose, lane, ts, kl, P.out + d);
                } else {
                    strip_copy_slow(tg, P.in, total, a, b - a, S.lastclose, lane, first[r], kl, P.out + d);
                }
            }
            __syncwarp();
        }
        __syncwarp();  // the next tile overwrites the warp's shared records and descriptors
        QB_PHASE(6);
    }
    QB_PHASE_FLUSH;
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
    TileRec* r = tt.rec + id;
    if (s == 0 || win(s - 1) != id) {
        const int64_t k0 = stream_off[s];
        r->u0p1 = s + 1;
        r->i0 = k0;
        r->b0 = chunk_off[k0];
    }
    if (s == n_streams - 1 || win(s + 1) != id) {
        const int64_t k1 = stream_off[s + 1];
        r->u1 = s + 1;
        r->i1 = k1;
        r->b1 = chunk_off[k1];
    }
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

struct __align__(16) WarpStreamShared {
    uint32_t rec[kFusedSteps * kRecPerStep];
    StreamDesc desc[kFusedUnits];
    int64_t bases[kFusedUnits];
    TokWarpScratch tok;
    uint8_t cnt[kFusedSteps + 6];
};

__global__ void __launch_bounds__(kThreads, 4)
stream_fused_kernel(const __grid_constant__ DevTags tg, const __grid_constant__ StreamFusedParams P) {
    __shared__ __align__(16) WarpStreamShared s_warp[kWarpsPerBlock];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    WarpStreamShared& S = s_warp[wib];
    const int64_t* __restrict__ chunk_off = P.chunk_off;
    const int64_t* __restrict__ stream_off = P.stream_off;
    const int64_t total = chunk_off[P.n_chunks];
    const int64_t cb = stream_off[0];
    const int64_t begin = chunk_off[cb], end = chunk_off[stream_off[P.n_streams]];
    int64_t ntiles = ((end - begin) >> P.tt.log2T) + 1;
    if (ntiles > P.tt.max_tiles) ntiles = P.tt.max_tiles;
    const int64_t init = P.chain ? P.out_chunk_off[cb] : 0;  // chained call: continue where the previous slice ended
    const uint32_t flags = P.flags;
    const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerBlock;  // tiles are dealt round-robin to the warps (see strip_fused_kernel)
    for (int64_t t = (int64_t)blockIdx.x * kWarpsPerBlock + wib; t < ntiles; t += nwarps) {
        TileRec tr;
        load_tile(P.tt, t, lane, tr);
        prefetch_tile(P.tt, t + nwarps, ntiles, P.bytes, begin, total, lane);
        if (tr.u0p1 == 0) {  // no stream starts in this window
            lookback2_exclusive(P.tt.lb, t, 0, ntiles, lane);
            continue;
        }
        const int64_t s0 = tr.u0p1 - 1, s1 = tr.u1, k0 = tr.i0, k1 = tr.i1, tb = tr.b0, te = tr.b1;
        const bool small = s1 - s0 <= kFusedUnits;
        StreamDesc* desc = small ? S.desc - s0 : P.desc;
        int64_t* bases = small ? S.bases - s0 : P.bases;

        // ---- 1: records ----
        const TokSrc ts = tokenize_tile(tg, P.bytes, total, tb, te, S.cnt, S.rec, S.tok, P.g_cnt, P.g_rec, P.g_max_steps, lane);

        // ---- 2 + 3: walk (lane per stream, the whole warp for the ones that need it), scan of the released totals ----
        int64_t run = 0;
        for (int64_t g0 = s0; g0 < s1; g0 += 32) {
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
            const int64_t incl = warp_incl_scan(tot, lane);
            if (have) bases[s] = run + incl - tot;  // tile-relative
            run += __shfl_sync(kFull, incl, 31);
        }
        __syncwarp();
        const int64_t base = init + lookback2_exclusive(P.tt.lb, t, run, ntiles, lane);

        // ---- 4: per-chunk output offsets, lane per feed boundary ----
        if (s0 == 0 && lane == 0) P.out_chunk_off[cb] = base;
        if (s1 - s0 == 1) {  // the common shape: one stream in the tile
            const StreamDesc* d = desc + s0;
            if (d->nruns >= 0) {
                const int64_t ob = base + bases[s0], us = d->ustart;
                for (int64_t k = k0 + lane; k < k1; k += 32) P.out_chunk_off[k + 1] = ob + released_before(d, (int)(chunk_off[k + 1] - us));
            }
        } else {
            for (int64_t k = k0 + lane; k < k1; k += 32) {
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
        }

        // ---- 5: finish: released carried bytes and kept runs to their final place ----
        for (int64_t s = s0; s < s1; s++) {
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
                if (nruns > 0) gather_unit<2>(d, P.bytes + d->ustart, P.out + ob, P.bytes + total, lane);
            }
            __syncwarp();
        }
        __syncwarp();  // the next tile overwrites the warp's shared records and descriptors
    }
}

}  // namespace qb
