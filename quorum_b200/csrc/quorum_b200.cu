// quorum_b200 — kernels + C ABI (include/quorum_b200.h).  sm_100a only; no CPU fallback.
//
// Default path (QB200_PATH=token), see DESIGN.md section 5.1:
//   stream_feed :  tokenize (one record per '<')  ->  count (lane per stream: automaton over the records, descriptor,
//                  carried state, prefix sum over the streams by look-back)  ->  offsets (thread per feed boundary)
//                  ->  finish (warp per stream: kept runs to their final place)
//   strip_concat:  tokenize  ->  count (lane per response: liveness, automaton, strip(), descriptor / run pool)  ->  request
//                  offsets (per-request lengths incl. separators, prefix sum by look-back)  ->  copy (warp per response)
// QB200_PATH=general: the byte-rescanning warp-per-unit walkers (count, scan, copy).
// Then the rows around the filter: SSE framing (qb_sse.cuh), SSE parsing (qb_ssein.cuh), aggregate prompts (qb_agg.cuh).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <vector>
#include <new>

#include "../../include/quorum_b200.h"
#include "qb_walk.cuh"
#include "qb_scan.cuh"
#include "qb_tok.cuh"
#include "qb_lane.cuh"
#include "qb_scanwalk.cuh"
#include "qb_sse.cuh"
#include "qb_agg.cuh"
#include "qb_ssein.cuh"
#include <stdlib.h>

using namespace qb;

struct qb200_tags {
    DevTags dev;
};

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

static int fail(int code, const char* fmt, const char* a = "") {
    snprintf(g_err, sizeof(g_err), fmt, a);
    return code;
}

#define CUDA_TRY(expr)                                                          \
    do {                                                                        \
        cudaError_t e_ = (expr);                                                \
        if (e_ != cudaSuccess) return fail(QB200_ECUDA, "CUDA error: %s", cudaGetErrorString(e_)); \
    } while (0)

// Every kernel launch goes through here: the launch counter, and (qb200_profile_begin) a CUDA event after the launch on
// the launching stream so that a benchmark can read the per-kernel split of a call without a profiler attached.
struct ProfMark { const char* name; cudaEvent_t ev; };
static std::atomic<bool> g_prof_on{false};
static std::mutex g_prof_mu;  // the mark list is shared by every thread that launches while profiling is on
static std::vector<ProfMark> g_prof;
static void prof_mark(const char* name, cudaStream_t stream) {
    ProfMark m{name, nullptr};
    if (cudaEventCreate(&m.ev) != cudaSuccess) return;
    cudaEventRecord(m.ev, stream);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof.push_back(m);
}
#define LAUNCHED(name) do { g_launches.fetch_add(1, std::memory_order_relaxed); if (g_prof_on.load(std::memory_order_relaxed)) prof_mark(name, stream); } while (0)

// ---------------------------------------------------------------------------------------------------------------
// Byte-rescanning walkers, out of line: used when no tokenizer records exist (QB200_PATH=general, the no-alloc entry
// point) or when a unit touches a step whose record slots overflowed.  Kept out of the record-driven kernels' register
// budget.
// ---------------------------------------------------------------------------------------------------------------
__device__ __noinline__ StripResult strip_count_slow(const DevTags& tg, const uint8_t* in, int64_t total, int64_t a, int64_t n,
                                                     unsigned long long* lc, int lane) {
    StripWalker<false> w(tg, in, total, a, n, lc, lane);
    w.run();
    return w.res;
}
__device__ __noinline__ void strip_copy_slow(const DevTags& tg, const uint8_t* in, int64_t total, int64_t a, int64_t n,
                                             unsigned long long* lc, int lane, int64_t first, int64_t kl, uint8_t* dst) {
    StripWalker<true> w(tg, in, total, a, n, lc, lane);
    w.res.first_pos = first;
    w.res.kept_len = kl;
    w.dst = dst;
    w.run();
}
// 9 or more kept runs: record-driven walk (32 runs at a time) + copy from the lanes, out of line for the same reason
__device__ __noinline__ void strip_copy_wide(const DevTags& tg, const uint8_t* in, int64_t total, int64_t a, int64_t n,
                                             unsigned long long* lc, int lane, TokSrc ts, int64_t kl, uint8_t* dst) {
    StripRunWalker<true, true> w(tg, in, total, a, n, lc, lane, ts);
    w.dst = dst;
    w.klimit = kl;
    w.run();
}
// count pass of a unit with more than 32 kept runs; returns false when even that cannot settle it
__device__ __noinline__ bool strip_count_wide(const DevTags& tg, const uint8_t* in, int64_t total, int64_t a, int64_t n,
                                              unsigned long long* lc, int lane, TokSrc ts, UnitDesc* desc, int64_t* kept_len) {
    StripRunWalker<false, true> w(tg, in, total, a, n, lc, lane, ts);
    w.desc = desc;
    w.run();
    *kept_len = w.kept_len;
    return !w.overflow;
}
// the record-driven warp walker of one response, out of line for the same reason (the lane walker is the common case)
__device__ __noinline__ StripResult strip_count_warp(const DevTags& tg, const uint8_t* in, int64_t total, int64_t a, int64_t n,
                                                     unsigned long long* lc, int lane, TokSrc ts, UnitDesc* desc) {
    StripResult res;
    res.first_pos = 0;
    StripRunWalker<false> w(tg, in, total, a, n, lc, lane, ts);
    w.desc = desc;
    if (n == 0 && lane == 0) desc->nruns = 0;  // empty response: nothing to walk, nothing to copy
    w.run();
    res.kept_len = w.kept_len;
    bool slow = w.overflow;
    if (w.many) slow = !strip_count_wide(tg, in, total, a, n, lc, lane, ts, desc, &res.kept_len);
    if (slow) {
        res = strip_count_slow(tg, in, total, a, n, lc, lane);
        if (lane == 0) desc->nruns = -1;  // redone from the bytes by the copy kernel
    }
    return res;
}
// one response by the whole warp, out of line: long responses go to the scan walker (lanes own steps, state carried by the
// scan of transition vectors, qb_scanwalk.cuh); what it cannot take goes to the run-list warp walker, then to the byte walker
__device__ __noinline__ StripResult strip_count_unit(const DevTags& tg, const uint8_t* in, int64_t total, int64_t a, int64_t n,
                                                     unsigned long long* lc, int2* stage, int lane, TokSrc ts, UnitDesc* desc,
                                                     const RunPool& pool, int scanwalk) {
    StripResult res;
    res.first_pos = 0;
    res.kept_len = 0;
    if (!ts.cnt) return strip_count_slow(tg, in, total, a, n, lc, lane);
    if (scanwalk && strip_scan_walk(tg, in, a, n, ts, desc, &res.kept_len, lc, stage, pool, lane)) return res;
    return strip_count_warp(tg, in, total, a, n, lc, lane, ts, desc);
}
__device__ __noinline__ StreamResult stream_count_slow(const DevTags& tg, const uint8_t* bytes, int64_t total, const int64_t* co,
                                                       int64_t m, int lane, uint32_t flags, const StreamState* st_in) {
    StreamWalker<false> w(tg, bytes, total, co, m, lane, flags);
    return w.run(st_in, nullptr);
}
// one stream by the whole warp, out of line (the lane walker is the common case; inline, the warp walkers cost the count
// kernel its registers): record-driven run-list walker, then the byte-rescanning walker
__device__ __noinline__ StreamResult stream_count_unit(const DevTags& tg, const uint8_t* bytes, int64_t total, const int64_t* co, int64_t m,
                                                       int lane, uint32_t flags, TokSrc ts, const StreamState* st_in, StreamState* st_out,
                                                       StreamDesc* desc) {
    StreamResult r;
    bool slow = ts.cnt == nullptr;
    if (!slow) {
        StreamRunWalker w(tg, bytes, total, co, m, lane, flags, ts);
        w.desc = desc;
        if (m <= 0 && lane == 0) desc->nruns = desc->held_kept = 0;  // nothing fed: nothing to copy
        // record-driven units get their carried state here (st_in may alias st_out; a unit that overflows returns before
        // it writes any and is redone, state included, by the finish kernel)
        r = w.run(st_in, st_out);
        slow = w.overflow;  // does not fit the descriptor: redo this unit from its bytes
    }
    if (slow) {
        r = stream_count_slow(tg, bytes, total, co, m, lane, flags, st_in);
        if (desc && lane == 0) desc->nruns = -1;
    }
    return r;
}
__device__ __noinline__ void stream_copy_slow(const DevTags& tg, const uint8_t* bytes, int64_t total, const int64_t* co, int64_t m,
                                              int lane, uint32_t flags, const StreamState* st_in, StreamState* st_out,
                                              int64_t base, uint8_t* out, int64_t* out_chunk_off, int64_t c0) {
    StreamWalker<true> w(tg, bytes, total, co, m, lane, flags);
    w.out_base = base;
    w.dst = out + base;
    w.out_chunk_off = out_chunk_off;
    w.chunk0 = c0;
    w.run(st_in, st_out);
}

// ---------------------------------------------------------------------------------------------------------------
// kernels: strip + concat
// ---------------------------------------------------------------------------------------------------------------
constexpr int kOffsetsPer = 8;                    // boundaries per thread in stream_offsets_kernel
constexpr int kOffsetsTile = 32 * kOffsetsPer;   // boundaries per warp there
constexpr int kWarpsPerBlock = 8;
constexpr int kThreads = kWarpsPerBlock * 32;

// Software pipelining of the warp-per-unit loops: while unit r is being walked, the lines the NEXT unit of this warp will
// touch first (its record counts, its records, the first and last bytes of its text) are already on their way.
__device__ __forceinline__ void prefetch_unit_meta(const TokSrc& ts, const uint8_t* bytes, int64_t a, int64_t b, int lane) {
    if (!ts.cnt || b <= a) return;
    const int64_t s0 = a >> 10, s1 = (b - 1) >> 10;
    const void* p = nullptr;
    if (lane == 0) p = ts.cnt + s0;
    else if (lane == 1) p = ts.cnt + s1;
    else if (lane == 2) p = bytes + a;
    else if (lane == 3) p = bytes + b - 1;
    else if (s0 + (lane - 4) <= s1) p = ts.head(s0 + (lane - 4));  // the first records of four steps per 128-byte line
    if (p) asm volatile("prefetch.L1 [%0];" ::"l"(p));
}

// scratch layout for strip_concat: int64 klen[n_resp] | int64 first[n_resp] | int64 dst[n_resp] | int64 reqlen[n_req+1]
// (five or six blocks per SM - 48 / 40 registers - change nothing: cfg3 0.221 / 0.226 ms against 0.217)
//
// A warp takes 32 consecutive responses, one per lane (qb_lane.cuh).  The ones a lane cannot take (longer than lane_steps,
// more classes than a lane tracks, pool exhausted, an overflowed step) need a whole warp each; they go to a queue that
// strip_count_long_kernel drains with every warp of the GPU, one response per warp - walked where they are found, 32 long
// responses would run one after the other on ONE warp (a batch of 4,096 responses of 256 KiB ran on 128 warps: 1.5 ms).
// (Tried: the long responses walked inside strip_count_kernel by its first 148..444 blocks - a contiguous share of the
// responses per block, its warps taking the long ones by ticket - so that the two kinds of walk overlap: cfg5's count
// 0.34 + 0.20 -> 0.61..0.75 ms, the static shares are uneven and fewer warps are left for the long walks.  And: the long
// responses listed by size and strip_count_long_kernel run on a second stream beside strip_count_kernel: cfg5 1.492 ->
// 1.501 ms - the two kernels use 50 % and 40 % of the issue slots at 2.4 and 13 active lanes per instruction, so side by
// side they take as long as one after the other; fewer instructions, not more overlap, is what the count phase needs.)
// (Tried: the record counts and first records of a warp's 32 consecutive responses fetched into shared memory with
// coalesced 16-byte loads, the lanes walking from there: cfg3's count 0.188 -> 0.234 ms at three blocks per SM; the
// generic-address loads and the lost block cost more than the sector-per-lane loads they replace.)
struct LongQueue {
    int64_t* items = nullptr;          // response indices (the `dst` array, not needed before the request offsets)
    unsigned long long* count = nullptr;
};
__global__ void __launch_bounds__(kThreads, 4)
strip_count_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off,
                   int64_t n_resp, int64_t* __restrict__ klen, int64_t* __restrict__ first, TokSrc ts,
                   UnitDesc* __restrict__ desc, int lane_steps, RunPool pool, LongQueue lq) {
    grid_dep_wait();
    __shared__ unsigned long long lastclose[kWarpsPerBlock][kMaxCls];
    __shared__ int s_lc[kLaneCls][kThreads];  // lane walker: last closer per class, one column per thread
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t total = in_off[n_resp];
    const int64_t ngroups = (n_resp + 31) >> 5;
    for (int64_t g = (int64_t)blockIdx.x * kWarpsPerBlock + wib; g < ngroups; g += (int64_t)gridDim.x * kWarpsPerBlock) {
        const int64_t r = (g << 5) + lane;
        const bool have = r < n_resp;
        int64_t a = 0, b = 0, kl = 0, fp = 0;
        bool done = false;
        if (have) {
            a = in_off[r];
            b = in_off[r + 1];
            if (ts.cnt && (b <= a || ((b - 1) >> 10) - (a >> 10) < lane_steps)) done = strip_walk_lane(tg, in, a, b - a, ts, desc + r, kl, &s_lc[0][threadIdx.x], kThreads, pool);
        }
        __syncwarp();
        unsigned todo = __ballot_sync(kFull, have && !done);
        if (ts.cnt) {  // record-driven call: the leftovers wait for strip_count_long_kernel
            if (todo) {
                unsigned long long at = 0;
                const int leader = __ffs(todo) - 1;
                if (lane == leader) at = atomicAdd(lq.count, (unsigned long long)__popc(todo));
                at = __shfl_sync(kFull, at, leader);
                if (have && !done) lq.items[at + __popc(todo & ((1u << lane) - 1u))] = r;
            }
            if (have && done) {
                klen[r] = kl;
                first[r] = fp;
            }
            continue;
        }
        while (todo) {  // no records (QB200_PATH=general): one warp per response, from the bytes
            const int j = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t aj = __shfl_sync(kFull, a, j), bj = __shfl_sync(kFull, b, j);
            const StripResult res = strip_count_slow(tg, in, total, aj, bj - aj, lastclose[wib], lane);
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
}

// the queued responses, one per warp, over the whole GPU
__global__ void __launch_bounds__(kThreads, 4)
strip_count_long_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off,
                        int64_t n_resp, int64_t* __restrict__ klen, int64_t* __restrict__ first, TokSrc ts,
                        UnitDesc* __restrict__ desc, RunPool pool, LongQueue lq, int scanwalk) {
    grid_dep_wait();
    __shared__ unsigned long long lastclose[kWarpsPerBlock][kMaxCls];
    __shared__ int2 s_stage[kWarpsPerBlock][kDescRuns + 1];  // scan walker: the runs of a unit that fits a descriptor
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t total = in_off[n_resp];
    const int64_t nq = (int64_t)*lq.count;
    for (int64_t i = (int64_t)blockIdx.x * kWarpsPerBlock + wib; i < nq; i += (int64_t)gridDim.x * kWarpsPerBlock) {
        const int64_t r = lq.items[i];
        const int64_t a = in_off[r], b = in_off[r + 1];
        const StripResult res = strip_count_unit(tg, in, total, a, b - a, lastclose[wib], s_stage[wib], lane, ts, desc + r, pool, scanwalk);
        if (lane == 0) {
            klen[r] = res.kept_len;
            first[r] = res.first_pos;
        }
        __syncwarp();
    }
}

// Exclusive scan of v[0..n) into out[0..n], out[n] = total, in three small launches:
//   scan_partials_kernel   block b sums its slice of kScanSlice elements            -> partial[b]
//   scan_kernel            one block scans the (few) partials in place               -> partial[b] = offset of slice b
//   scan_apply_kernel      block b rescans its slice with that offset
constexpr int kScanSlice = 2048;  // elements per block (256 threads x 8)

__global__ void __launch_bounds__(256) scan_partials_kernel(const int64_t* __restrict__ v, int64_t n, int64_t* __restrict__ partial) {
    __shared__ int64_t ws[8];
    const int64_t a = (int64_t)blockIdx.x * kScanSlice;
    int64_t sum = 0;
#pragma unroll 4
    for (int k = 0; k < kScanSlice / 256; k++) {
        const int64_t i = a + k * 256 + threadIdx.x;
        if (i < n) sum += v[i];
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(kFull, sum, d);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        int64_t t = 0;
        for (int w = 0; w < 8; w++) t += ws[w];
        partial[blockIdx.x] = t;
    }
}

// single-block exclusive scan (in place allowed: out may alias v); out[n] = total.  Warp w owns a contiguous slice.
__global__ void __launch_bounds__(1024) scan_kernel(const int64_t* v, int64_t n, int64_t* out, const int64_t* init_arr,
                                                    const int64_t* init_idx) {
    __shared__ int64_t warp_excl[32];
    const int64_t init = init_arr ? init_arr[init_idx ? *init_idx : 0] : 0;  // chained calls start at a given offset
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t per = ((n + 31) / 32 + 31) / 32 * 32;  // slice length, a multiple of 32
    const int64_t a = (int64_t)wid * per, b = a + per < n ? a + per : n;
    int64_t sum = 0;
    for (int64_t i = a + lane; i < b; i += 32) sum += v[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(kFull, sum, d);
    if (lane == 0) warp_excl[wid] = sum;
    __syncthreads();
    int64_t total = 0;
    if (wid == 0) {
        const int64_t ws = warp_excl[lane];
        int64_t wi = ws;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int64_t t = __shfl_up_sync(kFull, wi, d);
            if (lane >= d) wi += t;
        }
        warp_excl[lane] = wi - ws;
        total = wi;
    }
    __syncthreads();
    int64_t run = warp_excl[wid] + init;
    for (int64_t i0 = a; i0 < b; i0 += 32) {
        const int64_t i = i0 + lane;
        const int64_t x = i < b ? v[i] : 0;
        int64_t incl = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int64_t t = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += t;
        }
        if (i < b) out[i] = run + incl - x;
        run += __shfl_sync(kFull, incl, 31);
    }
    __syncthreads();
    if (wid == 0 && lane == 31) out[n] = total + init;
}

__global__ void __launch_bounds__(256) scan_apply_kernel(const int64_t* __restrict__ v, int64_t n, const int64_t* __restrict__ partial,
                                                         int64_t nblocks, int64_t* __restrict__ out, const int64_t* init_arr,
                                                         const int64_t* init_idx) {
    __shared__ int64_t ws[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t a = (int64_t)blockIdx.x * kScanSlice;
    const int64_t init = init_arr ? init_arr[init_idx ? *init_idx : 0] : 0;
    int64_t run = partial[blockIdx.x] + init;
    for (int k = 0; k < kScanSlice / 256; k++) {
        const int64_t i = a + k * 256 + threadIdx.x;
        const int64_t x = i < n ? v[i] : 0;
        int64_t incl = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int64_t t = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) ws[wid] = incl;
        __syncthreads();
        int64_t before = 0, all = 0;
        for (int w = 0; w < 8; w++) {
            const int64_t t = ws[w];
            if (w < wid) before += t;
            all += t;
        }
        if (i < n) out[i] = run + before + incl - x;
        run += all;
        __syncthreads();
    }
    if (blockIdx.x == nblocks - 1 && threadIdx.x == 0) out[n] = partial[nblocks] + init;
}

// host helper: out[0..n] = exclusive scan of v[0..n); `partial` needs n / kScanSlice + 2 entries
static void launch_scan(const int64_t* v, int64_t n, int64_t* out, int64_t* partial, cudaStream_t stream,
                        const int64_t* init_arr = nullptr, const int64_t* init_idx = nullptr) {
    if (n <= 4 * kScanSlice) {
        scan_kernel<<<1, 1024, 0, stream>>>(v, n, out, init_arr, init_idx);
        LAUNCHED("scan_kernel");
        return;
    }
    const int64_t nb = (n + kScanSlice - 1) / kScanSlice;
    scan_partials_kernel<<<(unsigned)nb, 256, 0, stream>>>(v, n, partial);
    LAUNCHED("scan_partials_kernel");
    scan_kernel<<<1, 1024, 0, stream>>>(partial, nb, partial, nullptr, nullptr);
    LAUNCHED("scan_kernel");
    scan_apply_kernel<<<(unsigned)nb, 256, 0, stream>>>(v, n, partial, nb, out, init_arr, init_idx);
    LAUNCHED("scan_apply_kernel");
}

// The join layout in one kernel (instead of layout, a three-kernel scan and finalize): a block takes 256 consecutive
// requests by ticket, a thread sums its request (kept lengths + separators), block scan + decoupled look-back over the
// blocks give the request's output offset, then the thread writes its responses' destinations.
struct ReqScan {
    unsigned long long* state;  // one look-back word per block of 256 requests (zeroed before the launch)
    unsigned int* ticket;
    int chain;                  // out_off[0] already holds where this call's output starts
};
__global__ void __launch_bounds__(256)
request_offsets_kernel(const int64_t* __restrict__ in_off, const int64_t* __restrict__ req_off, int64_t n_req,
                       const int64_t* __restrict__ klen, int hide, int stream_join, int32_t sep_len, int64_t* __restrict__ dst,
                       int64_t* __restrict__ out_off, ReqScan scan) {
    grid_dep_wait();
    __shared__ int64_t s_wsum[8];
    __shared__ int64_t s_excl, s_ticket;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t init = scan.chain ? out_off[0] : 0;
    const int64_t nblk = (n_req + 255) >> 8;
    for (;;) {
        if (threadIdx.x == 0) s_ticket = (int64_t)atomicAdd(scan.ticket, 1u);
        __syncthreads();
        const int64_t t = s_ticket;
        if (t >= nblk) break;
        const int64_t q = (t << 8) + threadIdx.x;
        int64_t r0 = 0, r1 = 0, len = 0;
        if (q < n_req) {
            r0 = req_off ? req_off[q] : q;
            r1 = req_off ? req_off[q + 1] : q + 1;
            bool first = true;
            for (int64_t r = r0; r < r1; r++) {
                const int64_t l = in_off[r + 1] - in_off[r];
                if (stream_join && l == 0) continue;  // oai_proxy.py:763 `if text`
                if (!first) len += sep_len;
                first = false;
                len += hide ? klen[r] : l;
            }
        }
        int64_t incl = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int64_t v = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) s_wsum[wib] = incl;
        __syncthreads();
        int64_t before = 0, block_tot = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            const int64_t v = s_wsum[w];
            if (w < wib) before += v;
            block_tot += v;
        }
        if (wib == 0) {
            const int64_t excl = lookback_exclusive(scan.state, t, block_tot, lane);
            if (lane == 0) s_excl = excl;
        }
        __syncthreads();
        if (q < n_req) {
            const int64_t base = init + s_excl + before + incl - len;
            out_off[q] = base;
            if (q == n_req - 1) out_off[n_req] = base + len;
            int64_t o = 0;
            bool first = true;
            for (int64_t r = r0; r < r1; r++) {
                const int64_t l = in_off[r + 1] - in_off[r];
                if (stream_join && l == 0) {
                    dst[r] = -1;
                    continue;
                }
                if (!first) o += sep_len;
                dst[r] = ((base + o) << 1) | (first ? 0 : 1);  // bit 0: a separator sits right before this response's text
                first = false;
                o += hide ? klen[r] : l;
            }
        }
    }
}

// Warp-cooperative copy of a unit's kept runs (descriptor) to dst: run i goes to dst + held_kept + (lengths before it).
// PHASE 0: prefetch the source lines and copy; 1: prefetch only; 2: copy only (the caller put work between the two).
// gather_runs: the runs are in registers (lane i < nruns holds run i as [rs, re) relative to src).
template <int PHASE>
__device__ __forceinline__ void gather_runs(int nruns, int rs, int re, int held_kept, const uint8_t* src, uint8_t* dst,
                                            const uint8_t* src_hi, int lane) {
    const int len = re - rs;
    int incl = len;
#pragma unroll
    for (int d = 1; d < kDescRuns; d <<= 1) {
        int t = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += t;
    }
    const int off = held_kept + incl - len;
    const bool tiny = __shfl_sync(kFull, incl, kDescRuns - 1) <= 64 * nruns;  // short runs: one run per lane, byte by byte
    if (PHASE != 2 && !tiny) {
        for (int i = 0; i < nruns; i++) {  // get every source line of every run moving before the first copy waits on one
            const int s_i = __shfl_sync(kFull, rs, i), l_i = __shfl_sync(kFull, len, i);
            const uint8_t* a = src + s_i;
            const uint8_t* base = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(a) & ~(uintptr_t)127);
            const int64_t bytes = (a - base) + l_i;
            for (int64_t o = (int64_t)lane * 128; o < bytes; o += 32 * 128) asm volatile("prefetch.L2 [%0];" ::"l"(base + o));
        }
    }
    if (PHASE == 1) return;
    if (tiny) {
        for (int k = 0; k < len; k++) dst[off + k] = src[rs + k];
        return;
    }
    for (int i = 0; i < nruns; i++) {
        const int s_i = __shfl_sync(kFull, rs, i), l_i = __shfl_sync(kFull, len, i), o_i = __shfl_sync(kFull, off, i);
        if (l_i > 0) warp_copy_fast(dst + o_i, src + s_i, l_i, src_hi, lane);
    }
}

// The same for a unit whose runs live in the pool (more than kDescRuns): 32 runs at a time, one per lane for the offsets,
// then run by run - long runs by the whole warp with 16-byte stores, short ones as lane-strided bytes (coalesced).
__device__ __forceinline__ void gather_pooled(const int2* __restrict__ runs, int nruns, const uint8_t* src, uint8_t* dst,
                                              const uint8_t* src_hi, int lane) {
    int done = 0;
    for (int base = 0; base < nruns; base += 32) {
        const int i = base + lane;
        int2 run = make_int2(0, 0);
        if (i < nruns) run = runs[i];
        const int len = run.y - run.x;
        if (len > 0) {  // the first lines of every run of the batch on their way to L2 before the first copy waits on one
            const uint8_t* a = src + run.x;  // (cfg5's copy 0.48 -> 0.44 ms)
            const uint8_t* base = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(a) & ~(uintptr_t)127);
            const int bytes = min((int)(a - base) + len, 8 * 128);
            for (int o = 0; o < bytes; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(base + o));
        }
        int incl = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(kFull, incl, d);
            if (lane >= d) incl += t;
        }
        const int off = done + incl - len;
        const int m = min(32, nruns - base);
        for (int j = 0; j < m; j++) {
            const int s_j = __shfl_sync(kFull, run.x, j), l_j = __shfl_sync(kFull, len, j), o_j = __shfl_sync(kFull, off, j);
            if (l_j >= 128) {
                warp_copy_fast(dst + o_j, src + s_j, l_j, src_hi, lane);
            } else {
                for (int k = lane; k < l_j; k += 32) dst[o_j + k] = src[s_j + k];
            }
        }
        done += __shfl_sync(kFull, incl, 31);
    }
}

// strip + concat, last kernel: every response's kept runs (and the separator in front of it) go to their final place.
__global__ void __launch_bounds__(kThreads, 5)
strip_copy_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ in, const int64_t* __restrict__ in_off,
                  int64_t n_resp, const int64_t* __restrict__ klen, const int64_t* __restrict__ first,
                  const int64_t* __restrict__ dst, int hide, const uint8_t* __restrict__ sep, int32_t sep_len,
                  uint8_t* __restrict__ out, int64_t* __restrict__ resp_out_len, const UnitDesc* __restrict__ desc,
                  TokSrc ts, RunPool pool) {
    grid_dep_wait();
    __shared__ unsigned long long lastclose[kWarpsPerBlock][kMaxCls];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t total = in_off[n_resp];
    for (int64_t r = (int64_t)blockIdx.x * kWarpsPerBlock + wib; r < n_resp; r += (int64_t)gridDim.x * kWarpsPerBlock) {
        // (prefetching the next response's metadata lines here made the kernel slower: cfg3 1.12 -> 1.17 ms; so did loading
        // all of a unit's metadata - offsets, destination, length, descriptor, runs - side by side up front: 1.38 ms; so
        // did a software pipeline that loads the NEXT response's metadata into registers before this one is copied:
        // 1.09 -> 1.31 ms at 48 registers, 1.17 at 64; and taking the long responses first by ticket (a queue filled by
        // the count kernels): cfg5 -3 %, cfg3 +3.5 %; blocks of 4 or 2 warps instead of 8, so that a block's slot is not
        // held by one long response: cfg5 -1.7 / -2 %, cfg3 -0.3 %, cfg4 +0.6 / +0.8 %)
        const int64_t a = in_off[r], b = in_off[r + 1];
        const int64_t de = dst[r];
        // the descriptor's first 32 bytes (run count, kept length, flags) and its runs fetched beside the offsets, not
        // after them: the kept length then needs no load of klen[r] behind the `de` test, the run count none behind that
        // and the runs none behind the run count - one round trip for a response's metadata instead of four (cfg3
        // 2.006 -> 1.940 ms; the same loads for the NEXT response, or prefetches of its lines, made it slower each time)
        int4 h0 = make_int4(-1, 0, 0, 0), h1 = make_int4(0, 0, 0, 0);
        int rs0 = 0, re0 = 0;
        if (hide && desc) {
            h0 = reinterpret_cast<const int4*>(desc + r)[0];  // nruns, total, held_kept, depth
            h1 = reinterpret_cast<const int4*>(desc + r)[1];  // last_tok_end, rlast, flags, held_len
            rs0 = desc[r].run_s[lane & (kDescRuns - 1)];       // and its runs, one per lane (pooled: the pool index)
            re0 = desc[r].run_e[lane & (kDescRuns - 1)];
        }
        if (de < 0) {  // excluded from the join
            if (resp_out_len && lane == 0) resp_out_len[r] = 0;
            continue;
        }
        const int64_t d = de >> 1;
        if (de & 1)  // the separator between this response and the previous included one of its request
            for (int i = lane; i < sep_len; i += 32) out[d - sep_len + i] = sep[i];
        const int nr = h0.x;
        const int64_t kl = !hide ? (b - a) : nr > 0 ? (int64_t)h0.y : klen[r];
        if (resp_out_len && lane == 0) resp_out_len[r] = kl;
        if (!hide) {
            warp_copy<true>(out + d, in + a, b - a, in + total, lane);
        } else if (kl > 0) {
            if (nr >= 0 && ((uint32_t)h1.z & kDescPooled)) {
                const unsigned long long at = (unsigned long long)(uint32_t)__shfl_sync(kFull, rs0, 0) |
                                              ((unsigned long long)(uint32_t)__shfl_sync(kFull, rs0, 1) << 32);
                gather_pooled(pool.runs + at, nr, in + a, out + d, in + total, lane);
            } else if (nr > 0) {
                int rs = 0, re = 0;
                if (lane < kDescRuns && lane < nr) {
                    rs = rs0;
                    re = re0;
                    if ((h1.z & 2) && lane == nr - 1 && h1.y < re) re = h1.y > rs ? h1.y : rs;  // open-ended final run
                }
                gather_runs<0>(nr, rs, re, h0.z, in + a, out + d, in + total, lane);
            } else if (nr == 0) {
            } else if (nr == -2) {  // 9..32 kept runs: walk the records again and copy from the lanes
                strip_copy_wide(tg, in, total, a, b - a, lastclose[wib], lane, ts, kl, out + d);
            } else {
                strip_copy_slow(tg, in, total, a, b - a, lastclose[wib], lane, first[r], kl, out + d);
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------------------
// kernels: stream feed
// ---------------------------------------------------------------------------------------------------------------
// First kernel after the tokenizer: every stream is walked (tag automaton over its token records), which gives the bytes
// it releases (totals -> scan -> bases), its descriptor (kept runs, hold intervals, carried bytes) and its carried state.
// A warp takes 32 consecutive streams, one per lane (qb_lane.cuh); streams that touch more than lane_steps steps, or do
// not fit the descriptor, are then walked one at a time by the whole warp.  No tokens (QB200_PATH=general): the
// byte-rescanning count of every stream.
//
// SCAN: the prefix sum over the streams' released bytes (their output bases) is done here as well, single pass: a block
// takes 256 consecutive streams by ticket, scans its totals (warp shuffles + shared memory) and gets its base by
// decoupled look-back over the blocks before it (lookback_exclusive, qb_scan.cuh; tickets give forward progress).
// scan.state / scan.ticket are zeroed by the tokenizer kernel, which runs just before on the same stream.
struct CountScan {
    int64_t* bases;                 // out: bases[s]
    unsigned long long* state;      // one look-back word per block of 256 streams
    unsigned int* ticket;
    const int64_t* init_arr;        // chained call: the first base is init_arr[*init_idx] (0 when null)
    const int64_t* init_idx;
};
// (two or three blocks per SM - 128 / 80 registers, no spills - change nothing: cfg2 46.2 / 47.7 us against 48.3)
template <bool SCAN>
__global__ void __launch_bounds__(kThreads, 4)
stream_count_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ bytes,
                    const int64_t* __restrict__ chunk_off, int64_t n_chunks, const int64_t* __restrict__ stream_off,
                    int64_t n_streams, const StreamState* st_in, StreamState* st_out, uint32_t flags,
                    int64_t* __restrict__ totals, TokSrc ts, StreamDesc* __restrict__ desc, int32_t* __restrict__ sid,
                    int lane_steps, CountScan scan) {
    grid_dep_wait();
    __shared__ int64_t s_wsum[kWarpsPerBlock];
    __shared__ int64_t s_excl, s_ticket;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t total = chunk_off[n_chunks];
    const int64_t cb = stream_off[0];
    const int64_t ngroups = (n_streams + 31) >> 5;
    int64_t init = 0;
    if (SCAN && scan.init_arr) init = scan.init_arr[scan.init_idx ? *scan.init_idx : 0];
    for (int64_t it = 0;; it++) {
        int64_t t = (int64_t)blockIdx.x + it * gridDim.x;
        if (SCAN) {
            if (threadIdx.x == 0) s_ticket = (int64_t)atomicAdd(scan.ticket, 1u);
            __syncthreads();
            t = s_ticket;
        }
        if (t * kWarpsPerBlock >= ngroups) break;
        const int64_t g = t * kWarpsPerBlock + wib;
        const int64_t s = (g << 5) + lane;
        const bool have = s < n_streams;
        int64_t c0 = 0, c1 = 0, tot = 0;
        bool done = false;
        if (have) {
            c0 = stream_off[s];
            c1 = stream_off[s + 1];
            if (ts.cnt) {
                bool small = true;
                if (c1 > c0) {
                    const int64_t a = chunk_off[c0], b = chunk_off[c1];
                    small = b <= a || ((b - 1) >> 10) - (a >> 10) < lane_steps;
                }
                if (small)
                    done = stream_walk_lane(tg, bytes, chunk_off + c0, c1 - c0, flags, ts, st_in ? st_in + s : nullptr,
                                            st_out ? st_out + s : nullptr, desc + s, tot);
            }
        }
        __syncwarp();
        unsigned todo = __ballot_sync(kFull, have && !done);
        while (todo) {  // one warp per stream
            const int j = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t sj = (g << 5) + j;
            const int64_t d0 = __shfl_sync(kFull, c0, j), d1 = __shfl_sync(kFull, c1, j);
            const StreamResult r = stream_count_unit(tg, bytes, total, chunk_off + d0, d1 - d0, lane, flags, ts, st_in ? st_in + sj : nullptr,
                                                     st_out ? st_out + sj : nullptr, desc ? desc + sj : nullptr);
            if (lane == j) tot = r.total_out;
            __syncwarp();
        }
        if (!SCAN) {
            if (have) totals[s] = tot;
        } else {
            int64_t incl = tot;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int64_t v = __shfl_up_sync(kFull, incl, d);
                if (lane >= d) incl += v;
            }
            if (lane == 31) s_wsum[wib] = incl;
            __syncthreads();
            int64_t before = 0, block_tot = 0;
#pragma unroll
            for (int w = 0; w < kWarpsPerBlock; w++) {
                const int64_t v = s_wsum[w];
                if (w < wib) before += v;
                block_tot += v;
            }
            if (wib == 0) {
                const int64_t excl = lookback_exclusive(scan.state, t, block_tot, lane);
                if (lane == 0) s_excl = excl;
            }
            __syncthreads();
            if (have) scan.bases[s] = init + s_excl + before + incl - tot;
        }
        if (sid && have && c1 > c0) {
            // tile -> stream map for the offsets kernel: the stream that owns the first boundary of every tile of
            // kOffsetsTile boundaries (boundary cb + 1 + i ends chunk cb + i).  A lane marks the tiles that start in its chunks.
            for (int64_t w = (c0 - cb + kOffsetsTile - 1) / kOffsetsTile; cb + w * kOffsetsTile < c1; w++) sid[w] = (int32_t)s;
        }
    }
}

// Byte-rescanning copy pass (QB200_PATH=general and the allocation-free entry point): warp per stream.
__global__ void __launch_bounds__(kThreads, 4)
stream_copy_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ bytes,
                   const int64_t* __restrict__ chunk_off, int64_t n_chunks, const int64_t* __restrict__ stream_off,
                   int64_t n_streams, const StreamState* st_in, StreamState* st_out, uint32_t flags,
                   const int64_t* __restrict__ bases, uint8_t* __restrict__ out, int64_t* __restrict__ out_chunk_off) {
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t total = chunk_off[n_chunks];
    for (int64_t s = (int64_t)blockIdx.x * kWarpsPerBlock + wib; s < n_streams; s += (int64_t)gridDim.x * kWarpsPerBlock) {
        const int64_t c0 = stream_off[s], c1 = stream_off[s + 1];
        // state_in may alias state_out: the walker reads its input state before writing the output state
        StreamState si;
        if (st_in) si = st_in[s];
        stream_copy_slow(tg, bytes, total, chunk_off + c0, c1 - c0, lane, flags, st_in ? &si : nullptr,
                         st_out ? st_out + s : nullptr, bases[s], out, out_chunk_off, c0);
        __syncwarp();
    }
}

// Per-chunk output offsets, a data-parallel map over the feed boundaries: out_chunk_off[k] = bases[s] + (bytes stream s
// has released by the end of chunk k-1).  Everything sequential was settled by the count kernel (descriptor: kept runs,
// hold intervals, carried bytes); what is left per boundary is a clamp-sum over at most kDescRuns runs and kDescHolds
// interval tests.  A thread takes four consecutive boundaries (their loads are issued together; they nearly always share
// one descriptor), threads are independent: no per-stream chain of dependent loads.
template <int PER>
__global__ void __launch_bounds__(256)
stream_offsets_kernel(const int64_t* __restrict__ chunk_off, const int64_t* __restrict__ stream_off, int64_t n_streams,
                      const int64_t* __restrict__ bases, int64_t* __restrict__ out_chunk_off, const StreamDesc* __restrict__ desc,
                      const int32_t* __restrict__ sid) {
    grid_dep_wait();
    // PER boundaries per thread; a warp takes 32 * PER consecutive ones, lane-interleaved (coalesced)
    const int64_t cb = stream_off[0], nb = stream_off[n_streams] - cb;  // boundary cb + 1 + i ends chunk cb + i, i < nb
    if (blockIdx.x == 0 && threadIdx.x == 0) out_chunk_off[cb] = bases[0];
    const int64_t* co = chunk_off + cb + 1;
    int64_t* oo = out_chunk_off + cb + 1;
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i0 = warp * (32 * PER) + lane; i0 < nb; i0 += nwarps * (32 * PER)) {
        int sv[PER];
        int64_t pos[PER];
#pragma unroll
        for (int j = 0; j < PER; j++) {
            const int64_t i = i0 + 32 * j;
            pos[j] = i < nb ? co[i] : 0;
        }
        // which stream owns each boundary: start from the owner of the tile's first boundary (written by the count
        // kernel) and step over the stream ends in between (a tile usually lies inside one or two streams)
        int sc = sid[(i0 - lane) / (32 * PER)];
        int64_t send = stream_off[sc + 1];  // first chunk of the next stream (kept in a register: one load per stream crossed)
#pragma unroll
        for (int j = 0; j < PER; j++) {
            const int64_t i = i0 + 32 * j;
            sv[j] = -1;
            if (i < nb) {
                const int64_t k = cb + i;  // chunk index
                int steps = 0;
                while (k >= send) {
                    sc++;
                    send = stream_off[sc + 1];
                    if (++steps == 16) {  // many short or empty streams: binary search for the last stream starting at or before k
                        int64_t lo = sc, hi = n_streams - 1;
                        while (lo < hi) {
                            const int64_t mid = (lo + hi + 1) >> 1;
                            if (stream_off[mid] <= k) lo = mid;
                            else hi = mid - 1;
                        }
                        sc = (int)lo;
                        send = stream_off[sc + 1];
                        break;
                    }
                }
                sv[j] = sc;
            }
        }
        int prev = -1, nruns = 0, held_kept = 0;
        int64_t ustart = 0, base = 0;
        int4 s0, s1, l0, l1, hx, he;
#pragma unroll
        for (int j = 0; j < PER; j++) {
            if (sv[j] < 0) continue;
            if (sv[j] != prev) {
                prev = sv[j];
                const StreamDesc* d = desc + prev;
                nruns = d->nruns;
                const int4 hd = *reinterpret_cast<const int4*>(&d->last_tok_end);  // last_tok_end, rlast, flags, held_len
                held_kept = (hd.z & 4) ? hd.w : 0;
                ustart = d->ustart;
                base = bases[prev];
                s0 = *reinterpret_cast<const int4*>(d->run_s), s1 = *reinterpret_cast<const int4*>(d->run_s + 4);
                l0 = *reinterpret_cast<const int4*>(d->run_e), l1 = *reinterpret_cast<const int4*>(d->run_e + 4);
                l0.x -= s0.x, l0.y -= s0.y, l0.z -= s0.z, l0.w -= s0.w, l1.x -= s1.x, l1.y -= s1.y, l1.z -= s1.z, l1.w -= s1.w;
                hx = *reinterpret_cast<const int4*>(d->hold_x), he = *reinterpret_cast<const int4*>(d->hold_e);
            }
            if (nruns < 0) continue;  // redone by the byte-rescanning walker, which writes its own offsets
            const int c = (int)(pos[j] - ustart);
            int P = held_kept;  // unused run slots are [0, 0): they add nothing
            P += min(max(c - s0.x, 0), l0.x) + min(max(c - s0.y, 0), l0.y) + min(max(c - s0.z, 0), l0.z) + min(max(c - s0.w, 0), l0.w);
            if (nruns > 4)
                P += min(max(c - s1.x, 0), l1.x) + min(max(c - s1.y, 0), l1.y) + min(max(c - s1.z, 0), l1.z) + min(max(c - s1.w, 0), l1.w);
            // a boundary inside a hold interval releases only the bytes before the held '<' (everything in between is text)
            if (hx.x != INT32_MAX) {
                if (c > hx.x && c <= he.x) P -= c - hx.x;
                if (c > hx.y && c <= he.y) P -= c - hx.y;
                if (c > hx.z && c <= he.z) P -= c - hx.z;
                if (c > hx.w && c <= he.w) P -= c - hx.w;
            }
            oo[i0 + 32 * j] = base + P;
        }
    }
}

// Last kernel of a stream feed: per stream, the kept runs and the released carried bytes go to their final place; all it
// reads besides the text is the stream's descriptor.  Units that did not fit the descriptor are redone here, offsets and
// carried state included, by the byte-rescanning walker.
__global__ void __launch_bounds__(kThreads, 5)
stream_finish_kernel(const __grid_constant__ DevTags tg, const uint8_t* __restrict__ bytes,
                     const int64_t* __restrict__ chunk_off, int64_t n_chunks, const int64_t* __restrict__ stream_off,
                     int64_t n_streams, const StreamState* st_in, StreamState* st_out, uint32_t flags,
                     const int64_t* __restrict__ bases, uint8_t* __restrict__ out, int64_t* __restrict__ out_chunk_off,
                     const StreamDesc* __restrict__ desc) {
    grid_dep_wait();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t total = chunk_off[n_chunks];
    for (int64_t s = (int64_t)blockIdx.x * kWarpsPerBlock + wib; s < n_streams; s += (int64_t)gridDim.x * kWarpsPerBlock) {
        const StreamDesc* d = desc + s;
        // everything the copy needs from the descriptor in one round trip (see strip_copy_kernel)
        const int4 h0 = reinterpret_cast<const int4*>(d)[0];  // nruns, total, held_kept, depth
        const int4 h1 = reinterpret_cast<const int4*>(d)[1];  // last_tok_end, rlast, flags, held_len
        const int rs0 = d->run_s[lane & (kDescRuns - 1)], re0 = d->run_e[lane & (kDescRuns - 1)];
        const int64_t ustart = d->ustart;
        const int nruns = h0.x;
        const int64_t base = bases[s];
        if (nruns < 0) {
            const int64_t c0 = stream_off[s], c1 = stream_off[s + 1];
            StreamState si;  // state_in may alias state_out: the input state is read before the output state is written
            if (st_in) si = st_in[s];
            stream_copy_slow(tg, bytes, total, chunk_off + c0, c1 - c0, lane, flags, st_in ? &si : nullptr,
                             st_out ? st_out + s : nullptr, base, out, out_chunk_off, c0);
        } else {
            const int hk = h0.z;
            if (lane < hk) out[base + lane] = d->held[lane];
            if (nruns > 0) {
                int rs = 0, re = 0;
                if (lane < kDescRuns && lane < nruns) {
                    rs = rs0;
                    re = re0;
                    if ((h1.z & 2) && lane == nruns - 1 && h1.y < re) re = h1.y > rs ? h1.y : rs;  // open-ended final run
                }
                gather_runs<0>(nruns, rs, re, hk, bytes + ustart, out + base, bytes + total, lane);
            }
        }
        __syncwarp();
    }
}

__global__ void stream_flush_kernel(StreamState* st, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) st[i].held_len = 0;
}

__global__ void set_i64_kernel(int64_t* p, int64_t v) { *p = v; }

// 32-bit lengths <-> 64-bit offsets (what a host ships over PCIe instead of 8 bytes per chunk)
__global__ void widen_i32_kernel(const int32_t* __restrict__ v, int64_t n, int64_t* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = v[i];
}
__global__ void lengths_i32_kernel(const int64_t* __restrict__ off, int64_t n, int32_t* __restrict__ len) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        len[i] = (int32_t)(off[i + 1] - off[i]);
}

// Launch with programmatic stream serialization (see grid_dep_wait): the next kernel's blocks are scheduled while the
// previous kernel's last blocks finish, which takes the ~2 us launch gap between the four or five kernels of a call
// off the critical path.  QB200_PDL=0 launches the ordinary way.
static bool pdl_on() {
    static const bool on = !(getenv("QB200_PDL") && atoi(getenv("QB200_PDL")) == 0);
    return on;
}
template <class... KArgs, class... Args>
static void launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_on() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
extern "C" {

const char* qb200_last_error(void) { return g_err; }

int64_t qb200_launch_count(void) { return g_launches.load(); }

int qb200_profile_begin(void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    {
        std::lock_guard<std::mutex> lk(g_prof_mu);
        for (auto& m : g_prof) cudaEventDestroy(m.ev);
        g_prof.clear();
    }
    g_prof_on = true;
    prof_mark("(begin)", stream);
    return QB200_OK;
}

int32_t qb200_profile_end(const char** names, float* ms, int32_t max_entries) {
    g_prof_on = false;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    int32_t n = 0;
    if (!g_prof.empty()) CUDA_TRY(cudaEventSynchronize(g_prof.back().ev));
    for (size_t i = 1; i < g_prof.size(); i++) {
        if (n >= max_entries) break;
        float t = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&t, g_prof[i - 1].ev, g_prof[i].ev));
        names[n] = g_prof[i].name;
        ms[n] = t;
        n++;
    }
    for (auto& m : g_prof) cudaEventDestroy(m.ev);
    g_prof.clear();
    return n;
}

int qb200_tags_create(const uint8_t* names, const int32_t* name_off, int32_t n_tags, qb200_tags** out) {
    if (!out || n_tags < 0 || (n_tags > 0 && (!names || !name_off))) return fail(QB200_EINVAL, "tags_create: bad argument");
    qb200_tags* t = new (std::nothrow) qb200_tags();
    if (!t) return fail(QB200_EINVAL, "tags_create: out of memory");
    DevTags& d = t->dev;
    memset(&d, 0, sizeof(d));
    d.hold_enabled = n_tags > 0;
    int ncls = 0;
    for (int i = 0; i < n_tags; i++) {
        int len = name_off[i + 1] - name_off[i];
        if (len < 0 || len > QB200_MAX_TAG_LEN) { delete t; return fail(QB200_ETAGS, "tag name longer than QB200_MAX_TAG_LEN"); }
        uint8_t low[kNameStride];
        int worst = 2, nvar = 0;
        for (int k = 0; k < len; k++) {
            uint8_t c = names[name_off[i] + k];
            if (c < 0x20 || c > 0x7E || strchr(".^$*+?{}[]\\|()<>/", c)) {
                delete t;
                return fail(QB200_ETAGS, "tag name must be printable ASCII without regex metacharacters or '<' '>' '/'");
            }
            if (c >= 'A' && c <= 'Z') c = (uint8_t)(c + 32);
            low[k] = c;
            worst += c == 'k' ? 3 : (c == 'i' || c == 's') ? 2 : 1;
            nvar += (c == 'i' || c == 's');
        }
        if (worst > QB200_HELD_MAX) { delete t; return fail(QB200_ETAGS, "tag name too long: a partial tag would exceed QB200_HELD_MAX bytes"); }
        bool dup = false;
        for (int j = 0; j < d.n; j++)
            if (d.len[j] == len && memcmp(d.name[j], low, (size_t)len) == 0) dup = true;
        if (dup) continue;
        if (d.n == QB200_MAX_TAGS) { delete t; return fail(QB200_ETAGS, "more than QB200_MAX_TAGS distinct tag names"); }
        if (ncls + (1 << nvar) > kMaxCls) { delete t; return fail(QB200_ETAGS, "too many backreference classes (i/s letters)"); }
        memcpy(d.name[d.n], low, (size_t)len);
        d.len[d.n] = (uint8_t)len;
        d.cls_base[d.n] = (uint8_t)ncls;
        ncls += 1 << nvar;
        d.n++;
    }
    if (d.n == 0) {  // "<()>": one empty name
        d.n = 1;
        d.len[0] = 0;
        d.cls_base[0] = 0;
        ncls = 1;
    }
    d.ncls = ncls;
    // fast-matcher tables
    auto set_ok = [&](unsigned c) { d.first_ok[c >> 5] |= 1u << (c & 31); };
    set_ok('/');
    for (int i = 0; i < d.n; i++) {
        uint8_t pat[24], orm[24], lenm[24];
        memset(pat, 0, 24);
        memset(orm, 0, 24);
        memset(lenm, 0, 24);
        const int L = d.len[i];
        for (int k = 0; k < L; k++) {
            const uint8_t c = d.name[i][k];
            pat[k] = c;
            orm[k] = (c >= 'a' && c <= 'z') ? 0x20 : 0;
            lenm[k] = 0xFF;
        }
        pat[L] = '>';
        lenm[L] = 0xFF;
        for (int k = 0; k < 6; k++) {
            memcpy(&d.pk[i][k], pat + 4 * k, 4);
            memcpy(&d.om[i][k], orm + 4 * k, 4);
            memcpy(&d.lm[i][k], lenm + 4 * k, 4);
        }
        const uint8_t f = L ? d.name[i][0] : (uint8_t)'>';
        set_ok(f);
        if (f >= 'a' && f <= 'z') set_ok(f - 32);
        if (f == 'i') set_ok(0xC4);
        if (f == 'k') set_ok(0xE2);
        if (f == 's') set_ok(0xC5);
    }
    *out = t;
    return QB200_OK;
}

void qb200_tags_destroy(qb200_tags* tags) { delete tags; }

// QB200_PATH selects the implementation (default "token"): "token" = tokenizer pass + record-driven kernels,
// "general" = walkers that rescan the bytes.
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

enum { kPathGeneral = 0, kPathToken = 2 };
static int path_choice() {
    const char* e = getenv("QB200_PATH");
    if (e && strcmp(e, "general") == 0) return kPathGeneral;
    return kPathToken;
}
static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

static int64_t steps_for(int64_t byte_bound) { return (byte_bound < 0 ? 0 : byte_bound) / 1024 + 2; }
static size_t tok_scratch_bytes(int64_t byte_bound) {
    const int64_t st = steps_for(byte_bound);
    return align256((size_t)st) + align256((size_t)st * kRecPerStep * 4) + 256;  // cnt | rec [st][kRecInline] | rec2 [st][the rest]
}

// tokenizer pass over the flat buffer; returns where the records are
static TokSrc launch_tokenizer(const DevTags& tg, const uint8_t* bytes, const int64_t* off, const int64_t* idx, int64_t n,
                               int64_t byte_bound, uint8_t* scratch, cudaStream_t stream, unsigned long long* zero = nullptr,
                               int64_t zero_words = 0) {
    const int64_t st = steps_for(byte_bound);
    TokSrc ts;
    uint8_t* cnt = scratch;
    uint32_t* rec = (uint32_t*)(scratch + align256((size_t)st));
    int64_t blocks = (st / 2 + 8) / 8;  // a warp takes two steps per iteration
    const int64_t cap = 148 * 32;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    launch_pdl(tokenize_kernel, (unsigned)((unsigned)blocks), 256, stream, tg, bytes, off, idx, n, cnt, rec, st, zero, zero_words);
    LAUNCHED("tokenize_kernel");
    ts.cnt = cnt;
    ts.rec = rec;
    ts.rec2 = rec + st * kRecInline;
    ts.nsteps = st;
    return ts;
}

static size_t strip_general_scratch_bytes(int64_t n_resp, int64_t n_req) {
    if (n_resp < 0) n_resp = 0;
    if (n_req < 0) n_req = 0;
    return 3 * align256((size_t)(n_resp + 1) * 8) + align256((size_t)(n_req + 2) * 8) + align256((size_t)(n_req / kScanSlice + 4) * 8) + 256;
}

static size_t desc_scratch_bytes(int64_t n_units) { return align256((size_t)(n_units < 0 ? 0 : n_units) * sizeof(UnitDesc) + 96); }
static size_t stream_desc_scratch_bytes(int64_t n_units) { return align256((size_t)(n_units < 0 ? 0 : n_units) * sizeof(StreamDesc) + 96); }

// run pool: one 8-byte slot per 128 bytes of input bound (a kept run plus the block it follows is rarely shorter), plus the
// allocation counter; units that do not get their slots fall back to the warp walkers
static int64_t pool_slots(int64_t byte_bound) { return (byte_bound < 0 ? 0 : byte_bound) / 128 + 64; }
static size_t pool_scratch_bytes(int64_t byte_bound) { return align256((size_t)pool_slots(byte_bound) * 8) + 256; }

size_t qb200_strip_concat_scratch_bytes(int64_t n_resp, int64_t n_req, int64_t out_capacity) {
    return strip_general_scratch_bytes(n_resp, n_req) + desc_scratch_bytes(n_resp) + tok_scratch_bytes(out_capacity) +
           pool_scratch_bytes(out_capacity) + 256;
}

// per-device facts, cached per device ordinal (one process may drive several GPUs)
static int sm_count() {
    static std::atomic<int> cache[64];
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    n = cache[dev].load(std::memory_order_relaxed);
    if (!n) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

static int grid_for(int64_t units) {
    int64_t g = (units + kWarpsPerBlock - 1) / kWarpsPerBlock;
    int64_t cap = 148 * 64;  // grid-stride beyond this
    if (const char* e = getenv("QB200_GRID_CAP")) {  // test knob: a small grid makes every block loop (and take many tickets)
        const int64_t v = atoll(e);
        if (v > 0) cap = v;
    }
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

int qb200_strip_concat_batch(const qb200_tags* tags, const uint8_t* in_bytes, const int64_t* in_off, int64_t n_resp,
                             const int64_t* req_off, int64_t n_req, const uint8_t* sep, int32_t sep_len, uint32_t flags,
                             uint8_t* out_bytes, int64_t out_capacity, int64_t* out_off, int64_t* resp_out_len,
                             void* scratch, size_t scratch_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!tags || !in_off || !out_off || n_resp < 0 || n_req < 0 || sep_len < 0 || (sep_len > 0 && !sep))
        return fail(QB200_EINVAL, "strip_concat_batch: bad argument");
    if (!req_off && n_req != n_resp) return fail(QB200_EINVAL, "strip_concat_batch: req_off == NULL needs n_req == n_resp");
    if (out_capacity < 0) return fail(QB200_EINVAL, "strip_concat_batch: negative out_capacity");
    if (((uintptr_t)in_bytes & 15) || ((uintptr_t)out_bytes & 15))
        return fail(QB200_EINVAL, "strip_concat_batch: in_bytes/out_bytes must be 16-byte aligned");
    if (scratch_bytes < qb200_strip_concat_scratch_bytes(n_resp, n_req, out_capacity) || !scratch)
        return fail(QB200_ESCRATCH, "strip_concat_batch: scratch too small (see qb200_strip_concat_scratch_bytes)");
    int dev_count = 0;
    CUDA_TRY(cudaGetDeviceCount(&dev_count));
    const bool chain = (flags & QB200_CHAIN) != 0;  // out_off[0] already holds where this call's output starts
    if (n_req == 0) {
        if (chain) return QB200_OK;
        set_i64_kernel<<<1, 1, 0, stream>>>(out_off, 0);
        LAUNCHED("set_i64_kernel");
        CUDA_TRY(cudaGetLastError());
        return QB200_OK;
    }
    uint8_t* sp = (uint8_t*)scratch;
    int64_t* klen = (int64_t*)sp;
    sp += align256((size_t)(n_resp + 1) * 8);
    int64_t* first = (int64_t*)sp;
    sp += align256((size_t)(n_resp + 1) * 8);
    int64_t* dst = (int64_t*)sp;
    sp += align256((size_t)(n_resp + 1) * 8);
    int64_t* reqlen = (int64_t*)sp;
    const int hide = (flags & QB200_HIDE) ? 1 : 0;
    const int sj = (flags & QB200_STREAM_JOIN) ? 1 : 0;

    TokSrc ts;
    UnitDesc* desc = (UnitDesc*)((uint8_t*)scratch + strip_general_scratch_bytes(n_resp, n_req));
    uint8_t* tok_scratch = (uint8_t*)scratch + strip_general_scratch_bytes(n_resp, n_req) + desc_scratch_bytes(n_resp);
    const int lane_steps = env_int("QB200_LANE_STEPS", 40);  // longer responses go to the queue; 0: warp per response only
    const int path = path_choice();
    // words that must start at zero: the look-back words + ticket of request_offsets_kernel, then the run pool's counter
    // (they live in the `reqlen` part of the scratch; the tokenizer kernel clears them, else a memset)
    unsigned long long* scan_words = (unsigned long long*)reqlen;
    const int64_t nblk_req = (n_req + 255) >> 8;
    RunPool pool;
    pool.top = scan_words + nblk_req + 2;
    LongQueue lq;
    lq.items = dst;
    lq.count = scan_words + nblk_req + 3;
    if (hide && n_resp > 0 && path != kPathGeneral) {
        uint8_t* pp = tok_scratch + tok_scratch_bytes(out_capacity);
        pool.runs = (int2*)(pp + 256);
        pool.cap = pool_slots(out_capacity);
        ts = launch_tokenizer(tags->dev, in_bytes, in_off, nullptr, n_resp, out_capacity, tok_scratch, stream, scan_words, nblk_req + 4);
    } else {
        CUDA_TRY(cudaMemsetAsync(scan_words, 0, (size_t)(nblk_req + 4) * 8, stream));
    }
    if (hide && n_resp > 0) {
        launch_pdl(strip_count_kernel, (unsigned)(grid_for((n_resp + 31) / 32)), kThreads, stream, tags->dev, in_bytes, in_off, n_resp, klen, first, ts,
                                                                                 ts.cnt ? desc : nullptr, lane_steps, pool, lq);
        LAUNCHED("strip_count_kernel");
        if (ts.cnt) {
            launch_pdl(strip_count_long_kernel, (unsigned)(4 * sm_count()), kThreads, stream, tags->dev, in_bytes, in_off, n_resp, klen, first, ts, desc, pool,
                                                                            lq, env_int("QB200_SCANWALK", 1));
            LAUNCHED("strip_count_long_kernel");
        }
    }
    {
        ReqScan rs;
        rs.state = scan_words;
        rs.ticket = (unsigned int*)(scan_words + nblk_req + 1);
        rs.chain = chain ? 1 : 0;
        int64_t g = nblk_req < 148 * 8 ? nblk_req : 148 * 8;
        launch_pdl(request_offsets_kernel, (unsigned)((unsigned)(g < 1 ? 1 : g)), 256, stream, in_off, req_off, n_req, klen, hide, sj, sep_len, dst, out_off, rs);
        LAUNCHED("request_offsets_kernel");
    }
    if (n_resp > 0) {
        launch_pdl(strip_copy_kernel, (unsigned)(grid_for(n_resp)), kThreads, stream, tags->dev, in_bytes, in_off, n_resp, klen, first, dst,
                                                                    hide, sep, sep_len, out_bytes, resp_out_len,
                                                                    (hide && ts.cnt) ? desc : nullptr, ts, pool);
        LAUNCHED("strip_copy_kernel");
    }
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

static size_t stream_general_scratch_bytes(int64_t n_streams) {
    if (n_streams < 0) n_streams = 0;
    return 2 * align256((size_t)(n_streams + 2) * 8) + align256((size_t)(n_streams / kScanSlice + 4) * 8) + 256;
}

static size_t sid_scratch_bytes(int64_t n_chunks) { return align256((size_t)((n_chunks < 0 ? 0 : n_chunks) / kOffsetsTile + 32) * 4); }

size_t qb200_stream_feed_scratch_bytes(int64_t n_chunks, int64_t n_streams, int64_t out_capacity) {
    return stream_general_scratch_bytes(n_streams) + stream_desc_scratch_bytes(n_streams) + tok_scratch_bytes(out_capacity) +
           sid_scratch_bytes(n_chunks) + 256;
}

static int stream_feed_check(const qb200_tags* tags, const uint8_t* chunk_bytes, const int64_t* chunk_off, int64_t n_chunks,
                             const int64_t* stream_off, int64_t n_streams, uint8_t* out_bytes, int64_t out_capacity,
                             int64_t* out_chunk_off) {
    if (!tags || !chunk_off || !stream_off || !out_chunk_off || n_chunks < 0 || n_streams < 0 || out_capacity < 0)
        return fail(QB200_EINVAL, "stream_feed_batch: bad argument");
    if (((uintptr_t)chunk_bytes & 15) || ((uintptr_t)out_bytes & 15))
        return fail(QB200_EINVAL, "stream_feed_batch: chunk_bytes/out_bytes must be 16-byte aligned");
    int dev_count = 0;
    CUDA_TRY(cudaGetDeviceCount(&dev_count));
    return QB200_OK;
}

// General path: count kernel -> scan over streams -> copy kernel.  Any unit size, input read twice.
static int stream_feed_general(const qb200_tags* tags, const uint8_t* chunk_bytes, const int64_t* chunk_off, int64_t n_chunks,
                               const int64_t* stream_off, int64_t n_streams, const qb200_stream_state* state_in,
                               qb200_stream_state* state_out, uint32_t flags, uint8_t* out_bytes, int64_t out_capacity,
                               int64_t* out_chunk_off, void* scratch, size_t scratch_bytes, cudaStream_t stream,
                               bool use_tokens = false) {
    int rc = stream_feed_check(tags, chunk_bytes, chunk_off, n_chunks, stream_off, n_streams, out_bytes, out_capacity,
                               out_chunk_off);
    if (rc != QB200_OK) return rc;
    if (scratch_bytes < stream_general_scratch_bytes(n_streams) || !scratch)
        return fail(QB200_ESCRATCH, "stream_feed_batch: scratch too small");
    const bool chain = (flags & QB200_CHAIN) != 0;
    if (n_streams == 0) {
        if (!chain) {
            set_i64_kernel<<<1, 1, 0, stream>>>(out_chunk_off, 0);
            LAUNCHED("set_i64_kernel");
        }
        CUDA_TRY(cudaGetLastError());
        return QB200_OK;
    }
    uint8_t* sp = (uint8_t*)scratch;
    int64_t* totals = (int64_t*)sp;
    sp += align256((size_t)(n_streams + 2) * 8);
    int64_t* bases = (int64_t*)sp;
    const StreamState* si = reinterpret_cast<const StreamState*>(state_in);
    StreamState* so = reinterpret_cast<StreamState*>(state_out);
    TokSrc ts;
    StreamDesc* desc = nullptr;
    size_t sid_offset = 0;
    unsigned long long* scan_state = nullptr;
    unsigned int* scan_ticket = nullptr;
    if (use_tokens) {  // descriptors + records (+ the chunk -> stream map) live behind the general scratch region
        uint8_t* tsp = (uint8_t*)scratch + stream_general_scratch_bytes(n_streams);
        desc = (StreamDesc*)tsp;
        scan_state = (unsigned long long*)totals;  // (n_streams / 256 + 2) words + the ticket fit in the totals array
        const int64_t nblk = (n_streams + 255) / 256 + 1;
        scan_ticket = (unsigned int*)(scan_state + nblk);
        ts = launch_tokenizer(tags->dev, chunk_bytes, chunk_off, stream_off, n_streams, out_capacity,
                              tsp + stream_desc_scratch_bytes(n_streams), stream, scan_state, nblk + 1);
        sid_offset = (size_t)(tsp - (uint8_t*)scratch) + stream_desc_scratch_bytes(n_streams) + tok_scratch_bytes(out_capacity);
    }
    int32_t* sid = use_tokens ? (int32_t*)((uint8_t*)scratch + sid_offset) : nullptr;
    const int lane_steps = env_int("QB200_LANE_STEPS", 40);  // 0: warp per stream only
    // chained call: this call's output starts where out_chunk_off[first chunk of the call] says (written by the call before)
    if (use_tokens) {  // prefix sum inside the count kernel (look-back words live where the totals would)
        CountScan cs{bases, scan_state, scan_ticket, chain ? out_chunk_off : nullptr, chain ? stream_off : nullptr};
        launch_pdl(stream_count_kernel<true>, (unsigned)(grid_for((n_streams + 31) / 32)), kThreads, stream, 
            tags->dev, chunk_bytes, chunk_off, n_chunks, stream_off, n_streams, si, so, flags, totals, ts, desc, sid, lane_steps, cs);
        LAUNCHED("stream_count_kernel");
    } else {
        stream_count_kernel<false><<<grid_for((n_streams + 31) / 32), kThreads, 0, stream>>>(
            tags->dev, chunk_bytes, chunk_off, n_chunks, stream_off, n_streams, si, nullptr, flags, totals, ts, desc, sid, lane_steps,
            CountScan{});
        LAUNCHED("stream_count_kernel");
        launch_scan(totals, n_streams, bases, (int64_t*)((uint8_t*)bases + align256((size_t)(n_streams + 2) * 8)), stream,
                    chain ? out_chunk_off : nullptr, chain ? stream_off : nullptr);
    }
    if (use_tokens) {
        // boundaries per thread measured on cfg2: 1 -> 606, 2 -> 588, 4 -> 565, 8 -> 563 us per call
        int64_t nb = (n_chunks + 256 * kOffsetsPer - 1) / (256 * kOffsetsPer);
        if (nb > 148 * 64) nb = 148 * 64;
        if (nb < 1) nb = 1;
        launch_pdl(stream_offsets_kernel<kOffsetsPer>, (unsigned)((unsigned)nb), 256, stream, chunk_off, stream_off, n_streams, bases, out_chunk_off, desc,
                                                                            sid);
        LAUNCHED("stream_offsets_kernel");
        launch_pdl(stream_finish_kernel, (unsigned)(grid_for(n_streams)), kThreads, stream, 
            tags->dev, chunk_bytes, chunk_off, n_chunks, stream_off, n_streams, si, so, flags, bases, out_bytes, out_chunk_off, desc);
        LAUNCHED("stream_finish_kernel");
    } else {
        stream_copy_kernel<<<grid_for(n_streams), kThreads, 0, stream>>>(tags->dev, chunk_bytes, chunk_off, n_chunks, stream_off,
                                                                        n_streams, si, so, flags, bases, out_bytes, out_chunk_off);
        LAUNCHED("stream_copy_kernel");
    }
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

int qb200_stream_feed_batch(const qb200_tags* tags, const uint8_t* chunk_bytes, const int64_t* chunk_off, int64_t n_chunks,
                            const int64_t* stream_off, int64_t n_streams, const qb200_stream_state* state_in,
                            qb200_stream_state* state_out, uint32_t flags, uint8_t* out_bytes, int64_t out_capacity,
                            int64_t* out_chunk_off, void* scratch, size_t scratch_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool tok = path_choice() != kPathGeneral && n_streams > 0;
    if (tok && (scratch_bytes < qb200_stream_feed_scratch_bytes(n_chunks, n_streams, out_capacity) || !scratch))
        return fail(QB200_ESCRATCH, "stream_feed_batch: scratch too small (see qb200_stream_feed_scratch_bytes)");
    return stream_feed_general(tags, chunk_bytes, chunk_off, n_chunks, stream_off, n_streams, state_in, state_out, flags,
                               out_bytes, out_capacity, out_chunk_off, scratch, scratch_bytes, stream, tok);
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of qb200_stream_feed_batch.
//
// Who needs it: a host that cannot hand over scratch memory at call time.  Examples: a serving process that gave all
// of HBM to its own pools at start-up and has no allocator left to ask; a binding with nothing to allocate with (a
// bare ctypes / cgo / JNI caller that only owns its input and output buffers); a latency-critical loop in which
// cudaMalloc is forbidden because it synchronises the whole device and stalls every other stream.
//
// The scratch (two int64 arrays over the streams: released bytes per stream and their prefix sums) lives in the
// file-scope device array below instead.  Consequences the caller must know:
//   * the batch is limited to QB200_NOALLOC_MAX_STREAMS streams; more is rejected with QB200_EINVAL;
//   * the array costs QB200_NOALLOC_SCRATCH_BYTES of device memory on every GPU that loads this library, until exit;
//   * every call on a device uses the SAME array.  Two calls whose kernels can overlap on the GPU (different streams,
//     different host threads) race on it and corrupt each other's offsets: serialise them.  The array still holds the
//     previous call's values when a call starts; nothing in it has to start at zero, because the counting kernel
//     overwrites every entry the later kernels read before they read it;
//   * the first call on a device resolves the array's device address (set-up, once per device, with that device
//     current).  Make that first call outside any stream capture.
// ---------------------------------------------------------------------------------------------------------------
#define QB200_NOALLOC_SCRATCH_BYTES_ (2 * (((QB200_NOALLOC_MAX_STREAMS + 2) * 8 + 255) / 256 * 256) + 1024)
static __device__ __align__(256) unsigned char g_noalloc_scratch[QB200_NOALLOC_SCRATCH_BYTES_];

static constexpr int kMaxDevices = 64;
static std::atomic<void*> g_noalloc_ptr[kMaxDevices];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: ask the runtime for the
// address, on the device that is current, and cache it per device.
static int noalloc_scratch_ptr(void** out) {
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(QB200_EINVAL, "stream_feed_batch_noalloc: device ordinal out of range");
    void* p = g_noalloc_ptr[dev].load(std::memory_order_acquire);
    if (!p) {
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(QB200_ECUDA, "stream_feed_batch_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        g_noalloc_ptr[dev].store(p, std::memory_order_release);  // same value from every thread: a benign race
    }
    *out = p;
    return QB200_OK;
}

int qb200_stream_feed_batch_noalloc(const qb200_tags* tags, const uint8_t* chunk_bytes, const int64_t* chunk_off,
                                    int64_t n_chunks, const int64_t* stream_off, int64_t n_streams,
                                    const qb200_stream_state* state_in, qb200_stream_state* state_out, uint32_t flags,
                                    uint8_t* out_bytes, int64_t out_capacity, int64_t* out_chunk_off, void* stream_) {
    if (n_streams > QB200_NOALLOC_MAX_STREAMS)
        return fail(QB200_EINVAL,
                    "stream_feed_batch_noalloc: n_streams exceeds QB200_NOALLOC_MAX_STREAMS (65536); split the batch or "
                    "use qb200_stream_feed_batch with caller-owned scratch");
    void* scratch = nullptr;
    int rc = noalloc_scratch_ptr(&scratch);
    if (rc != QB200_OK) return rc;
    // the general three-kernel path (count, scan, copy) with its scratch taken from the static array
    return stream_feed_general(tags, chunk_bytes, chunk_off, n_chunks, stream_off, n_streams, state_in, state_out, flags,
                               out_bytes, out_capacity, out_chunk_off, scratch, (size_t)QB200_NOALLOC_SCRATCH_BYTES_,
                               (cudaStream_t)stream_);
}

size_t qb200_sse_frame_scratch_bytes(int64_t n_items) {
    if (n_items < 0) n_items = 0;
    return align256((size_t)(n_items + 2) * 8) + align256((size_t)(n_items / kScanSlice + 4) * 8) + 256;
}

int qb200_sse_frame_batch(const uint8_t* in_bytes, const int64_t* in_off, int64_t n_items, const int32_t* item_prefix,
                          const uint8_t* prefix_bytes, const int64_t* prefix_off, int32_t n_prefix, const uint8_t* suffix,
                          int32_t suffix_len, uint32_t flags, uint8_t* out_bytes, int64_t out_capacity, int64_t* out_off,
                          void* scratch, size_t scratch_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!in_off || !out_off || !prefix_off || n_items < 0 || n_prefix < 1 || suffix_len < 0 || out_capacity < 0 ||
        (suffix_len > 0 && !suffix))
        return fail(QB200_EINVAL, "sse_frame_batch: bad argument");
    if (scratch_bytes < qb200_sse_frame_scratch_bytes(n_items) || !scratch)
        return fail(QB200_ESCRATCH, "sse_frame_batch: scratch too small");
    int dev_count = 0;
    CUDA_TRY(cudaGetDeviceCount(&dev_count));
    if (n_items == 0) {
        set_i64_kernel<<<1, 1, 0, stream>>>(out_off, 0);
        LAUNCHED("set_i64_kernel");
        CUDA_TRY(cudaGetLastError());
        return QB200_OK;
    }
    int64_t* lens = (int64_t*)scratch;
    int64_t* partial = (int64_t*)((uint8_t*)scratch + align256((size_t)(n_items + 2) * 8));
    const int64_t groups = (n_items + 31) / 32;  // a warp takes 32 consecutive items
    int64_t blocks = (groups + 7) / 8;
    if (blocks > 148 * 32) blocks = 148 * 32;
    sse_count_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in_bytes, in_off, n_items, item_prefix, prefix_off, suffix_len, flags,
                                                          lens);
    LAUNCHED("sse_count_kernel");
    launch_scan(lens, n_items, out_off, partial, stream);
    int64_t wblocks = (groups + kSseWarps - 1) / kSseWarps;
    if (wblocks > 148 * 32) wblocks = 148 * 32;
    sse_write_kernel<<<(unsigned)wblocks, kSseWarps * 32, 0, stream>>>(in_bytes, in_off, n_items, item_prefix, prefix_bytes, prefix_off,
                                                                      n_prefix, suffix, suffix_len, out_off, out_bytes);
    LAUNCHED("sse_write_kernel");
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

size_t qb200_sse_split_scratch_bytes(int64_t n_http) {
    if (n_http < 0) n_http = 0;
    return align256((size_t)(n_http + 2) * 8) + align256((size_t)(n_http / kScanSlice + 4) * 8) + 256;
}

int qb200_sse_split_batch(const uint8_t* http_bytes, const int64_t* http_off, int64_t n_http, int64_t* ev_first, void* scratch,
                          size_t scratch_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!http_off || !ev_first || n_http < 0) return fail(QB200_EINVAL, "sse_split_batch: bad argument");
    if (scratch_bytes < qb200_sse_split_scratch_bytes(n_http) || !scratch) return fail(QB200_ESCRATCH, "sse_split_batch: scratch too small");
    if (n_http == 0) {
        set_i64_kernel<<<1, 1, 0, stream>>>(ev_first, 0);
        LAUNCHED("set_i64_kernel");
        CUDA_TRY(cudaGetLastError());
        return QB200_OK;
    }
    int64_t* counts = (int64_t*)scratch;
    int64_t* partial = (int64_t*)((uint8_t*)scratch + align256((size_t)(n_http + 2) * 8));
    sse_split_count_kernel<<<(unsigned)((n_http + 127) / 128), 128, 0, stream>>>(http_bytes, http_off, n_http, counts);
    LAUNCHED("sse_split_count_kernel");
    launch_scan(counts, n_http, ev_first, partial, stream);
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

size_t qb200_sse_parse_scratch_bytes(int64_t n_events) {
    if (n_events < 0) n_events = 0;
    return 5 * align256((size_t)(n_events + 2) * 8) + align256((size_t)(n_events / kScanSlice + 4) * 8) + 256;
}

int qb200_sse_parse_batch(const uint8_t* http_bytes, const int64_t* http_off, int64_t n_http, const int64_t* ev_first,
                          int64_t n_events, uint8_t* out_bytes, int64_t out_capacity, int64_t* out_off, uint8_t* ev_kind,
                          int64_t* n_host, void* scratch, size_t scratch_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!http_off || !ev_first || !out_off || !n_host || n_http < 0 || n_events < 0 || out_capacity < 0 || (n_events > 0 && !ev_kind))
        return fail(QB200_EINVAL, "sse_parse_batch: bad argument");
    if (scratch_bytes < qb200_sse_parse_scratch_bytes(n_events) || !scratch) return fail(QB200_ESCRATCH, "sse_parse_batch: scratch too small");
    set_i64_kernel<<<1, 1, 0, stream>>>(n_host, 0);
    LAUNCHED("set_i64_kernel");
    if (n_events == 0) {
        set_i64_kernel<<<1, 1, 0, stream>>>(out_off, 0);
        LAUNCHED("set_i64_kernel");
        CUDA_TRY(cudaGetLastError());
        return QB200_OK;
    }
    uint8_t* sp = (uint8_t*)scratch;
    const size_t slot = align256((size_t)(n_events + 2) * 8);
    int64_t* ev_s = (int64_t*)sp;
    int64_t* ev_e = (int64_t*)(sp + slot);
    int64_t* cs = (int64_t*)(sp + 2 * slot);
    int64_t* ce = (int64_t*)(sp + 3 * slot);
    int64_t* lens = (int64_t*)(sp + 4 * slot);
    int64_t* partial = (int64_t*)(sp + 5 * slot);
    if (n_http > 0) {
        sse_split_write_kernel<<<(unsigned)((n_http + 127) / 128), 128, 0, stream>>>(http_bytes, http_off, n_http, ev_first, ev_s, ev_e);
        LAUNCHED("sse_split_write_kernel");
    }
    const unsigned eg = (unsigned)((n_events + 127) / 128);
    sse_parse_kernel<<<eg, 128, 0, stream>>>(http_bytes, ev_s, ev_e, n_events, ev_kind, cs, ce, lens, (unsigned long long*)n_host);
    LAUNCHED("sse_parse_kernel");
    launch_scan(lens, n_events, out_off, partial, stream);
    sse_unescape_kernel<<<eg, 128, 0, stream>>>(http_bytes, n_events, ev_kind, cs, ce, out_off, out_bytes);
    LAUNCHED("sse_unescape_kernel");
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

size_t qb200_assemble_scratch_bytes(int64_t n_resp, int64_t n_req) {
    if (n_resp < 0) n_resp = 0;
    if (n_req < 0) n_req = 0;
    return 2 * align256((size_t)(n_resp + 2) * 8) + align256((size_t)(n_resp + 2) * 4) + align256((size_t)(n_req + 2) * 8) +
           align256((size_t)(n_req / kScanSlice + 4) * 8) + 256;
}

int qb200_assemble_prompts_batch(const uint8_t* resp_bytes, const int64_t* resp_off, int64_t n_resp, const int64_t* req_off,
                                 int64_t n_req, const uint8_t* head_bytes, const int64_t* head_off, const uint8_t* label_bytes,
                                 const int32_t* label_off, int32_t n_labels, const uint8_t* sep, int32_t sep_len,
                                 const uint8_t* tail, int32_t tail_len, uint8_t* out_bytes, int64_t out_capacity, int64_t* out_off,
                                 void* scratch, size_t scratch_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!resp_off || !req_off || !out_off || n_resp < 0 || n_req < 0 || n_labels < 0 || sep_len < 0 || tail_len < 0 ||
        out_capacity < 0 || (n_labels > 0 && (!label_bytes || !label_off)) || (sep_len > 0 && !sep) || (tail_len > 0 && !tail))
        return fail(QB200_EINVAL, "assemble_prompts_batch: bad argument");
    if (scratch_bytes < qb200_assemble_scratch_bytes(n_resp, n_req) || !scratch)
        return fail(QB200_ESCRATCH, "assemble_prompts_batch: scratch too small (see qb200_assemble_scratch_bytes)");
    if (n_req == 0) {
        set_i64_kernel<<<1, 1, 0, stream>>>(out_off, 0);
        LAUNCHED("set_i64_kernel");
        CUDA_TRY(cudaGetLastError());
        return QB200_OK;
    }
    uint8_t* sp = (uint8_t*)scratch;
    int64_t* rel = (int64_t*)sp;
    sp += align256((size_t)(n_resp + 2) * 8);
    int64_t* req_of = (int64_t*)sp;
    sp += align256((size_t)(n_resp + 2) * 8);
    int32_t* idx = (int32_t*)sp;
    sp += align256((size_t)(n_resp + 2) * 4);
    int64_t* reqlen = (int64_t*)sp;
    sp += align256((size_t)(n_req + 2) * 8);
    int64_t* partial = (int64_t*)sp;
    const unsigned rg = (unsigned)((n_req + 255) / 256);
    assemble_layout_kernel<<<rg, 256, 0, stream>>>(resp_off, req_off, n_req, head_off, label_off, n_labels, sep_len, tail_len, rel,
                                                  idx, reqlen);
    LAUNCHED("assemble_layout_kernel");
    assemble_owner_kernel<<<rg, 256, 0, stream>>>(req_off, n_req, req_of);
    LAUNCHED("assemble_owner_kernel");
    launch_scan(reqlen, n_req, out_off, partial, stream);
    assemble_frame_kernel<<<(unsigned)((n_req + 7) / 8), 256, 0, stream>>>(req_off, n_req, head_bytes, head_off, tail, tail_len, out_off,
                                                                          out_bytes);
    LAUNCHED("assemble_frame_kernel");
    if (n_resp > 0) {
        assemble_copy_kernel<<<grid_for(n_resp), 256, 0, stream>>>(resp_bytes, resp_off, n_resp, rel, idx, req_of, label_bytes, label_off,
                                                                  n_labels, sep, sep_len, out_off, out_bytes);
        LAUNCHED("assemble_copy_kernel");
    }
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

size_t qb200_offsets_scratch_bytes(int64_t n) {
    if (n < 0) n = 0;
    return align256((size_t)(n + 2) * 8) + align256((size_t)(n / kScanSlice + 4) * 8) + 256;
}

int qb200_offsets_from_lengths32(const int32_t* len, int64_t n, int64_t* off, void* scratch, size_t scratch_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n < 0 || !off || (n > 0 && !len)) return fail(QB200_EINVAL, "offsets_from_lengths32: bad argument");
    if (n == 0) return QB200_OK;
    if (scratch_bytes < qb200_offsets_scratch_bytes(n) || !scratch) return fail(QB200_ESCRATCH, "offsets_from_lengths32: scratch too small");
    int64_t* wide = (int64_t*)scratch;
    int64_t* partial = (int64_t*)((uint8_t*)scratch + align256((size_t)(n + 2) * 8));
    int64_t g = (n + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    widen_i32_kernel<<<(unsigned)g, 256, 0, stream>>>(len, n, wide);
    LAUNCHED("widen_i32_kernel");
    launch_scan(wide, n, off, partial, stream, off, nullptr);  // continues from the value already in off[0]
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

int qb200_lengths32_from_offsets(const int64_t* off, int64_t n, int32_t* len, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n < 0 || (n > 0 && (!off || !len))) return fail(QB200_EINVAL, "lengths32_from_offsets: bad argument");
    if (n == 0) return QB200_OK;
    int64_t g = (n + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    lengths_i32_kernel<<<(unsigned)g, 256, 0, stream>>>(off, n, len);
    LAUNCHED("lengths_i32_kernel");
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

int qb200_usage_sums_batch(const int64_t* usage, const int64_t* req_off, int64_t n_req, int64_t* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (n_req < 0 || (n_req > 0 && (!usage || !req_off || !out))) return fail(QB200_EINVAL, "usage_sums_batch: bad argument");
    if (n_req == 0) return QB200_OK;
    usage_sums_kernel<<<(unsigned)((n_req + 255) / 256), 256, 0, stream>>>(usage, req_off, n_req, out);
    LAUNCHED("usage_sums_kernel");
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

int qb200_stream_flush_batch(qb200_stream_state* state, int64_t n_streams, void* stream_) {
    if (n_streams < 0 || (n_streams > 0 && !state)) return fail(QB200_EINVAL, "stream_flush_batch: bad argument");
    if (n_streams == 0) return QB200_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    stream_flush_kernel<<<(unsigned)((n_streams + 255) / 256), 256, 0, stream>>>(
        reinterpret_cast<StreamState*>(state), n_streams);
    LAUNCHED("stream_flush_kernel");
    CUDA_TRY(cudaGetLastError());
    return QB200_OK;
}

}  // extern "C"
