// Fused single-pass tile kernels: every input byte is read from HBM once (TMA bulk copy into shared memory), every output
// byte is written once, output offsets come from a decoupled look-back over tiles.
//
// Work assignment: the flat input is cut into 16 KiB windows; tile t owns the units (streams / responses) whose first
// byte lies in window t, so units are never split and a tile's bytes are contiguous (a small pre-pass kernel writes
// every tile's unit / byte / chunk range).  A CTA takes tiles from a ticket counter (tiles start in index order, which
// the look-back needs for forward progress).
//
// Inside a tile (fast path, everything in shared memory):
//   A  all warps, one 1 KiB step each: '<' bitmap (16-byte shared loads + SWAR), tag matching at the candidates
//      against the tag table -> token records in per-step slots (popc ranks by warp scan)
//   B  one warp per unit: sequential automaton over the (sparse) token list -> kept segments, kept bitmap painted
//   C  block scan of kept-byte counts per word; one thread per feed boundary: release boundary + hold-back test
//   D  block scan of per-unit output lengths, tile aggregate, decoupled look-back -> tile output base
//   E  per-chunk offsets, carried state, and the kept segments copied out with 16-byte stores
// A tile that does not fit the fast path's tables (too many units / tokens / segments / chunks, or a unit larger than
// the staging buffer) runs the general warp-per-unit walkers of qb_walk.cuh inside the same kernel instead.
#pragma once
#include "qb_walk.cuh"

namespace qb {

constexpr int kWindow = 16384;            // bytes of input per tile window
constexpr int kStageCap = 32768 + 64;     // staged bytes per tile: a unit <= 16 KiB starting anywhere in the window fits,
                                          // a 32 KiB unit fits when it starts at the window start
constexpr int kTileWarps = 8;
constexpr int kTileThreads = kTileWarps * 32;
constexpr int kMaxSteps = (kStageCap + 1023) / 1024;
constexpr int kStageWords = kMaxSteps * 32;
constexpr int kMaxUnits = 64;
constexpr int kSlots = 8;                 // token slots per 1 KiB step
constexpr int kSegCap = 192;
constexpr int kChunkCap = 576;
constexpr int kSegPiece = 1024;           // long kept runs are split so the copy spreads over warps

// ---------------------------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + TMA bulk copy (cp.async.bulk -> UBLKCP)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// ---------------------------------------------------------------------------------------------------------------
// Decoupled look-back over tiles (single pass prefix sum).  state word = status (2 bits) | value (62 bits).
// Called by one full warp.  Returns the exclusive prefix of `aggregate` over all earlier tiles.
// ---------------------------------------------------------------------------------------------------------------
constexpr unsigned long long kStAgg = 1ull << 62, kStPre = 2ull << 62, kStMask = (1ull << 62) - 1;

__device__ inline int64_t lookback_exclusive(unsigned long long* tile_state, int64_t tile, int64_t aggregate, int lane) {
    if (tile == 0) {
        if (lane == 0) atomicExch(&tile_state[0], kStPre | (unsigned long long)aggregate);
        return 0;
    }
    if (lane == 0) atomicExch(&tile_state[tile], kStAgg | (unsigned long long)aggregate);
    int64_t excl = 0;
    int64_t j = tile - 1;
    for (;;) {
        const int64_t idx = j - lane;
        unsigned long long v = idx >= 0 ? ld_volatile_u64(&tile_state[idx]) : kStPre;
        const unsigned st = (unsigned)(v >> 62);
        const unsigned has_p = __ballot_sync(kFull, st == 2);
        const unsigned inval = __ballot_sync(kFull, st == 0);
        const int firstp = has_p ? __ffs(has_p) - 1 : 32;
        const unsigned need = firstp >= 31 ? 0xffffffffu : ((2u << firstp) - 1u);
        if (inval & need) {  // a predecessor has not published yet: poll again, without hogging the issue slots
            __nanosleep(40);
            continue;
        }
        int64_t val = (lane <= firstp) ? (int64_t)(v & kStMask) : 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(kFull, val, d);
        excl += val;
        if (has_p) break;
        j -= 32;
    }
    if (lane == 0) atomicExch(&tile_state[tile], kStPre | (unsigned long long)(excl + aggregate));
    return excl;
}

// token record: x = tile-relative position of '<';  y = len | close<<8 | cls<<9 | exrel<<16 (0xFF = no exotic form)
struct Tok {
    uint32_t x, y;
};
// kept segment: copy `len` bytes from tile-relative `src` to unit-relative output offset `dst`
struct Seg {
    uint32_t src, len, dst;
    uint16_t unit, final_;
};
// Written by the pre-pass: the units / bytes / items (chunks or requests) a tile owns.  u0p1 == 0: no unit starts here.
struct __align__(16) TileRec {
    int64_t u0p1, u1, b0, b1, c0, c1;
};

struct TileShared {
    uint64_t mbar;
    uint64_t pad_;
    TileRec next_rec;   // prefetched record of the CTA's next tile (cp.async)
    int64_t g0;         // absolute offset of the first staged byte (multiple of 32)
    int64_t u0, u1;     // unit range of this tile
    int64_t c0;         // (streams) absolute index of the tile's first chunk
    int64_t tile_base;  // output offset of the tile
    unsigned long long agg;  // general path: sum of unit totals
    int32_t staged;     // staged bytes
    int32_t nunits, nchunks, nsteps;
    int32_t fast;       // 1 = fast path
    int32_t nseg;
    int32_t overflow;
    int32_t tile;
    int32_t tile_total;
    int32_t step_tok[kMaxSteps + 2];   // tokens recorded per step
    int32_t step_cum[kMaxSteps + 2];   // kept bytes before the step
    int32_t u_beg[kMaxUnits + 1], u_end[kMaxUnits + 1];  // tile-relative byte range
    int32_t u_cend[kMaxUnits + 1];     // cumulative chunk count up to and including the unit (tile-local)
    int32_t u_depth[kMaxUnits], u_last_tok_end[kMaxUnits], u_total[kMaxUnits], u_out_rel[kMaxUnits + 1];
    int32_t u_rlast[kMaxUnits];
    uint8_t u_held_kept[kMaxUnits], u_held_unc[kMaxUnits];
    Tok tok[kMaxSteps * kSlots];
    Seg seg[kSegCap];
    int32_t bnd[kChunkCap];            // feed boundaries: tile-relative position, later (unit<<24 | released bytes before)
    uint32_t ltbm[kStageWords];        // '<' bitmap, one word per 32 bytes
    uint32_t keptbm[kStageWords];      // kept-byte bitmap
    uint16_t wpre[kStageWords];        // kept bytes before the word inside its step
    __align__(128) uint8_t data[kStageCap + 64];
};

// exclusive scan of a[0..n), n <= 64, by one full warp; returns the total
__device__ inline int warp_scan64(int32_t* a, int n, int lane) {
    const int v0 = lane < n ? a[lane] : 0, v1 = lane + 32 < n ? a[lane + 32] : 0;
    int i0 = v0, i1 = v1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t0 = __shfl_up_sync(kFull, i0, d), t1 = __shfl_up_sync(kFull, i1, d);
        if (lane >= d) {
            i0 += t0;
            i1 += t1;
        }
    }
    const int tot0 = __shfl_sync(kFull, i0, 31), tot1 = __shfl_sync(kFull, i1, 31);
    if (lane < n) a[lane] = i0 - v0;
    if (lane + 32 < n) a[lane + 32] = tot0 + i1 - v1;
    return tot0 + tot1;
}

// ---------------------------------------------------------------------------------------------------------------
// pre-pass: which units does every 16 KiB window own?  One thread per unit; exactly one writer per field.
// ---------------------------------------------------------------------------------------------------------------
// streams: unit start = chunk_off[stream_off[u]], item index = chunk index
__global__ void stream_tile_map_kernel(const int64_t* __restrict__ chunk_off, int64_t n_chunks,
                                       const int64_t* __restrict__ stream_off, int64_t n_streams, TileRec* __restrict__ rec,
                                       int64_t max_tiles) {
    const int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_streams) return;
    const int64_t c = stream_off[u], cn = stream_off[u + 1];
    const int64_t b = chunk_off[c], bn = chunk_off[cn];
    const int64_t t = b / kWindow;
    if (t >= max_tiles) return;
    const int64_t tp = u > 0 ? chunk_off[stream_off[u - 1]] / kWindow : -1;
    const int64_t tn = u + 1 < n_streams ? bn / kWindow : -2;
    if (tp != t) {
        rec[t].u0p1 = u + 1;
        rec[t].b0 = b;
        rec[t].c0 = c;
    }
    if (tn != t) {
        rec[t].u1 = u + 1;
        rec[t].b1 = bn;
        rec[t].c1 = cn;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// stream feed, fused
// ---------------------------------------------------------------------------------------------------------------
struct StreamTileParams {
    DevTags tg;
    const uint8_t* bytes;
    const int64_t* chunk_off;
    int64_t n_chunks;
    const int64_t* stream_off;
    int64_t n_streams;
    const StreamState* st_in;
    StreamState* st_out;
    uint32_t flags;
    uint8_t* out;
    int64_t* out_chunk_off;
    unsigned long long* tile_state;  // [max_tiles], zeroed
    const TileRec* rec;              // [max_tiles], zeroed then filled by the pre-pass
    unsigned int* ticket;            // zeroed
    int64_t* unit_totals;            // [n_streams] scratch (general path)
    int64_t* unit_bases;             // [n_streams] scratch (general path)
    int64_t max_tiles;               // tiles the scratch has room for (>= total / kWindow + 1)
};

// General path inside a tile: warp-per-unit walkers (qb_walk.cuh), reading a unit from shared memory when it was staged
// and from global memory otherwise; per-unit results go through global scratch because the unit count is unbounded.
__device__ __noinline__ void stream_tile_general(const StreamTileParams& P, TileShared& S, int64_t tile, int64_t g0,
                                                 int staged, int64_t total) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const DevTags& tg = P.tg;
    const int64_t u0 = S.u0, u1 = S.u1;
    const uint8_t* sbuf = S.data - g0;  // sbuf + absolute offset addresses the staged copy
    for (int64_t s = u0 + wid; s < u1; s += kTileWarps) {
        const int64_t cf = P.stream_off[s], m = P.stream_off[s + 1] - cf;
        const int64_t ue = P.chunk_off[cf + m];
        const bool in_smem = ue - g0 <= staged;
        StreamWalker<false> w(tg, in_smem ? sbuf : P.bytes, in_smem ? g0 + staged : total, P.chunk_off + cf, m, lane, P.flags);
        StreamResult r = w.run(P.st_in ? P.st_in + s : nullptr, nullptr);
        if (lane == 0) {
            P.unit_totals[s] = r.total_out;
            atomicAdd(&S.agg, (unsigned long long)r.total_out);
        }
        __syncwarp();
    }
    __syncthreads();
    if (wid == 0) {
        int64_t base = lookback_exclusive(P.tile_state, tile, (int64_t)S.agg, lane);
        if (lane == 0) S.tile_base = base;
    }
    if (wid == 1 % kTileWarps) {  // exclusive offsets of the tile's units
        int64_t run = 0;
        for (int64_t b = u0; b < u1; b += 32) {
            const int64_t s = b + lane;
            const int64_t v = s < u1 ? P.unit_totals[s] : 0;
            int64_t incl = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int64_t t = __shfl_up_sync(kFull, incl, d);
                if (lane >= d) incl += t;
            }
            if (s < u1) P.unit_bases[s] = run + incl - v;
            run += __shfl_sync(kFull, incl, 31);
        }
    }
    __syncthreads();
    const int64_t tbase = S.tile_base;
    if (tile == 0 && tid == 0) P.out_chunk_off[0] = 0;
    for (int64_t s = u0 + wid; s < u1; s += kTileWarps) {
        const int64_t cf = P.stream_off[s], m = P.stream_off[s + 1] - cf;
        const int64_t ue = P.chunk_off[cf + m];
        const bool in_smem = ue - g0 <= staged;
        StreamWalker<true> w(tg, in_smem ? sbuf : P.bytes, in_smem ? g0 + staged : total, P.chunk_off + cf, m, lane, P.flags);
        w.out_base = tbase + P.unit_bases[s];
        w.dst = P.out + w.out_base;
        w.out_chunk_off = P.out_chunk_off;
        w.chunk0 = cf;
        StreamState si;
        if (P.st_in) si = P.st_in[s];
        w.run(P.st_in ? &si : nullptr, P.st_out ? P.st_out + s : nullptr);
        __syncwarp();
    }
}

__global__ void __launch_bounds__(kTileThreads, 4) stream_tile_kernel(const __grid_constant__ StreamTileParams P) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    TileShared& S = *reinterpret_cast<TileShared*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const DevTags& tg = P.tg;
    // the call's streams end at chunk stream_off[n_streams] (a first slice of a chained batch stops short of n_chunks)
    const int64_t total = P.chunk_off[P.stream_off[P.n_streams]];
    const int64_t n_tiles = total / kWindow + 1 <= P.max_tiles ? total / kWindow + 1 : P.max_tiles;
    uint32_t parity = 0;
    if (tid == 0) mbar_init(&S.mbar, 1);
    __syncthreads();

    // thread 0 keeps one ticket ahead: `cur` is this iteration's tile, `nxt` the next one, whose record is prefetched
    int64_t cur = 0, nxt = 0;
    bool have_rec = false;
    if (tid == 0) cur = (int64_t)atomicAdd(P.ticket, 1u);
#define QB_PREFETCH_REC()                                                     \
    do {                                                                      \
        have_rec = nxt < n_tiles;                                             \
        if (have_rec) {                                                       \
            const char* src_ = reinterpret_cast<const char*>(P.rec + nxt);    \
            char* dst_ = reinterpret_cast<char*>(&S.next_rec);                \
            cp_async16(dst_, src_);                                           \
            cp_async16(dst_ + 16, src_ + 16);                                 \
            cp_async16(dst_ + 32, src_ + 32);                                 \
        }                                                                     \
        cur = nxt;                                                            \
    } while (0)

    for (;;) {
        // ---- take a tile ----
        __syncthreads();
        if (tid == 0) {
            const int64_t tile = cur;
            S.tile = (int32_t)(tile < n_tiles ? tile : n_tiles);
            if (tile < n_tiles) {
                TileRec r;
                if (have_rec) {
                    cp_async_wait_all();
                    r = S.next_rec;
                } else {
                    r = P.rec[tile];
                }
                const int64_t nun = r.u0p1 ? r.u1 - (r.u0p1 - 1) : 0;
                const int64_t g0 = r.b0 & ~(int64_t)31;
                const bool fits = (r.b1 - g0) <= kStageCap;
                S.g0 = g0;
                S.u0 = r.u0p1 - 1;
                S.u1 = r.u1;
                S.c0 = r.c0;
                S.nunits = (int32_t)(nun > 0x7fffffff ? 0x7fffffff : nun);
                S.nchunks = (int32_t)((r.c1 - r.c0) > 0x7fffffff ? 0x7fffffff : (r.c1 - r.c0));
                int64_t stg = fits ? (r.b1 - g0) : kStageCap;
                if (g0 + stg > total) stg = total - g0;
                if (nun == 0) stg = 0;
                S.staged = (int32_t)stg;
                S.nsteps = (int32_t)((stg + 1023) >> 10);
                S.fast = (fits && nun <= kMaxUnits && nun > 0 && (r.c1 - r.c0) <= kChunkCap) ? 1 : 0;
                S.nseg = 0;
                S.overflow = 0;
                S.agg = 0;
                // stage [g0, g0 + staged): TMA bulk copy of the 16-byte multiple, tail bytes by hand below
                const uint32_t bulk = (uint32_t)(stg & ~(int64_t)15);
                if (bulk > 0) {
                    fence_proxy_async();
                    mbar_expect_tx(&S.mbar, bulk);
                    bulk_g2s(S.data, P.bytes + g0, bulk, &S.mbar);
                }
                nxt = (int64_t)atomicAdd(P.ticket, 1u);  // consumed later, by QB_PREFETCH_REC
            }
        }
        __syncthreads();
        const int64_t tile = S.tile;
        if (tile >= n_tiles) break;
        const int nunits = S.nunits;
        if (nunits == 0) {
            if (tid == 0) {
                if (tile == 0) P.out_chunk_off[0] = 0;
                QB_PREFETCH_REC();
            }
            if (wid == 0) lookback_exclusive(P.tile_state, tile, 0, lane);
            continue;
        }
        const int64_t g0 = S.g0;
        const int staged = S.staged;
        const bool fast = S.fast != 0;
        // per-unit geometry + the tile's feed boundaries while the bulk copy is in flight
        if (fast) {
            for (int u = tid; u <= nunits; u += kTileThreads) {
                const int64_t s = S.u0 + u;
                const int64_t cs = P.stream_off[s < P.n_streams ? s : P.n_streams];
                const int64_t b = P.chunk_off[cs];
                if (u < nunits) {
                    S.u_beg[u] = (int32_t)(b - g0);
                    S.u_rlast[u] = 0;
                    S.u_total[u] = 0;
                }
                if (u > 0) {
                    S.u_end[u - 1] = (int32_t)(b - g0);
                    S.u_cend[u - 1] = (int32_t)(cs - S.c0);
                }
            }
            for (int i = tid; i < S.nchunks; i += kTileThreads) S.bnd[i] = (int32_t)(P.chunk_off[S.c0 + 1 + i] - g0);
            for (int i = tid; i < kStageWords; i += kTileThreads) S.keptbm[i] = 0;
        }
        {   // tail bytes the bulk copy could not take (size not a multiple of 16), then wait for the copy
            const int bulk = staged & ~15;
            if (tid < staged - bulk) S.data[bulk + tid] = P.bytes[g0 + bulk + tid];
            if (bulk > 0) {
                while (!mbar_try_wait(&S.mbar, parity)) {}
                parity ^= 1;
            }
            if (tid == 0) QB_PREFETCH_REC();  // the next ticket has long arrived: fetch that tile's record in the background
        }
        __syncthreads();

        if (fast) {
            const int nsteps = S.nsteps;
            const int lo_pos = S.u_beg[0], hi_pos = S.u_end[nunits - 1];
            auto unit_end_of = [&](int x) -> int {  // end of the unit containing tile-relative position x
                int lo = 0, hi = nunits - 1;
                while (lo < hi) {
                    int mid = (lo + hi) >> 1;
                    if (S.u_end[mid] > x) hi = mid;
                    else lo = mid + 1;
                }
                return S.u_end[lo];
            };
            // ---- A: '<' bitmap + token records, one 1 KiB step per warp at a time ----
            for (int st = wid; st < nsteps; st += kTileWarps) {
                const int wi = st * 32 + lane;
                const int wb = wi * 32;
                uint32_t lt = 0;
                if (wb < staged) {
                    const uint4* p = reinterpret_cast<const uint4*>(S.data + wb);
                    uint4 a = p[0], b = p[1];
                    lt = lt_mask16(a) | (lt_mask16(b) << 16);
                    if (wb < lo_pos || wb + 32 > hi_pos) lt &= mask_ge(wb, lo_pos) & ~mask_ge(wb, hi_pos);
                }
                S.ltbm[wi] = lt;
                Tok r0{0, 0}, r1{0, 0};
                int cnt = 0;
                bool many = false;
                while (lt) {
                    const int b = __ffs(lt) - 1;
                    lt &= lt - 1;
                    const int x = wb + b;
                    int is_close, cls, exrel;
                    const int len = fast_match_tag(tg, S.data, x, nunits == 1 ? hi_pos : unit_end_of(x), is_close, cls, exrel);
                    if (len) {
                        Tok t;
                        t.x = (uint32_t)x;
                        t.y = (uint32_t)len | (is_close << 8) | (cls << 9) | ((uint32_t)exrel << 16);
                        if (cnt == 0) r0 = t;
                        else if (cnt == 1) r1 = t;
                        else many = true;
                        cnt++;
                    }
                }
                int incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    int t = __shfl_up_sync(kFull, incl, d);
                    if (lane >= d) incl += t;
                }
                const int tot = __shfl_sync(kFull, incl, 31);
                const bool bad = __any_sync(kFull, many) || tot > kSlots;
                if (bad) {
                    if (lane == 0) S.overflow = 1;
                } else {
                    const int slot = st * kSlots + incl - cnt;
                    if (cnt > 0) S.tok[slot] = r0;
                    if (cnt > 1) S.tok[slot + 1] = r1;
                }
                if (lane == 0) S.step_tok[st] = bad ? 0 : tot;
            }
            __syncthreads();
        }
        const bool fastok = fast && !S.overflow;

        if (fastok) {
            // ---- B: one warp per unit walks its tokens (all lanes in lock step, state uniform) ----
            for (int u = wid; u < nunits; u += kTileWarps) {
                const int ub = S.u_beg[u], ue = S.u_end[u];
                const int n = ue - ub;
                const int64_t s = S.u0 + u;
                const int cprev = u > 0 ? S.u_cend[u - 1] : 0;
                const int64_t m = S.u_cend[u] - cprev;
                const int64_t* co = P.chunk_off + (S.c0 + cprev);
                const int64_t uabs = g0 + ub;
                int32_t depth = 0;
                int h = 0;
                const uint8_t* heldp = S.data;
                if (P.st_in) {
                    depth = P.st_in[s].depth;
                    h = P.st_in[s].held_len;
                    heldp = P.st_in[s].held;
                }
                const HeldAcc A{S.data + ub, heldp, h};
                int last_tok_end = -1, seg_start = 0, held_kept = 0, held_unc = 0, cum = 0;
                auto first_boundary_after = [&](int64_t pos) -> int64_t {
                    int64_t lo = 1, hi = m + 1;
                    while (lo < hi) {
                        int64_t mid = (lo + hi) >> 1;
                        if (co[mid] - uabs > pos) hi = mid;
                        else lo = mid + 1;
                    }
                    return lo;
                };
                auto recognised = [&](bool is_close, int64_t x, int64_t tend, int64_t ex) -> bool {
                    if (is_close) return depth > 0;
                    if (depth > 0) return true;
                    if (tg.hold_enabled && ex < 0) return true;
                    const int64_t lo = tg.hold_enabled ? ex : x;
                    int64_t k = first_boundary_after(lo);
                    return !(k <= m && co[k] - uabs < tend);
                };
                auto emit = [&](int sa, int sb, int fin) {  // kept run [sa, sb) in unit coordinates
                    if (sb <= sa) return;
                    const int a = ub + sa;
                    const int len = sb - sa;
                    const int npieces = (len + kSegPiece - 1) / kSegPiece;
                    int slot = 0;
                    if (lane == 0) slot = atomicAdd(&S.nseg, npieces);
                    slot = __shfl_sync(kFull, slot, 0);
                    if (slot + npieces > kSegCap) {
                        S.overflow = 1;
                    } else {
                        for (int pi = lane; pi < npieces; pi += 32) {
                            Seg sg;
                            sg.src = (uint32_t)(a + pi * kSegPiece);
                            sg.len = (uint32_t)((pi == npieces - 1) ? len - pi * kSegPiece : kSegPiece);
                            sg.dst = (uint32_t)(cum + pi * kSegPiece);
                            sg.unit = (uint16_t)u;
                            sg.final_ = (uint16_t)fin;
                            S.seg[slot + pi] = sg;
                        }
                    }
                    cum += len;
                };
                if (m > 0) {
                    if (h > 0) {  // the carried partial tag
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
                            last_tok_end = seg_start = (int)tend;
                        } else {
                            held_unc = 1;
                            if (depth == 0) held_kept = h;
                        }
                    }
                    cum = held_kept;
                    if (n > 0) {
                        const int st0 = ub >> 10, st1 = (ue - 1) >> 10;
                        // steps of this unit that recorded tokens (lane j looks at steps st0+j and st0+32+j)
                        const int ca = st0 + lane <= st1 ? S.step_tok[st0 + lane] : 0;
                        const int cb = st0 + 32 + lane <= st1 ? S.step_tok[st0 + 32 + lane] : 0;
                        unsigned long long live = (unsigned long long)__ballot_sync(kFull, ca > 0) |
                                                  ((unsigned long long)__ballot_sync(kFull, cb > 0) << 32);
                        while (live) {
                            const int j = __ffsll((long long)live) - 1;
                            live &= live - 1;
                            const int st = st0 + j;
                            const int cnt = S.step_tok[st];
                            for (int k = 0; k < cnt; k++) {
                                const Tok tk = S.tok[st * kSlots + k];
                                if ((int)tk.x < ub || (int)tk.x >= ue) continue;
                                const int x = (int)tk.x - ub;
                                const int tend = x + (int)(tk.y & 0xff);
                                const bool is_close = (tk.y >> 8) & 1;
                                const int exr = (tk.y >> 16) & 0xff;
                                const int64_t ex = exr == 0xff ? -1 : x + exr;
                                if (recognised(is_close, x, tend, ex)) {
                                    if (depth == 0) {
                                        emit(seg_start, x, 0);
                                        depth = 1;
                                    } else {
                                        depth += is_close ? -1 : 1;
                                        if (depth == 0) seg_start = tend;
                                    }
                                    last_tok_end = tend;
                                }
                            }
                        }
                    }
                    if (depth == 0) emit(seg_start, n, 1);
                }
                if (lane == 0) {
                    S.u_depth[u] = depth;
                    S.u_last_tok_end[u] = last_tok_end;
                    S.u_held_kept[u] = (uint8_t)held_kept;
                    S.u_held_unc[u] = (uint8_t)held_unc;
                }
            }
            __syncthreads();
        }
        const bool fastok2 = fastok && !S.overflow;

        if (fastok2) {
            const int nsteps = S.nsteps;
            // ---- C: paint the kept bitmap from the segment pieces, kept bytes before every word (step_cum + wpre),
            //         then one thread per feed boundary ----
            {
                const int nseg0 = S.nseg;
                for (int i = wid; i < nseg0; i += kTileWarps) {
                    const int a = (int)S.seg[i].src, b = a + (int)S.seg[i].len;
                    const int w0 = a >> 5, w1 = (b - 1) >> 5;
                    for (int w = w0 + lane; w <= w1; w += 32) {
                        const uint32_t msk = mask_ge(w * 32, a) & ~mask_ge(w * 32, b);
                        if (msk == 0xffffffffu) S.keptbm[w] = msk;
                        else atomicOr(&S.keptbm[w], msk);
                    }
                }
            }
            __syncthreads();
            for (int st = wid; st < nsteps; st += kTileWarps) {
                const int wi = st * 32 + lane;
                const int pc = __popc(S.keptbm[wi]);
                int incl = pc;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    int t = __shfl_up_sync(kFull, incl, d);
                    if (lane >= d) incl += t;
                }
                S.wpre[wi] = (uint16_t)(incl - pc);
                if (lane == 31) S.step_cum[st] = incl;
            }
            __syncthreads();
            if (wid == 0) {
                const int tot = warp_scan64(S.step_cum, nsteps, lane);
                if (lane == 0) S.step_cum[nsteps] = tot;
            }
            __syncthreads();
            auto G = [&](int x) -> int {  // kept bytes in tile positions [0, x)
                const int w = x >> 5, b = x & 31;
                if ((w >> 5) >= nsteps) return S.step_cum[nsteps];
                return S.step_cum[w >> 5] + S.wpre[w] + (b ? __popc(S.keptbm[w] & ((1u << b) - 1u)) : 0);
            };
            const int nb = S.nchunks;
            for (int i = tid; i < nb; i += kTileThreads) {
                const int cpos = S.bnd[i];
                int lo = 0, hi = nunits - 1;  // unit of boundary i: first unit with u_cend > i
                while (lo < hi) {
                    int mid = (lo + hi) >> 1;
                    if (S.u_cend[mid] > i) hi = mid;
                    else lo = mid + 1;
                }
                const int u = lo;
                const int ub = S.u_beg[u];
                const int c = cpos - ub;  // unit coordinates
                const int hk = S.u_held_kept[u];
                int Pv = hk + G(cpos) - G(ub), R = c;
                bool found_lt = false;
                if (c > 0) {
                    const int w = (cpos - 1) >> 5;
                    const int nbits = cpos - w * 32;  // 1..32
                    const uint32_t low = nbits >= 32 ? 0xffffffffu : ((1u << nbits) - 1u);
                    const uint32_t ltw = S.ltbm[w] & low & mask_ge(w * 32, ub);
                    int p = 0;
                    bool p_kept = false;
                    if (ltw) {
                        const int b = 31 - __clz(ltw);
                        p = w * 32 + b;
                        p_kept = (S.keptbm[w] >> b) & 1u;
                        found_lt = true;
                    } else if (w > 0) {
                        const uint32_t ltp = S.ltbm[w - 1] & mask_ge((w - 1) * 32, ub);
                        if (ltp) {
                            const int b = 31 - __clz(ltp);
                            p = (w - 1) * 32 + b;
                            p_kept = (S.keptbm[w - 1] >> b) & 1u;
                            found_lt = true;
                        }
                    }
                    if (found_lt && p_kept && tg.hold_enabled && cpos - p <= kHeldMax &&
                        is_tag_prefix(tg, PtrAcc{S.data}, p, cpos, true, false)) {
                        R = p - ub;
                        Pv -= (cpos - p);
                    }
                }
                if (!found_lt && hk) {
                    const int64_t s = S.u0 + u;
                    const int h = P.st_in[s].held_len;
                    if (tg.hold_enabled && c + h <= kHeldMax &&
                        is_tag_prefix(tg, HeldAcc{S.data + ub, P.st_in[s].held, h}, -h, c, true, false)) {
                        R = -h;
                        Pv = 0;
                    }
                }
                S.bnd[i] = (u << 24) | Pv;
                if (i == S.u_cend[u] - 1) {  // the unit's last feed boundary
                    S.u_rlast[u] = R;
                    S.u_total[u] = Pv;
                }
            }
            __syncthreads();
            // ---- D: warp 0: unit output offsets inside the tile, tile aggregate, look-back;
            //         meanwhile the other warps work out the carried state of every unit ----
            if (wid == 0) {
                if (lane < nunits) S.u_out_rel[lane] = S.u_total[lane];
                if (lane + 32 < nunits) S.u_out_rel[lane + 32] = S.u_total[lane + 32];
                __syncwarp();
                const int tot = warp_scan64(S.u_out_rel, nunits, lane);
                const int64_t base = lookback_exclusive(P.tile_state, tile, tot, lane);
                if (lane == 0) S.tile_base = base;
            }
            static_assert(kMaxUnits <= kTileThreads - 32, "one unit per thread of warps 1..");
            StreamState so;
            const int my_u = tid - 32;  // warps 1.. : one unit per thread
            if (my_u >= 0 && my_u < nunits) {
                const int u = my_u;
                const int64_t s = S.u0 + u;
                const int ub = S.u_beg[u], n = S.u_end[u] - ub;
                const int m = S.u_cend[u] - (u > 0 ? S.u_cend[u - 1] : 0);
                StreamState si;
                if (P.st_in) {
                    si = P.st_in[s];
                } else {
                    si.depth = 0;
                    si.tail_len = -1;
                    si.held_len = 0;
                    for (int k = 0; k < kHeldMax; k++) si.held[k] = 0;
                }
                const int h = si.held_len;
                if (m == 0) {
                    so = si;
                    if (P.flags & 4u) so.held_len = 0;
                    for (int k = so.held_len; k < kHeldMax; k++) so.held[k] = 0;
                    so.tail_len = -1;
                } else {
                    const HeldAcc A{S.data + ub, si.held, h};
                    const int depth = S.u_depth[u], lte = S.u_last_tok_end[u], rl = S.u_rlast[u];
                    int hp = INT32_MIN;
                    if (!(P.flags & 4u)) {
                        if (depth == 0) {
                            if (rl < n) hp = rl;
                        } else {
                            int lo = n - kHeldMax;
                            if (lo < 0) lo = 0;
                            if (lte > lo) lo = lte;
                            int p = INT32_MIN;
                            for (int x = n - 1; x >= lo; x--)
                                if (S.data[ub + x] == '<') { p = x; break; }
                            if (p == INT32_MIN && S.u_held_unc[u] && lte < 0 && n + h <= kHeldMax) p = -h;
                            if (p != INT32_MIN && is_tag_prefix(tg, A, p, n, false, true)) hp = p;
                        }
                    }
                    so.depth = depth;
                    so.tail_len = lte >= 0 ? (n - lte) : -1;
                    so.held_len = 0;
                    for (int k = 0; k < kHeldMax; k++) so.held[k] = 0;
                    if (hp != INT32_MIN) {
                        so.held_len = (uint8_t)(n - hp);
                        for (int k = 0; k < so.held_len; k++) so.held[k] = A(hp + k);
                    }
                }
            }
            __syncthreads();
            // ---- E: per-chunk offsets, carried literal bytes, kept segments ----
            const int64_t tbase = S.tile_base;
            if (tile == 0 && tid == 0) P.out_chunk_off[0] = 0;
            for (int i = tid; i < nb; i += kTileThreads) {
                const int v = S.bnd[i];
                P.out_chunk_off[S.c0 + 1 + i] = tbase + S.u_out_rel[v >> 24] + (v & 0xffffff);
            }
            const int nseg = S.nseg;
            for (int i = wid; i < nseg; i += kTileWarps) {
                const Seg sg = S.seg[i];
                const int u = sg.unit;
                int len = (int)sg.len;
                if (sg.final_) {  // the unit's last kept run is clipped at the last release boundary
                    const int lim = S.u_beg[u] + S.u_rlast[u] - (int)sg.src;
                    if (lim < len) len = lim;
                }
                if (len > 0)
                    warp_copy(P.out + tbase + S.u_out_rel[u] + sg.dst, S.data + sg.src, len, S.data + kStageCap + 64, lane);
            }
            if (my_u >= 0 && my_u < nunits) {
                // carried bytes that turned out to be literal text come first in the unit's output; read them before the
                // state is overwritten (state_out may alias state_in)
                const int h = S.u_held_kept[my_u];
                if (h && S.u_rlast[my_u] > -h) {
                    const uint8_t* hb = P.st_in[S.u0 + my_u].held;
                    uint8_t* d = P.out + tbase + S.u_out_rel[my_u];
                    for (int k = 0; k < h; k++) d[k] = hb[k];
                }
                if (P.st_out) P.st_out[S.u0 + my_u] = so;
            }
        } else {
            stream_tile_general(P, S, tile, g0, staged, total);
        }
    }
}

}  // namespace qb
