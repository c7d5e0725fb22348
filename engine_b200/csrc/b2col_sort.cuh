// b2col_sort.cuh — hand-written LSD radix sort of (64-bit Morton key, 32-bit primitive index) pairs
// for sm_100a: 8-bit digits, one kernel per digit ("onesweep": every tile ranks its keys locally
// and obtains the number of equal-digit keys in all earlier tiles by decoupled look-back over a
// per-tile status word, so each pass reads and writes every pair exactly once).  The per-pass digit
// histograms are produced upstream by the Morton kernel (b2col.cu), which already touches every key.
//
// HBM traffic per pass: 12 B read + 12 B written per pair.  Stable, deterministic.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2 {

constexpr int RS_BITS = 8;
constexpr int RS_RADIX = 1 << RS_BITS;
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
#ifndef B2_RS_ITEMS
#define B2_RS_ITEMS 16
#endif
constexpr int RS_ITEMS = B2_RS_ITEMS;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 pairs per tile
constexpr int RS_MAX_PASSES = 8;

enum : uint32_t {
    LB_FLAG_AGG = 1u << 30,     // own digit count is posted
    LB_FLAG_PREFIX = 2u << 30,  // inclusive count over everything up to and including this one is posted
    LB_VALUE_MASK = (1u << 30) - 1u
};
#ifndef B2_LB_BATCH
#define B2_LB_BATCH 8
#endif
constexpr int LB_BATCH = B2_LB_BATCH;
// Tiles per look-back group, a template parameter of the pass: 8 from 128 tiles up (-2.8 us per update at 1M, -25 us at 10M),
// 16 below (200k colliders: 70.8 vs 73.6 us with 8).  The status array is sized for the smaller group.
#ifndef B2_LB_GROUP_SMALL
#define B2_LB_GROUP_SMALL 8
#endif
constexpr uint32_t LB_GROUP_SMALL = B2_LB_GROUP_SMALL, LB_GROUP_LARGE = 16;
static_assert(RS_THREADS == RS_RADIX, "one thread per digit in the scan phases");

#ifdef B2_SORT_TIMING  // dev builds only: per-tile phase timestamps (ns, %globaltimer) of the LAST pass launched
__device__ unsigned long long g_sort_timing[8192][8];
#define RS_STAMP(k)                                                                              \
    do {                                                                                         \
        if (threadIdx.x == 0 && blockIdx.x < 8192) {                                             \
            unsigned long long t_;                                                               \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                               \
            g_sort_timing[blockIdx.x][k] = t_;                                                   \
        }                                                                                        \
    } while (0)
#else
#define RS_STAMP(k) ((void)0)
#endif

// Exclusive scan of one value per thread across the 256-thread block. `warp_tot` is 8 words of smem.
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t* warp_tot)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t base = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w)
        if (w < warp) base += warp_tot[w];
    __syncthreads();
    return base + inc - v;
}

// One digit pass. Grid = number of tiles; tiles are claimed through `ticket` so that a tile can only
// ever wait on tiles that already started (forward progress of the look-back).
//   ghist    [256]  total count of each digit value in this pass (over all n keys)
//   lookback [tiles + ceil(tiles / LB_GROUP)][256] zero-initialised status words for this pass
//
// MODE 0: 64-bit keys and 32-bit values in separate arrays (keys wider than 32 bits).
// MODE 1, 2: the key is at most 32 bits wide and travels as ONE 64-bit word (key << 32 | value) — 8 instead of 12 bytes per
//   pair and pass, one staging round and one scatter instead of two; `shift` then addresses the upper half (32 + 8 * pass).
//   MODE 1 writes packed words again (vin / vout unused); MODE 2, the last pass, writes kout = key, vout = value, so that
//   everything downstream reads the same arrays whatever the key width.
#ifndef B2_RS_MIN_BLOCKS
#define B2_RS_MIN_BLOCKS 2  // 128 registers: with 1 the compiler takes ~170 and a 10M-pair pass slows from 111 to 162 us
#endif
template <int MODE, uint32_t LB_GROUP>
__global__ void __launch_bounds__(RS_THREADS, B2_RS_MIN_BLOCKS)
k_onesweep(const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin, uint64_t* __restrict__ kout, uint32_t* __restrict__ vout,
           uint32_t n, uint32_t n_tiles, int shift, const uint32_t* __restrict__ ghist, uint32_t* lookback, uint32_t* ticket)
{
    __shared__ uint64_t skeys[RS_TILE];  // reused for the values after the keys are written out
    __shared__ uint32_t whist[RS_WARPS][RS_RADIX];
    __shared__ uint32_t sbase[RS_RADIX];  // first tile-local sorted position of each digit
    __shared__ uint32_t gbase[RS_RADIX];  // global position of that first element
    __shared__ uint32_t warp_tot[RS_WARPS];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();  // frame path launches with programmatic stream serialization (b2col.cu, pdl_wait)
#endif
    RS_STAMP(0);
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) whist[w][tid] = 0u;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t tile_base = tile * (uint32_t)RS_TILE;
    const uint32_t valid = min((uint32_t)RS_TILE, n - tile_base);

    // ---- load, warp-striped: item k of lane l sits at warp*512 + k*32 + l (coalesced, order-preserving)
    uint64_t key[RS_ITEMS];
    uint32_t val[RS_ITEMS];
    uint32_t rank[RS_ITEMS];
    const uint32_t wbase = (uint32_t)warp * (RS_ITEMS * 32u) + (uint32_t)lane;
#ifdef B2_RS_BULK_LOAD
    // dev variant: the tile's keys are ONE contiguous range of up to 32 KB — moved into shared memory by the
    // bulk-copy engine (cp.async.bulk + mbarrier), then read from there in warp-striped order (tiles start 32 KB apart: 16-byte aligned;
    // the size is rounded up to 16 bytes: at most one key beyond the array's last, inside the buffer's allocation slack)
    {
        __shared__ alignas(8) uint64_t s_bar;
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            const uint32_t bytes = (valid * 8u + 15u) & ~15u;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"((uint32_t)__cvta_generic_to_shared(skeys)),
                         "l"(kin + tile_base), "r"(bytes), "r"(bar)
                         : "memory");
        }
        __syncthreads();  // the barrier is initialised
        asm volatile("{\n\t.reg .pred P1;\n\tRS_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0;\n\t@P1 bra RS_DONE;\n\tbra RS_WAIT;\n\tRS_DONE:\n\t}" ::"r"(bar) : "memory");
#pragma unroll
        for (int k = 0; k < RS_ITEMS; ++k) {
            const uint32_t idx = wbase + (uint32_t)k * 32u;
            const bool ok = idx < valid;
            key[k] = ok ? skeys[idx] : ~0ull;
            if (MODE == 0) val[k] = ok ? vin[tile_base + idx] : 0u;
        }
        __syncthreads();  // skeys is reused by the scatter
    }
#else
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const uint32_t idx = wbase + (uint32_t)k * 32u;
        const bool ok = idx < valid;
        key[k] = ok ? kin[tile_base + idx] : ~0ull;
        if (MODE == 0) val[k] = ok ? vin[tile_base + idx] : 0u;
    }
#endif

#ifdef B2_SORT_TIMING
    if (key[0] == 0x123456789abcdefull && key[RS_ITEMS - 1] == 77ull) s_tile = 0;  // wait for the loads
#endif
    RS_STAMP(1);
    // ---- rank within the warp: lanes with equal digits find each other with one ballot per digit bit
    // (measured on B200: MATCH.ANY runs at roughly one per 55 cycles per SM and made this phase 8 us of a
    // 21 us pass; eight VOTE + LOP3 pairs are several times cheaper).
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const uint32_t idx = wbase + (uint32_t)k * 32u;
        const bool ok = idx < valid;
        const uint32_t d = (uint32_t)(key[k] >> shift) & (RS_RADIX - 1);
        uint32_t peers = __ballot_sync(0xffffffffu, ok);
#pragma unroll
        for (int b = 0; b < RS_BITS; ++b) {
            const uint32_t m = __ballot_sync(0xffffffffu, (d >> b) & 1u);
            peers &= ((d >> b) & 1u) ? m : ~m;
        }
        const int leader = __ffs(peers) - 1;  // peers contains this lane whenever ok
        uint32_t old = 0u;
        if (ok && lane == leader) {
            old = whist[warp][d];
            whist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, ok ? leader : 0);
        rank[k] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    RS_STAMP(2);

    // ---- thread d: digit d. Exclusive offsets of each warp inside the tile's digit-d run.
    uint32_t tile_count = 0u;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
        const uint32_t c = whist[w][tid];
        whist[w][tid] = tile_count;
        tile_count += c;
    }
    // ---- publish the tile's count, then sum the counts of all earlier tiles.  Two levels, because at ~1M
    // pairs every tile of the pass is resident and starts at the same moment: a plain decoupled look-back
    // then degenerates into a serial walk over up to `tile` status words (dozens of dependent L2 round trips).
    //   level 1: tiles form groups of LB_GROUP; a tile sums the counts of the earlier tiles of its own group
    //            directly (<= LB_GROUP-1 independent loads, one round trip once they are posted);
    //   level 2: the last tile of a group posts the group's count, then — by decoupled look-back over GROUP
    //            status words, batched — the inclusive count over all groups up to its own; every tile
    //            obtains the count of the groups before its own the same way.
    // A tile only ever waits on tiles with a smaller ticket, which have started: forward progress holds for any
    // grid size.
    uint32_t excl = 0u;
    {
        volatile uint32_t* ta = lookback;                                   // [tiles][256] tile counts
        volatile uint32_t* gs = lookback + (size_t)n_tiles * RS_RADIX;      // [groups][256] group status
        const uint32_t grp = tile / LB_GROUP, first = grp * LB_GROUP;
        ta[(size_t)tile * RS_RADIX + tid] = LB_FLAG_AGG | tile_count;
        uint32_t in_group = 0u;
        {
            uint32_t v[LB_GROUP - 1];
#pragma unroll
            for (int j = 0; j < (int)LB_GROUP - 1; ++j) {
                const uint32_t t = first + (uint32_t)j;
                v[j] = t < tile ? ta[(size_t)t * RS_RADIX + tid] : LB_FLAG_AGG;
            }
#pragma unroll
            for (int j = 0; j < (int)LB_GROUP - 1; ++j) {
                const uint32_t t = first + (uint32_t)j;
                uint32_t w = v[j];
                while ((w >> 30) == 0u) w = ta[(size_t)t * RS_RADIX + tid];  // not posted yet: that tile is running (ticket order)
                in_group += w & LB_VALUE_MASK;
            }
        }
        const bool closes_group = (tile % LB_GROUP) == LB_GROUP - 1u;
        if (closes_group) gs[(size_t)grp * RS_RADIX + tid] = (grp == 0u ? LB_FLAG_PREFIX : LB_FLAG_AGG) | (in_group + tile_count);
        uint32_t before = 0u;  // count over all groups before mine
        if (grp > 0u) {
            int g = (int)grp - 1;
            bool done = false;
            while (!done) {
                uint32_t v[LB_BATCH];
#pragma unroll
                for (int j = 0; j < LB_BATCH; ++j) v[j] = (g - j >= 0) ? gs[(size_t)(g - j) * RS_RADIX + tid] : LB_FLAG_PREFIX;  // PREFIX with value 0 ends the walk
#pragma unroll
                for (int j = 0; j < LB_BATCH; ++j) {
                    if (done) break;
                    uint32_t w = v[j];
                    while ((w >> 30) == 0u) w = gs[(size_t)(g - j) * RS_RADIX + tid];
                    before += w & LB_VALUE_MASK;
                    done = (w >> 30) == 2u;
                }
                g -= LB_BATCH;
            }
            if (closes_group) gs[(size_t)grp * RS_RADIX + tid] = LB_FLAG_PREFIX | (before + in_group + tile_count);
        }
        excl = before + in_group;
    }
    const uint32_t g_excl = block_excl_scan_256(ghist[tid], warp_tot);    // keys with a smaller digit, globally
    const uint32_t t_excl = block_excl_scan_256(tile_count, warp_tot);    // ... within this tile
    sbase[tid] = t_excl;
    gbase[tid] = g_excl + excl;
    __syncthreads();
    RS_STAMP(3);

    // ---- scatter keys to their tile-local sorted position
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const uint32_t idx = wbase + (uint32_t)k * 32u;
        if (idx < valid) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & (RS_RADIX - 1);
            rank[k] += sbase[d] + whist[warp][d];
            skeys[rank[k]] = key[k];
        }
    }
    __syncthreads();
    // ---- write keys: consecutive threads write consecutive addresses within each digit run
    uint32_t dst[RS_ITEMS];
#pragma unroll
    for (int m = 0; m < RS_ITEMS; ++m) {
        const uint32_t j = (uint32_t)tid + (uint32_t)m * RS_THREADS;
        if (j < valid) {
            const uint64_t kk = skeys[j];
            const uint32_t d = (uint32_t)(kk >> shift) & (RS_RADIX - 1);
            dst[m] = gbase[d] + (j - sbase[d]);
            if (MODE == 2) {
                kout[dst[m]] = kk >> 32;
                vout[dst[m]] = (uint32_t)kk;
            }
            else {
                kout[dst[m]] = kk;
            }
        }
    }
    RS_STAMP(4);
    if (MODE != 0) {
        RS_STAMP(5);
        return;
    }
    __syncthreads();
    uint32_t* svals = reinterpret_cast<uint32_t*>(skeys);
#pragma unroll
    for (int k = 0; k < RS_ITEMS; ++k) {
        const uint32_t idx = wbase + (uint32_t)k * 32u;
        if (idx < valid) svals[rank[k]] = val[k];
    }
    __syncthreads();
#pragma unroll
    for (int m = 0; m < RS_ITEMS; ++m) {
        const uint32_t j = (uint32_t)tid + (uint32_t)m * RS_THREADS;
        if (j < valid) vout[dst[m]] = svals[j];
    }
    RS_STAMP(5);
}


// =================================================================================================
// MSD + block-local sort (keys of at most 32 bits, packed as key << 32 | index; up to ~1.2M pairs)
// =================================================================================================
// Four LSD passes at ~1M pairs are four dependent, latency-bound launches of ~15 us (look-back across tiles that all
// start at the same moment: profiles/r1_e).  Instead: ONE onesweep pass on the TOP digit (k_onesweep with shift 56)
// partitions the array into 256 buckets — each a contiguous range of the final order — and k_local_sort finishes every
// bucket inside one block: the bucket's pairs live in registers / shared memory through three LSD passes on the
// remaining 24 bits, with no communication between blocks at all.  Two launches instead of four, the second one free
// of any look-back.
//
// A bucket larger than LS_TILE pairs (a scene whose colliders crowd into one 256th of the Morton range) is still
// sorted correctly, by the same block, with three LSD passes through GLOBAL memory over its own segment (tile after
// tile with running digit offsets — slow: one block, but exact); the largest bucket of the frame is reported to the
// host, which then chooses the plain LSD pipeline for the following frames (b2col_update).
constexpr int LS_THREADS = 512;
constexpr int LS_ITEMS = 16;
constexpr int LS_WARPS = LS_THREADS / 32;
constexpr int LS_TILE = LS_THREADS * LS_ITEMS;  // 8192 pairs: twice the mean bucket at 1M

struct LocalSortSmem {
    uint64_t keys[LS_TILE];
    uint32_t whist[LS_WARPS][RS_RADIX];  // per-warp digit counts, then each warp's offset inside the tile's digit run
    uint32_t base[RS_RADIX];             // first tile-local position of each digit
    uint32_t cnt[RS_RADIX];              // tile count of each digit
    uint32_t run[RS_RADIX];              // oversized buckets: where the next pair of each digit goes in the segment
    uint32_t warp_tot[LS_WARPS];
};

// exclusive scan of one value per thread over the first 256 threads of the block (threads >= 256 pass 0 and get garbage)
__device__ __forceinline__ uint32_t ls_scan_256(uint32_t v, uint32_t* warp_tot)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t before = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w)
        if (w < warp) before += warp_tot[w];
    __syncthreads();
    return before + inc - v;
}

// Stable tile-local position of every pair held in registers for the digit at `shift`; leaves sm.base / sm.cnt for the
// tile.  Layout: warp-striped over `items` rounds — item k < items of lane l is element warp * items * 32 + k * 32 + l —
// where items = ceil(valid / LS_THREADS) is block-uniform: a bucket of 4096 pairs keeps all 16 warps busy for 8 rounds
// instead of half of the warps for 16.
__device__ __forceinline__ void ls_rank_tile(LocalSortSmem& sm, const uint64_t* key, uint32_t valid, int shift, uint32_t* pos, int items)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t wbase = (uint32_t)warp * ((uint32_t)items * 32u) + (uint32_t)lane;
    for (int i = tid; i < LS_WARPS * RS_RADIX; i += LS_THREADS) (&sm.whist[0][0])[i] = 0u;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < LS_ITEMS; ++k) {
        if (k >= items) break;  // block-uniform
        const bool ok = wbase + (uint32_t)k * 32u < valid;
        const uint32_t d = (uint32_t)(key[k] >> shift) & (RS_RADIX - 1);
        uint32_t peers = __ballot_sync(0xffffffffu, ok);
#pragma unroll
        for (int b = 0; b < RS_BITS; ++b) {
            const uint32_t m = __ballot_sync(0xffffffffu, (d >> b) & 1u);
            peers &= ((d >> b) & 1u) ? m : ~m;
        }
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0u;
        if (ok && lane == leader) {
            old = sm.whist[warp][d];
            sm.whist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, ok ? leader : 0);
        pos[k] = old + __popc(peers & lt_mask);
        __syncwarp();
    }
    __syncthreads();
    uint32_t tile_count = 0u;
    if (tid < RS_RADIX) {
#pragma unroll
        for (int w = 0; w < LS_WARPS; ++w) {
            const uint32_t c = sm.whist[w][tid];
            sm.whist[w][tid] = tile_count;
            tile_count += c;
        }
    }
    const uint32_t excl = ls_scan_256(tid < RS_RADIX ? tile_count : 0u, sm.warp_tot);
    if (tid < RS_RADIX) {
        sm.base[tid] = excl;
        sm.cnt[tid] = tile_count;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < LS_ITEMS; ++k) {
        if (k < items && wbase + (uint32_t)k * 32u < valid) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & (RS_RADIX - 1);
            pos[k] += sm.base[d] + sm.whist[warp][d];
        }
    }
}

// One block per bucket of the MSD pass.  in: packed pairs partitioned by the top digit (counts in top_hist);
// scratch: the other packed buffer (same size; only oversized buckets use their segment of it);
// out: kout[i] = key, vout[i] = index, sorted.  stat[0] = largest bucket of the frame (atomicMax; host-visible).
__global__ void __launch_bounds__(LS_THREADS, 2)
k_local_sort(uint64_t* in, uint64_t* scratch, uint64_t* kout, uint32_t* vout, const uint32_t* __restrict__ top_hist, uint32_t* stat)  // kout may alias scratch
{
    extern __shared__ __align__(16) unsigned char ls_raw[];
    LocalSortSmem& sm = *reinterpret_cast<LocalSortSmem*>(ls_raw);
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 900
    cudaGridDependencySynchronize();
#endif
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // this bucket's segment: exclusive scan of the top-digit histogram
    const uint32_t mine = tid < RS_RADIX ? top_hist[tid] : 0u;
    const uint32_t excl = ls_scan_256(mine, sm.warp_tot);
    if (tid == (int)blockIdx.x) {
        sm.run[0] = excl;
        sm.run[1] = mine;
    }
    __syncthreads();
    const uint32_t start = sm.run[0], count = sm.run[1];
    __syncthreads();
    if (count == 0u) return;
    if (tid == 0) atomicMax(stat, count);
    uint64_t key[LS_ITEMS];
    uint32_t pos[LS_ITEMS];
    if (count <= (uint32_t)LS_TILE) {
        // ---- the bucket fits the block: three LSD passes through registers and shared memory
        const int items = (int)((count + LS_THREADS - 1u) / LS_THREADS);  // rounds this bucket needs (block-uniform)
        const uint32_t wbase = (uint32_t)warp * ((uint32_t)items * 32u) + (uint32_t)lane;
#pragma unroll
        for (int k = 0; k < LS_ITEMS; ++k) {
            const uint32_t idx = wbase + (uint32_t)k * 32u;
            key[k] = (k < items && idx < count) ? in[start + idx] : ~0ull;
        }
#pragma unroll 1
        for (int pass = 0; pass < 3; ++pass) {
            ls_rank_tile(sm, key, count, 32 + RS_BITS * pass, pos, items);
#pragma unroll
            for (int k = 0; k < LS_ITEMS; ++k)
                if (k < items && wbase + (uint32_t)k * 32u < count) sm.keys[pos[k]] = key[k];
            __syncthreads();
#pragma unroll
            for (int k = 0; k < LS_ITEMS; ++k) {
                const uint32_t idx = wbase + (uint32_t)k * 32u;
                if (k < items && idx < count) key[k] = sm.keys[idx];
            }
            __syncthreads();
        }
#pragma unroll
        for (int k = 0; k < LS_ITEMS; ++k) {
            const uint32_t idx = wbase + (uint32_t)k * 32u;
            if (k < items && idx < count) {
                kout[start + idx] = key[k] >> 32;
                vout[start + idx] = (uint32_t)key[k];
            }
        }
        return;
    }
    const uint32_t wbase = (uint32_t)warp * (LS_ITEMS * 32u) + (uint32_t)lane;
    // ---- oversized bucket: the same three passes over the segment in global memory, tile after tile (exact, slow)
#pragma unroll 1
    for (int pass = 0; pass < 3; ++pass) {
        const int shift = 32 + RS_BITS * pass;
        const uint64_t* src = (pass == 1) ? scratch : in;  // in -> scratch -> in -> (kout, vout)
        uint64_t* dst = (pass == 0) ? scratch : in;
        if (tid < RS_RADIX) sm.run[tid] = 0u;
        __syncthreads();
        for (uint32_t i = tid; i < count; i += LS_THREADS) atomicAdd(&sm.run[(uint32_t)(__ldcg(src + start + i) >> shift) & (RS_RADIX - 1)], 1u);
        __syncthreads();
        const uint32_t c = tid < RS_RADIX ? sm.run[tid] : 0u;
        const uint32_t e = ls_scan_256(c, sm.warp_tot);
        if (tid < RS_RADIX) sm.run[tid] = e;
        __syncthreads();
        for (uint32_t tile = 0; tile < count; tile += LS_TILE) {
            const uint32_t valid = min((uint32_t)LS_TILE, count - tile);
#pragma unroll
            for (int k = 0; k < LS_ITEMS; ++k) {
                const uint32_t idx = wbase + (uint32_t)k * 32u;
                key[k] = idx < valid ? __ldcg(src + start + tile + idx) : ~0ull;  // written by this block in the pass before: read at L2
            }
            ls_rank_tile(sm, key, valid, shift, pos, LS_ITEMS);
#pragma unroll
            for (int k = 0; k < LS_ITEMS; ++k) {
                if (wbase + (uint32_t)k * 32u < valid) {
                    const uint32_t d = (uint32_t)(key[k] >> shift) & (RS_RADIX - 1);
                    const uint32_t at = start + sm.run[d] + (pos[k] - sm.base[d]);
                    if (pass == 2) {
                        kout[at] = key[k] >> 32;
                        vout[at] = (uint32_t)key[k];
                    }
                    else {
                        dst[at] = key[k];
                    }
                }
            }
            __syncthreads();
            if (tid < RS_RADIX) sm.run[tid] += sm.cnt[tid];
            __syncthreads();
        }
        __threadfence_block();
        __syncthreads();
    }
}

}  // namespace b2
