// One stable 8-bit radix pass executed by a persistent cooperative grid, for `segs` independent
// segments of the same length.  Shared by the presort (presort.cuh) and the fused selection tail
// (tail.cuh): both are launch-latency problems at these sizes, so every pass of every sort runs inside
// one launch, separated by grid-wide syncs instead of kernel boundaries.
//
// A pass = tile histograms -> grid sync -> offsets by summation over the tile histograms and a stable
// scatter -> grid sync.  Keys are not stored: `key(seg, row)` recomputes the 32-bit key of the row an
// index refers to (the rows are small and L2-resident), so only indices move.
// Index buffers are laid out [seg][cap]; the source of the first pass may be one list shared by all
// segments (`src_shared`) or absent (identity: position p of segment s stands for index s * cap + p).
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"
#include "sort.cuh"

namespace evoxb {

constexpr int kCoopItems = 8;
constexpr int kCoopTile = 256 * kCoopItems;  // 2048 positions per tile

struct CoopSortSmem {
    uint32_t hist[kRadix];
    uint32_t base[kRadix];
    uint32_t wc[8][kRadix];
    uint32_t warp[8];
};

template <class KeyFn>
__device__ void coop_radix_pass(cooperative_groups::grid_group &grid, CoopSortSmem &sm, KeyFn key, int segs, int n,
                                int64_t cap, const uint32_t *__restrict__ src, bool src_shared,
                                uint32_t *__restrict__ dst, uint32_t *__restrict__ hist_g, int shift) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tiles = (n + kCoopTile - 1) / kCoopTile;  // per segment
    const int items = segs * tiles;
    const uint32_t lt_mask = (1u << lane) - 1u;

    // ---- phase 1: tile histograms ----
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
        const int seg = w / tiles, tile = w - seg * tiles;
        const uint32_t *s = src ? src + (src_shared ? 0 : (int64_t)seg * cap) : nullptr;
        sm.hist[threadIdx.x] = 0u;
        __syncthreads();
#pragma unroll
        for (int it = 0; it < kCoopItems; ++it) {
            const int p = tile * kCoopTile + it * 256 + threadIdx.x;
            if (p < n) {
                const uint32_t row = s ? __ldcg(s + p) : (uint32_t)(seg * cap + p);
                atomicAdd(&sm.hist[(key(seg, row) >> shift) & 0xffu], 1u);
            }
        }
        __syncthreads();
        hist_g[(int64_t)w * kRadix + threadIdx.x] = sm.hist[threadIdx.x];
        __syncthreads();
    }
    grid.sync();
    // ---- phase 2: offsets by summation over the segment's tile histograms, then stable scatter ----
    for (int w = blockIdx.x; w < items; w += gridDim.x) {
        const int seg = w / tiles, tile = w - seg * tiles;
        const uint32_t *s = src ? src + (src_shared ? 0 : (int64_t)seg * cap) : nullptr;
        uint32_t *d = dst + (int64_t)seg * cap;
        const uint32_t *hseg = hist_g + (int64_t)seg * tiles * kRadix;
        uint32_t before = 0u, total = 0u;
#pragma unroll 8
        for (int t = 0; t < tiles; ++t) {
            const uint32_t v = __ldcg(hseg + (int64_t)t * kRadix + threadIdx.x);
            total += v;
            before += (t < tile) ? v : 0u;
        }
        uint32_t incl = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) sm.warp[warp] = incl;
        for (int q = threadIdx.x; q < 8 * kRadix; q += 256) (&sm.wc[0][0])[q] = 0u;
        __syncthreads();
        uint32_t wbefore = 0u;
        for (int q = 0; q < warp; ++q) wbefore += sm.warp[q];
        sm.base[threadIdx.x] = wbefore + incl - total + before;  // first output slot of (digit, this tile)
        __syncthreads();

        // warp `warp` owns 256 consecutive positions of the tile and walks them in 8 rounds of 32, so
        // "earlier in the input" == (earlier warp, earlier round, lower lane)
        uint32_t rows[kCoopItems], off[kCoopItems], dig[kCoopItems];
        const int wbase = tile * kCoopTile + warp * (32 * kCoopItems);
#pragma unroll
        for (int it = 0; it < kCoopItems; ++it) {
            const int p = wbase + it * 32 + lane;
            const bool valid = p < n;
            rows[it] = valid ? (s ? __ldcg(s + p) : (uint32_t)(seg * cap + p)) : 0u;
            dig[it] = valid ? ((key(seg, rows[it]) >> shift) & 0xffu) : 0x100u;
            const uint32_t peers = __match_any_sync(0xffffffffu, dig[it]);
            const uint32_t dd = dig[it] & 0xffu;
            const uint32_t prev = sm.wc[warp][dd];
            __syncwarp();
            if (valid && (peers & lt_mask) == 0u) sm.wc[warp][dd] = prev + __popc(peers);
            __syncwarp();
            off[it] = prev + __popc(peers & lt_mask);
        }
        __syncthreads();
        {
            uint32_t run = sm.base[threadIdx.x];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const uint32_t c = sm.wc[q][threadIdx.x];
                sm.wc[q][threadIdx.x] = run;
                run += c;
            }
        }
        __syncthreads();
#pragma unroll
        for (int it = 0; it < kCoopItems; ++it) {
            const int p = wbase + it * 32 + lane;
            if (p < n) d[sm.wc[warp][dig[it]] + off[it]] = rows[it];
        }
        __syncthreads();
    }
    grid.sync();
}

}  // namespace evoxb
