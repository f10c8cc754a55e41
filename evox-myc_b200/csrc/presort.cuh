// Presort of the population, so that the dominance matrix becomes strictly upper triangular and one
// objective compare per pair disappears.
//
// No counterpart in the reference (its matrix at non_dominate.py:60 is built in input order).
// Rows are sorted by the pair (key_0, S') where key_k is the order-preserving u32 image of objective
// k (lax.sort order: -0 == +0) and S' = sum_{k>=1} key_k as an exact integer (< 2^36 for M <= 16).
// Why that is enough: if row a dominates row b then key_k(a) <= key_k(b) for every k with at least
// one strict, hence (key_0, S') of a is lexicographically smaller: a sits BEFORE b, in a different
// run of equal pairs.  Conversely, with uid = 1-based start position of a row's run of equal pairs,
//     D[i][j] = (uid_i < uid_j) & all_{k>=1}(x_ik <= x_jk)
// holds for ALL ordered pairs of NaN-free rows: uid_i < uid_j gives key_0(i) <= key_0(j), i.e.
// x_i0 <= x_j0, which completes all(<=); and rows with all(<=) that are not identical have a strictly
// smaller pair, while identical rows share a uid -- exactly all(<=) & any(<) of common.py:98.
// The sort costs 9 radix passes whatever M is (a full lexicographic sort would need 4 M).
// Rows holding a NaN never dominate and are never dominated (every compare with NaN is false):
// they get uid_row = INT_MAX, uid_col = 0, which makes the uid test false in both roles.
// Ranks are produced in sorted positions and written back through the permutation.
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"
#include "sort.cuh"

namespace evoxb {

__device__ __forceinline__ unsigned long long sum_tail_keys(const float *__restrict__ row, int m) {
    unsigned long long acc = 0ull;
    for (int k = 1; k < m; ++k) acc += float_to_key(row[k]);
    return acc;
}

// First sort key: low 32 bits of S'.  vals = identity.
__global__ void presort_keys_sumlo_kernel(const float *__restrict__ x, int64_t n, int m, uint32_t *__restrict__ keys,
                                          uint32_t *__restrict__ vals) {
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= n) return;
    keys[p] = (uint32_t)sum_tail_keys(x + p * m, m);
    vals[p] = (uint32_t)p;
}

// which = 0: high bits of S' (bits 32..39); which = 1: key of objective 0.  Keys follow the
// current permutation `perm`; vals = perm.
__global__ void presort_keys_gather_kernel(const float *__restrict__ x, int64_t n, int m, int which,
                                           const uint32_t *__restrict__ perm, uint32_t *__restrict__ keys,
                                           uint32_t *__restrict__ vals) {
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t row = perm[p];
    keys[p] = which == 0 ? (uint32_t)(sum_tail_keys(x + (int64_t)row * m, m) >> 32) : float_to_key(x[(int64_t)row * m]);
    vals[p] = row;
}

constexpr int kPrepTile = 4096;  // 256 threads x 16 positions

// Gathers the sorted population into objective-major xt[k][p] (NaN padding past n), writes the
// permutation as int, marks heads of runs of equal (key_0, S') pairs (hp[p] = p for a head, -1
// otherwise), flags NaN rows, and reduces the largest head position of every 4096-tile.
__global__ void __launch_bounds__(256)
sorted_gather_kernel(const float *__restrict__ x, int64_t n, int m, int64_t npad, const uint32_t *__restrict__ perm_u,
                     float *__restrict__ xt, int *__restrict__ perm, int *__restrict__ hp, uint8_t *__restrict__ nanrow,
                     int *__restrict__ tile_max) {
    __shared__ int wmax[8];
    const float nanv = __int_as_float(0x7fc00000);
    int best = -1;
#pragma unroll 1
    for (int it = 0; it < 16; ++it) {
        const int64_t p = (int64_t)blockIdx.x * kPrepTile + it * 256 + threadIdx.x;
        if (p >= npad) continue;
        if (p >= n) {
            for (int k = 0; k < m; ++k) xt[(int64_t)k * npad + p] = nanv;
            hp[p] = -1;
            nanrow[p] = 1;
            continue;
        }
        const uint32_t row = perm_u[p];
        bool has_nan = false;
        for (int k = 0; k < m; ++k) {
            const float v = x[(int64_t)row * m + k];
            xt[(int64_t)k * npad + p] = v;
            has_nan = has_nan || (v != v);
        }
        bool differs = (p == 0);
        if (p > 0) {
            const float *a = x + (int64_t)row * m, *b = x + (int64_t)perm_u[p - 1] * m;
            differs = float_to_key(a[0]) != float_to_key(b[0]) || sum_tail_keys(a, m) != sum_tail_keys(b, m);
        }
        perm[p] = (int)row;
        nanrow[p] = has_nan ? 1 : 0;
        hp[p] = differs ? (int)p : -1;
        if (differs) best = max(best, (int)p);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((threadIdx.x & 31) == 0) wmax[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        int b = -1;
        for (int w = 0; w < 8; ++w) b = max(b, wmax[w]);
        tile_max[blockIdx.x] = b;
    }
}

// uid = 1 + position of the latest head at or before p (inclusive max-scan of hp).
// Thread t owns the 16 consecutive positions [tile0 + 16 t, tile0 + 16 t + 16).
__global__ void __launch_bounds__(256)
uid_kernel(const int *__restrict__ hp, const uint8_t *__restrict__ nanrow, const int *__restrict__ tile_max, int64_t npad,
           int *__restrict__ uid_row, int *__restrict__ uid_col) {
    __shared__ int wmax[8];
    __shared__ int carry_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int c = -1;
    for (int t = threadIdx.x; t < (int)blockIdx.x; t += 256) c = max(c, tile_max[t]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c = max(c, __shfl_xor_sync(0xffffffffu, c, o));
    if (lane == 0) wmax[warp] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int b = -1;
        for (int w = 0; w < 8; ++w) b = max(b, wmax[w]);
        carry_s = b;
    }
    __syncthreads();
    const int64_t first = (int64_t)blockIdx.x * kPrepTile + threadIdx.x * 16;
    int loc[16];
    int run = -1;
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        const int64_t p = first + e;
        const int h = (p < npad) ? hp[p] : -1;
        run = max(run, h);
        loc[e] = run;
    }
    int incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl = max(incl, t);
    }
    int excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = -1;
    __syncthreads();
    if (lane == 31) wmax[warp] = incl;
    __syncthreads();
    int before = carry_s;
    for (int w = 0; w < warp; ++w) before = max(before, wmax[w]);
    before = max(before, excl);
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        const int64_t p = first + e;
        if (p < npad) {
            const int u = max(before, loc[e]) + 1;
            const bool bad = nanrow[p] != 0;
            uid_row[p] = bad ? 0x7fffffff : u;
            uid_col[p] = bad ? 0 : u;
        }
    }
}

// ---- the whole presort in ONE cooperative launch ------------------------------------------------------
// The 3-kernels-per-pass radix sort of sort.cuh is launch-bound at these sizes (9 passes = 27 launches
// plus key gathers).  Here a persistent grid runs every pass back to back: per pass a tile histogram,
// a grid-wide sync, then prefix-by-summation over the tile histograms and a stable scatter, then a
// second sync.  Keys are never stored: each pass recomputes the digit from the fitness row addressed
// through the current permutation (rows are a few bytes and L2-resident).  After the last pass the same
// launch marks the heads of equal-(key_0, S') runs, scans them into uids and writes the sorted
// objective-major copy, so the build kernel can start right behind it.
constexpr int kCoopItems = 8;
constexpr int kCoopTile = 256 * kCoopItems;  // 2048 positions per tile
constexpr int kMaxPasses = 12;

struct PresortPasses {
    int count;
    int kind[kMaxPasses];   // 0: low 32 bits of S', 1: high bits of S', 2: key of objective 0
    int shift[kMaxPasses];
};

__device__ __forceinline__ uint32_t presort_digit(const float *__restrict__ row, int m, int kind, int shift) {
    uint32_t key;
    if (kind == 2) key = float_to_key(row[0]);
    else {
        const unsigned long long sp = sum_tail_keys(row, m);
        key = kind == 0 ? (uint32_t)sp : (uint32_t)(sp >> 32);
    }
    return (key >> shift) & 0xffu;
}

__global__ void __launch_bounds__(256)
presort_coop_kernel(const float *__restrict__ x, int n, int m, int npad, PresortPasses passes,
                    uint32_t *__restrict__ perm_a, uint32_t *__restrict__ perm_b, uint32_t *__restrict__ hist_g,
                    int *__restrict__ tile_max, int *__restrict__ hp, float *__restrict__ xt, int *__restrict__ perm_out,
                    int *__restrict__ uid_row, int *__restrict__ uid_col) {
    __shared__ uint32_t sh_hist[kRadix];
    __shared__ uint32_t sh_base[kRadix];
    __shared__ uint32_t wc[8][kRadix];
    __shared__ uint32_t sh_warp[8];
    __shared__ int sh_carry;
    namespace cgx = cooperative_groups;
    cgx::grid_group grid = cgx::this_grid();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tiles = (n + kCoopTile - 1) / kCoopTile;
    const uint32_t lt_mask = (1u << lane) - 1u;

    const uint32_t *src = nullptr;  // nullptr = identity permutation
    uint32_t *dst = perm_a;
    for (int ps = 0; ps < passes.count; ++ps) {
        const int kind = passes.kind[ps], shift = passes.shift[ps];
        // ---- phase 1: tile histograms ----
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            sh_hist[threadIdx.x] = 0u;
            __syncthreads();
#pragma unroll
            for (int it = 0; it < kCoopItems; ++it) {
                const int p = tile * kCoopTile + it * 256 + threadIdx.x;
                if (p < n) {
                    const uint32_t row = src ? __ldcg(src + p) : (uint32_t)p;
                    atomicAdd(&sh_hist[presort_digit(x + (int64_t)row * m, m, kind, shift)], 1u);
                }
            }
            __syncthreads();
            hist_g[tile * kRadix + threadIdx.x] = sh_hist[threadIdx.x];
            __syncthreads();
        }
        grid.sync();
        // ---- phase 2: offsets by summation over the tile histograms, then stable scatter ----
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            uint32_t before = 0u, total = 0u;
#pragma unroll 8
            for (int t = 0; t < tiles; ++t) {
                const uint32_t v = __ldcg(hist_g + t * kRadix + threadIdx.x);
                total += v;
                before += (t < tile) ? v : 0u;
            }
            uint32_t incl = total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            if (lane == 31) sh_warp[warp] = incl;
            for (int q = threadIdx.x; q < 8 * kRadix; q += 256) (&wc[0][0])[q] = 0u;
            __syncthreads();
            uint32_t wbefore = 0u;
            for (int w = 0; w < warp; ++w) wbefore += sh_warp[w];
            sh_base[threadIdx.x] = wbefore + incl - total + before;  // first output slot of (digit, this tile)
            __syncthreads();

            uint32_t rows[kCoopItems], off[kCoopItems], dig[kCoopItems];
            const int wbase = tile * kCoopTile + warp * (32 * kCoopItems);
#pragma unroll
            for (int it = 0; it < kCoopItems; ++it) {
                const int p = wbase + it * 32 + lane;
                const bool valid = p < n;
                rows[it] = valid ? (src ? __ldcg(src + p) : (uint32_t)p) : 0u;
                dig[it] = valid ? presort_digit(x + (int64_t)rows[it] * m, m, kind, shift) : 0x100u;
                const uint32_t peers = __match_any_sync(0xffffffffu, dig[it]);
                const uint32_t d = dig[it] & 0xffu;
                const uint32_t prev = wc[warp][d];
                __syncwarp();
                if (valid && (peers & lt_mask) == 0u) wc[warp][d] = prev + __popc(peers);
                __syncwarp();
                off[it] = prev + __popc(peers & lt_mask);
            }
            __syncthreads();
            {
                uint32_t run = sh_base[threadIdx.x];
#pragma unroll
                for (int w = 0; w < 8; ++w) {
                    const uint32_t c = wc[w][threadIdx.x];
                    wc[w][threadIdx.x] = run;
                    run += c;
                }
            }
            __syncthreads();
#pragma unroll
            for (int it = 0; it < kCoopItems; ++it) {
                const int p = wbase + it * 32 + lane;
                if (p < n) dst[wc[warp][dig[it]] + off[it]] = rows[it];
            }
            __syncthreads();
        }
        grid.sync();
        src = dst;
        dst = (dst == perm_a) ? perm_b : perm_a;
    }

    // ---- heads of equal-(key_0, S') runs, NaN flags (kept in uid_col for now), tile maxima ----
    const int ptiles = (npad + kCoopTile - 1) / kCoopTile;
    for (int tile = blockIdx.x; tile < ptiles; tile += gridDim.x) {
        int best = -1;
#pragma unroll 1
        for (int it = 0; it < kCoopItems; ++it) {
            const int p = tile * kCoopTile + it * 256 + threadIdx.x;
            if (p >= npad) continue;
            if (p >= n) {
                hp[p] = -1;
                uid_col[p] = 1;  // "NaN row" flag until the uid phase
                continue;
            }
            const float *a = x + (int64_t)(src ? __ldcg(src + p) : (uint32_t)p) * m;
            bool has_nan = false;
            for (int k = 0; k < m; ++k) has_nan = has_nan || (a[k] != a[k]);
            bool differs = (p == 0);
            if (p > 0) {
                const float *b = x + (int64_t)(src ? __ldcg(src + p - 1) : (uint32_t)(p - 1)) * m;
                differs = float_to_key(a[0]) != float_to_key(b[0]) || sum_tail_keys(a, m) != sum_tail_keys(b, m);
            }
            uid_col[p] = has_nan ? 1 : 0;
            hp[p] = differs ? p : -1;
            if (differs) best = max(best, p);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
        __syncthreads();
        if (lane == 0) sh_warp[warp] = (uint32_t)best;
        __syncthreads();
        if (threadIdx.x == 0) {
            int b = -1;
            for (int w = 0; w < 8; ++w) b = max(b, (int)sh_warp[w]);
            tile_max[tile] = b;
        }
    }
    grid.sync();
    // ---- uid = 1 + latest head at or before p; sorted objective-major copy; permutation as int ----
    const float nanv = __int_as_float(0x7fc00000);
    for (int tile = blockIdx.x; tile < ptiles; tile += gridDim.x) {
        int c = -1;
        for (int t = threadIdx.x; t < tile; t += 256) c = max(c, __ldcg(tile_max + t));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c = max(c, __shfl_xor_sync(0xffffffffu, c, o));
        __syncthreads();
        if (lane == 0) sh_warp[warp] = (uint32_t)c;
        __syncthreads();
        if (threadIdx.x == 0) {
            int b = -1;
            for (int w = 0; w < 8; ++w) b = max(b, (int)sh_warp[w]);
            sh_carry = b;
        }
        __syncthreads();
        const int first = tile * kCoopTile + threadIdx.x * kCoopItems;  // thread owns kCoopItems consecutive positions
        int loc[kCoopItems];
        int run = -1;
#pragma unroll
        for (int e = 0; e < kCoopItems; ++e) {
            const int p = first + e;
            run = max(run, (p < npad) ? __ldcg(hp + p) : -1);
            loc[e] = run;
        }
        int incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl = max(incl, t);
        }
        int excl = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) excl = -1;
        __syncthreads();
        if (lane == 31) sh_warp[warp] = (uint32_t)incl;
        __syncthreads();
        int before = sh_carry;
        for (int w = 0; w < warp; ++w) before = max(before, (int)sh_warp[w]);
        before = max(before, excl);
#pragma unroll
        for (int e = 0; e < kCoopItems; ++e) {
            const int p = first + e;
            if (p >= npad) continue;
            const bool bad = __ldcg(uid_col + p) != 0;
            const int u = max(before, loc[e]) + 1;
            uid_row[p] = bad ? 0x7fffffff : u;
            uid_col[p] = bad ? 0 : u;
            if (p < n) {
                const uint32_t row = src ? __ldcg(src + p) : (uint32_t)p;
                perm_out[p] = (int)row;
                for (int k = 0; k < m; ++k) xt[(int64_t)k * npad + p] = x[(int64_t)row * m + k];
            } else {
                for (int k = 0; k < m; ++k) xt[(int64_t)k * npad + p] = nanv;
            }
        }
        __syncthreads();
    }
}

}  // namespace evoxb
