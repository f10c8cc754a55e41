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
#include "coop_sort.cuh"
#include "sort.cuh"

namespace evoxb {

__device__ __forceinline__ unsigned long long sum_tail_keys(const float *__restrict__ row, int m) {
    unsigned long long acc = 0ull;
    for (int k = 1; k < m; ++k) acc += float_to_key(row[k]);
    return acc;
}

// ---- the whole presort in ONE cooperative launch ------------------------------------------------------
// The 3-kernels-per-pass radix sort of sort.cuh is launch-bound at these sizes (9 passes = 27 launches
// plus key gathers).  Here a persistent grid runs every pass back to back: per pass a tile histogram,
// a grid-wide sync, then prefix-by-summation over the tile histograms and a stable scatter, then a
// second sync.  Keys are never stored: each pass recomputes the digit from the fitness row addressed
// through the current permutation (rows are a few bytes and L2-resident).  After the last pass the same
// launch marks the heads of equal-(key_0, S') runs, scans them into uids and writes the sorted
// objective-major copy, so the build kernel can start right behind it.
constexpr int kMaxPasses = 12;
constexpr int kSlab = 1024;  // positions per slab = columns per build tile

struct PresortPasses {
    int count;
    int kind[kMaxPasses];   // 0: low 32 bits of S', 1: high bits of S', 2: key of objective 0
    int shift[kMaxPasses];
};

__global__ void __launch_bounds__(256)
presort_coop_kernel(const float *__restrict__ x, int n, int m, int npad, PresortPasses passes,
                    uint32_t *__restrict__ perm_a, uint32_t *__restrict__ perm_b, uint32_t *__restrict__ hist_g,
                    int *__restrict__ tile_max, int *__restrict__ hp, float *__restrict__ xt, int *__restrict__ perm_out,
                    int *__restrict__ uid_row, int *__restrict__ uid_col, int *__restrict__ uid_row_q,
                    int *__restrict__ uid_col_q) {
    __shared__ CoopSortSmem csm;
    __shared__ uint32_t sh_warp[8];
    __shared__ int sh_carry;
    __shared__ uint32_t slab_key[kSlab];
    __shared__ uint16_t slab_idx[kSlab];
    namespace cgx = cooperative_groups;
    cgx::grid_group grid = cgx::this_grid();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    const uint32_t *src = nullptr;  // nullptr = identity permutation
    uint32_t *dst = perm_a;
    for (int ps = 0; ps < passes.count; ++ps) {
        const int kind = passes.kind[ps];
        auto key = [=](int, uint32_t row) -> uint32_t {
            const float *r = x + (int64_t)row * m;
            if (kind == 2) return float_to_key(r[0]);
            const unsigned long long sp = sum_tail_keys(r, m);
            return kind == 0 ? (uint32_t)sp : (uint32_t)(sp >> 32);
        };
        coop_radix_pass(grid, csm, key, 1, n, npad, src, false, dst, hist_g, passes.shift[ps]);
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
    // ---- uid = 1 + latest head at or before p (still indexed by the (key_0, S') position) ----
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
            uid_row_q[p] = bad ? 0x7fffffff : u;
            uid_col_q[p] = bad ? 0 : u;
        }
        __syncthreads();
    }
    grid.sync();

    // ---- order inside every 1024-slab by objective 1 ----
    // The predicate (uid_i < uid_j) & all_{k>=1}(x_ik <= x_jk) does not depend on where rows sit, and a
    // row can only dominate rows of its own slab or of later slabs (uid is monotone in the (key_0, S')
    // order), so the order INSIDE a slab is free.  Sorting it by objective 1 gives every run of 32 / 128
    // consecutive positions a narrow band of x_1, which lets the build kernel skip (rows, columns)
    // blocks whose bands cannot satisfy x_i1 <= x_j1, and drop that compare where they always do.
    // One CTA per slab, bitonic network in shared memory on (key, position) pairs -- positions make the
    // order total, so the result is deterministic.  Positions >= n carry the largest key and therefore
    // stay at the end of the last slab.
    const float nanv = __int_as_float(0x7fc00000);
    const int nslabs = npad / kSlab;
    for (int slab = blockIdx.x; slab < nslabs; slab += gridDim.x) {
        const int base = slab * kSlab;
        auto row_of = [&](int q) -> uint32_t { return src ? __ldcg(src + q) : (uint32_t)q; };
        for (int level = 1; level <= 1 && level < m; ++level) {  // (a second level on objective 2 was measured: no gain)
            // keys of this level, gathered through the order reached so far
            for (int t = threadIdx.x; t < kSlab; t += 256) {
                const int q = base + (level == 1 ? t : (int)slab_idx[t]);
                slab_key[t] = (q < n) ? float_to_key(x[(int64_t)row_of(q) * m + level]) : 0xffffffffu;
                if (level == 1) slab_idx[t] = (uint16_t)t;
            }
            __syncthreads();
            const int span = (level == 1) ? kSlab : 128;  // level 2 sorts every 128-band on its own
            for (int k = 2; k <= span; k <<= 1) {
                for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int u = threadIdx.x + h * 256;  // 512 compare-exchanges per step
                        const int i = ((u & ~(j - 1)) << 1) | (u & (j - 1));
                        const int l = i | j;
                        const uint32_t ki = slab_key[i], kl = slab_key[l];
                        const uint16_t ii = slab_idx[i], il = slab_idx[l];
                        const bool up = (i & k) == 0 || k == span;  // the last merge leaves every block ascending
                        const bool gt = ki > kl || (ki == kl && ii > il);
                        if (gt == up) {
                            slab_key[i] = kl; slab_key[l] = ki;
                            slab_idx[i] = il; slab_idx[l] = ii;
                        }
                    }
                    __syncthreads();
                }
            }
        }
        // ---- final layout of the slab: sorted objective-major copy, permutation, uids ----
        for (int t = threadIdx.x; t < kSlab; t += 256) {
            const int p = base + t;
            const int q = (m >= 2) ? base + (int)slab_idx[t] : p;
            uid_row[p] = __ldcg(uid_row_q + q);
            uid_col[p] = __ldcg(uid_col_q + q);
            if (q < n) {
                const uint32_t row = row_of(q);
                perm_out[p] = (int)row;
                for (int k = 0; k < m; ++k) xt[(int64_t)k * npad + p] = x[(int64_t)row * m + k];
            } else {
                perm_out[p] = -1;
                for (int k = 0; k < m; ++k) xt[(int64_t)k * npad + p] = nanv;
            }
        }
        __syncthreads();
    }
}

}  // namespace evoxb
