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

}  // namespace evoxb
