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
constexpr int kSlab = 1024;   // columns per build tile
constexpr int kChunk = 4096;  // positions per chunk (= csrc/peel.cuh kChunkCols): reordering stays inside a chunk

struct PresortPasses {
    int count;
    int kind[kMaxPasses];   // 0: low 32 bits of S', 1: high bits of S', 2: key of objective 0
    int shift[kMaxPasses];
    int fast_first;         // index of the first key_0 pass: passes [fast_first, count) alone sort by key_0
};

constexpr int kPresortRunMax = 32;  // longest run of equal key_0 the fast path orders in place

__global__ void __launch_bounds__(256)
presort_coop_kernel(const float *__restrict__ x, int n, int m, int npad, PresortPasses passes,
                    uint32_t *__restrict__ perm_a, uint32_t *__restrict__ perm_b, uint32_t *__restrict__ hist_g,
                    int *__restrict__ tile_max, int *__restrict__ hp, float *__restrict__ xt, int *__restrict__ perm_out,
                    int *__restrict__ uid_row, int *__restrict__ uid_col, int *__restrict__ uid_row_q,
                    int *__restrict__ uid_col_q, float *__restrict__ tile_min1, float *__restrict__ tile_max1,
                    int *__restrict__ slow_flag) {
    __shared__ CoopSortSmem csm;
    __shared__ uint32_t sh_warp[8];
    __shared__ int sh_carry;
    __shared__ uint32_t slab_key[kChunk];
    __shared__ uint16_t slab_idx[kChunk];
    __shared__ float band_lo[8], band_hi[8];
    namespace cgx = cooperative_groups;
    cgx::grid_group grid = cgx::this_grid();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // Fast path: rows are first sorted by key_0 alone (4 of the 9 passes).  The order by (key_0, S') differs
    // from that only inside runs of equal key_0, which are short for real-valued objectives (a few thousand
    // pairs among 262144 uniform floats): every run of at most kPresortRunMax rows is put in S' order in place
    // by the thread that finds its head (stable, so the result is the permutation the full sort would give).
    // A longer run (tie-heavy data: integer grids, duplicated rows) raises *slow_flag (zeroed by the host)
    // and the whole grid falls back to all passes.
    const uint32_t *src = nullptr;  // nullptr = identity permutation
    uint32_t *dst = perm_a;
    auto run_passes = [&](int first, int last) {
        for (int ps = first; ps < last; ++ps) {
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
    };
    // Which path: the fast one needs every value of key_0 to occur at most kPresortRunMax times.  Real populations
    // break that (objectives clipped at a bound: hundreds of rows with f_0 == 0 in a converged NSGA-II run), and
    // finding it out after four wasted passes costs more than asking first: a hashed multiplicity count of key_0
    // (buckets in perm_b, which is free until the first pass) raises *slow_flag when any bucket exceeds the limit
    // -- a bucket holds every copy of a value, so no flag means no long run (collisions only err to the slow side).
    bool fast = passes.fast_first > 0;
    if (fast) {
        int nb = 4096;
        while (nb * 2 <= npad && nb < 262144) nb *= 2;
        for (int q = blockIdx.x * 256 + threadIdx.x; q < nb; q += gridDim.x * 256) perm_b[q] = 0u;
        grid.sync();
        for (int p = blockIdx.x * 256 + threadIdx.x; p < n; p += gridDim.x * 256) {
            uint32_t k0 = float_to_key(x[(int64_t)p * m]);
            k0 = (k0 ^ (k0 >> 15)) * 0x9E3779B1u;
            if (atomicAdd(perm_b + (k0 >> 8) % (uint32_t)nb, 1u) >= (uint32_t)kPresortRunMax) atomicOr(slow_flag, 1);
        }
        grid.sync();
        fast = *reinterpret_cast<volatile int *>(slow_flag) == 0;
    }
    run_passes(fast ? passes.fast_first : 0, passes.count);
    if (fast) {
        uint32_t *cur = const_cast<uint32_t *>(src);
        for (int p = blockIdx.x * 256 + threadIdx.x; p < n; p += gridDim.x * 256) {
            const uint32_t k0 = float_to_key(x[(int64_t)__ldcg(cur + p) * m]);
            if (p > 0 && float_to_key(x[(int64_t)__ldcg(cur + p - 1) * m]) == k0) continue;  // not a head
            int len = 1;
            while (p + len < n && len <= kPresortRunMax && float_to_key(x[(int64_t)__ldcg(cur + p + len) * m]) == k0) ++len;
            if (len == 1) continue;
            if (len > kPresortRunMax) {
                atomicOr(slow_flag, 1);
                continue;
            }
            // stable insertion sort of the run by S' (rows of equal (key_0, S') keep their input order)
            uint32_t rows[kPresortRunMax];
            unsigned long long sp[kPresortRunMax];
            for (int i = 0; i < len; ++i) {
                const uint32_t row = __ldcg(cur + p + i);
                const unsigned long long v = sum_tail_keys(x + (int64_t)row * m, m);
                int j = i;
                while (j > 0 && sp[j - 1] > v) {
                    sp[j] = sp[j - 1];
                    rows[j] = rows[j - 1];
                    --j;
                }
                sp[j] = v;
                rows[j] = row;
            }
            for (int i = 0; i < len; ++i) cur[p + i] = rows[i];
        }
        grid.sync();
        if (*reinterpret_cast<volatile int *>(slow_flag) != 0) {
            src = nullptr;
            dst = perm_a;
            run_passes(0, passes.count);
        }
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

    // ---- order inside every 4096-chunk: by objective 1, then inside every 1024-tile by objective 2 ----
    // The predicate (uid_i < uid_j) & all_{k>=1}(x_ik <= x_jk) does not depend on where rows sit, and a
    // row can only dominate rows of its own chunk or of later chunks (uid is monotone in the (key_0, S')
    // order; the peel only assumes "no bits left of the row's own chunk"), so the order INSIDE a chunk is
    // free.  Sorting it by objective 1 makes each of its four 1024-tiles a band of x_1; sorting each tile
    // by objective 2 makes every run of 32 / 128 positions in it a narrow band of x_2.  The build kernel
    // then (a) discards a whole (row tile, column tile) whose x_1 bands cannot match, (b) discards
    // (32 rows x 128 columns) blocks whose x_2 bands cannot match, (c) drops a compare that always holds.
    // One CTA per chunk, bitonic network in shared memory on (key, position) pairs -- positions make the
    // order total, so the result is deterministic.  Positions >= n carry the largest key and therefore
    // stay at the end of the last chunk.
    const float nanv = __int_as_float(0x7fc00000);
    const int nchunks = npad / kChunk;
    for (int chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
        const int base = chunk * kChunk;
        auto row_of = [&](int q) -> uint32_t { return src ? __ldcg(src + q) : (uint32_t)q; };
        for (int level = 1; level <= 2 && level < m; ++level) {
            // keys of this level, gathered through the order reached so far
            for (int t = threadIdx.x; t < kChunk; t += 256) {
                const int q = base + (level == 1 ? t : (int)slab_idx[t]);
                slab_key[t] = (q < n) ? float_to_key(x[(int64_t)row_of(q) * m + level]) : 0xffffffffu;
                if (level == 1) slab_idx[t] = (uint16_t)t;
            }
            __syncthreads();
            const int span = (level == 1) ? kChunk : kSlab;  // level 2 sorts every 1024-tile on its own
            for (int k = 2; k <= span; k <<= 1) {
                for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
                    for (int h = 0; h < kChunk / 512; ++h) {
                        const int u = threadIdx.x + h * 256;  // kChunk / 2 compare-exchanges per step
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
        // ---- final layout of the chunk: sorted objective-major copy, permutation, uids, x_1 band per tile ----
        for (int tile = 0; tile < kChunk / kSlab; ++tile) {
            float lo = __int_as_float(0x7fc00000), hi = lo;  // NaN: "no value yet" (fminf / fmaxf skip it)
            for (int t = tile * kSlab + threadIdx.x; t < (tile + 1) * kSlab; t += 256) {
                const int p = base + t;
                const int q = (m >= 2) ? base + (int)slab_idx[t] : p;
                uid_row[p] = __ldcg(uid_row_q + q);
                uid_col[p] = __ldcg(uid_col_q + q);
                if (q < n) {
                    const uint32_t row = row_of(q);
                    perm_out[p] = (int)row;
                    for (int k = 0; k < m; ++k) {
                        const float v = x[(int64_t)row * m + k];
                        xt[(int64_t)k * npad + p] = v;
                        if (k == 1) { lo = fminf(lo, v); hi = fmaxf(hi, v); }
                    }
                } else {
                    perm_out[p] = -1;
                    for (int k = 0; k < m; ++k) xt[(int64_t)k * npad + p] = nanv;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            }
            __syncthreads();
            if (lane == 0) { band_lo[warp] = lo; band_hi[warp] = hi; }
            __syncthreads();
            if (threadIdx.x == 0) {
                for (int w = 1; w < 8; ++w) { lo = fminf(lo, band_lo[w]); hi = fmaxf(hi, band_hi[w]); }
                tile_min1[chunk * (kChunk / kSlab) + tile] = lo;
                tile_max1[chunk * (kChunk / kSlab) + tile] = hi;
            }
        }
        __syncthreads();
    }
}

}  // namespace evoxb
