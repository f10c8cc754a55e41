// Levels for the level form of K1 (dominance.cuh: dominance_build_lvl_kernel) and the fix-up that makes it exact.
//
// Replaces nothing of the reference by itself: it is a faster evaluation of `dominate_relation`
// (src/evox/utils/common.py:95-104) for the presorted population.  Every objective k >= 1 is mapped to a 15-bit
// level by a monotone function of the float order (-0 == +0, NaN excluded):
//   order key (sort.cuh: float_to_key) -> first-level bin = its top 12 bits (sign, exponent, 3 mantissa bits);
//   a histogram over the bins gives every non-empty bin a share of the 32768 levels in proportion to its rows
//   (at least one); inside a bin the level grows linearly with the remaining 20 key bits.
// x_ik <= x_jk  =>  level_ik <= level_jk, so "all levels <=" is a superset of "all floats <="; the two differ only
// for pairs that share a level in some objective and have x_ik > x_jk there.  Those pairs are enumerated from a
// counting sort of the rows by (objective, level) and their bits cleared with atomicAnd (level_fixup_kernel).
// Levels whose rows all hold the same value ("pure", e.g. integer-valued or heavily tied objectives) need no
// fix-up and cost nothing.  If the mixed levels are too crowded for that to be cheap (values packed far more
// densely than 32768 levels can resolve), a flag is raised on the device and the float kernel runs instead.
#pragma once

#include "common.cuh"
#include "dominance.cuh"
#include "peel.cuh"
#include "sort.cuh"

namespace evoxb {

constexpr int kLvlBins = 4096;
constexpr int kLvlCount = 32768;
constexpr int kLvlSpare = kLvlCount - kLvlBins;  // levels handed out in proportion; every non-empty bin gets >= 1 on top
constexpr int kLvlMaxBucket = 512;               // a mixed level with more rows than this -> float kernel
constexpr int kLvlBudgetPerRow = 64;             // sum over mixed levels of rows^2 <= this * n, per objective

struct LevelBufs {
    int K, W;           // compared objectives (m - 1) and packed words per row
    uint32_t *q;        // [W][npad] level words, row form (guard bits clear); objective 2w+1 low half, 2w+2 high half
    uint32_t *zero0;    // start of the region cleared before every call: cost, list_count, hist1, cnt, kmax
    size_t zero_bytes;
    unsigned long long *cost;  // [K] sum over mixed levels of rows^2
    uint32_t *list_count;      // entries of list
    unsigned long long *list;  // [list_cap] bits to clear after the build: (word index << 5) | bit
    int *list_col;             // [list_cap] column position of the entry (its dominator count drops with the bit)
    uint32_t list_cap;
    uint32_t *hist1;    // [K][kLvlBins]
    uint32_t *cnt;      // [K][kLvlCount] rows per level
    uint32_t *kmax;     // [K][kLvlCount] largest key of the level
    uint32_t *kmin;     // [K][kLvlCount] smallest key (set to all ones before every call)
    uint2 *bins;        // [K][kLvlBins] {first level, levels} of a bin
    uint32_t *start;    // [K][kLvlCount + 1] first slot of a level in items
    uint32_t *cursor;   // [K][kLvlCount]
    int2 *items;        // [K][npad] (position, float bits) grouped by level (mixed levels only); W > 2
    int4 *items4;       // the same buffer for W <= 2: (position, float bits, level word 0, level word 1), unused slots -1
};

__device__ __forceinline__ uint32_t level_of(uint32_t key, const uint2 *__restrict__ bins_k) {
    const uint2 b = __ldg(bins_k + (key >> 20));
    return b.x + __umulhi((key & 0xfffffu) << 12, b.y);
}

// Rows per first-level bin.  grid = (CTAs, K)
__global__ void __launch_bounds__(256)
level_hist_kernel(const float *__restrict__ xt, int npad, int n, uint32_t *__restrict__ hist1) {
    __shared__ uint32_t h[kLvlBins];
    const int k = blockIdx.y;
    for (int i = threadIdx.x; i < kLvlBins; i += 256) h[i] = 0u;
    __syncthreads();
    const float *x = xt + (int64_t)(k + 1) * npad;
    for (int p = blockIdx.x * 256 + threadIdx.x; p < n; p += gridDim.x * 256) {
        const uint32_t key = float_to_key(x[p]);
        if (key != 0xffffffffu) atomicAdd(&h[key >> 20], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kLvlBins; i += 256)
        if (h[i]) atomicAdd(&hist1[k * kLvlBins + i], h[i]);
}

// Block-wide exclusive scan of one value per thread (1024 threads); returns the exclusive prefix, *total = sum.
__device__ __forceinline__ uint32_t block_scan_1024(uint32_t v, uint32_t *sh_warp, uint32_t *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    __syncthreads();
    if (lane == 31) sh_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = sh_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        sh_warp[lane] = wi - w;
        if (lane == 31) sh_warp[32] = wi;
    }
    __syncthreads();
    *total = sh_warp[32];
    return sh_warp[warp] + incl - v;
}

// Levels per bin and their first level.  One CTA of 1024 threads per objective, 4 bins per thread.
__global__ void __launch_bounds__(1024) level_plan_kernel(const uint32_t *__restrict__ hist1, uint2 *__restrict__ bins) {
    __shared__ uint32_t sh_warp[33];
    const int k = blockIdx.x;
    uint32_t h[4], rows = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        h[e] = hist1[k * kLvlBins + threadIdx.x * 4 + e];
        rows += h[e];
    }
    uint32_t total;
    block_scan_1024(rows, sh_warp, &total);
    uint32_t ns[4], mine = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        ns[e] = h[e] ? 1u + (uint32_t)(((unsigned long long)h[e] * kLvlSpare) / total) : 0u;
        mine += ns[e];
    }
    uint32_t levels;
    uint32_t base = block_scan_1024(mine, sh_warp, &levels);
    // sum <= kLvlSpare + non-empty bins <= kLvlCount
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        bins[k * kLvlBins + threadIdx.x * 4 + e] = make_uint2(base, ns[e]);
        base += ns[e];
    }
}

// Level words of every position; rows, smallest and largest key per level.  One thread per position.
template <int W>
__global__ void __launch_bounds__(256)
level_assign_kernel(const float *__restrict__ xt, int npad, int n, int K, const uint2 *__restrict__ bins,
                    uint32_t *__restrict__ q, uint32_t *__restrict__ cnt, uint32_t *__restrict__ kmin,
                    uint32_t *__restrict__ kmax) {
    const int p = blockIdx.x * 256 + threadIdx.x;
    if (p >= npad) return;
#pragma unroll
    for (int w = 0; w < W; ++w) {
        uint32_t word = 0u;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int k = 2 * w + hh;  // objective k + 1
            if (k < K && p < n) {
                const uint32_t key = float_to_key(xt[(int64_t)(k + 1) * npad + p]);
                if (key != 0xffffffffu) {
                    const uint32_t lv = level_of(key, bins + k * kLvlBins);
                    word |= lv << (16 * hh);
                    atomicAdd(&cnt[k * kLvlCount + lv], 1u);
                    atomicMin(&kmin[k * kLvlCount + lv], key);
                    atomicMax(&kmax[k * kLvlCount + lv], key);
                }
            }
        }
        q[(int64_t)w * npad + p] = word;
    }
}

// First slot of every level (exclusive scan of the rows per level), and the decision whether the fix-up is
// affordable.  grid = (kLvlScanParts, K), 256 threads: a CTA scans one quarter of an objective's levels; instead of
// waiting for the quarters in front of it, it sums them itself (at most 96 KB of coalesced reads).  A warp owns
// 1024 consecutive levels, lane l the levels l, l + 32, ... of them, so every access is coalesced and the scan is
// 32 warp scans with a running carry.
constexpr int kLvlScanParts = 4;
__global__ void __launch_bounds__(256)
level_scan_kernel(const uint32_t *__restrict__ cnt, const uint32_t *__restrict__ kmin, const uint32_t *__restrict__ kmax,
                  uint32_t *__restrict__ start, uint32_t *__restrict__ cursor, int n, unsigned long long *__restrict__ cost_acc,
                  int *__restrict__ fallback) {
    __shared__ uint32_t sh_warp[8];
    __shared__ uint32_t sh_front[8];
    const int k = blockIdx.y, part = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kPart = kLvlCount / kLvlScanParts;  // 8192 levels = 8 warps x 1024
    const uint32_t *ck = cnt + (size_t)k * kLvlCount;
    // rows in the quarters in front of this one
    uint32_t front = 0;
    {
        const uint4 *c4 = reinterpret_cast<const uint4 *>(ck);  // 16-byte aligned: kLvlCount * 4 bytes per objective
#pragma unroll 8
        for (int i = threadIdx.x; i < part * kPart / 4; i += 256) {
            const uint4 v = c4[i];
            front += v.x + v.y + v.z + v.w;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) front += __shfl_xor_sync(0xffffffffu, front, o);
    if (lane == 0) sh_front[warp] = front;

    const int l0 = part * kPart + warp * 1024;
    uint32_t c[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) c[j] = ck[l0 + j * 32 + lane];
    uint32_t big = 0;
    unsigned long long cost = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int i = k * kLvlCount + l0 + j * 32 + lane;
        if (c[j] > 1 && kmin[i] != kmax[i]) {
            cost += (unsigned long long)c[j] * c[j];
            big = max(big, c[j]);
        }
    }
    uint32_t carry = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        uint32_t incl = c[j];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
        c[j] = carry + incl - c[j];  // exclusive prefix inside the warp's 1024 levels
        carry += tot;
    }
    if (lane == 0) sh_warp[warp] = carry;
    __syncthreads();
    uint32_t base = 0;
    for (int w = 0; w < 8; ++w) {
        base += sh_front[w];
        if (w < warp) base += sh_warp[w];
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const uint32_t st = base + c[j];
        start[k * (kLvlCount + 1) + l0 + j * 32 + lane] = st;
        cursor[k * kLvlCount + l0 + j * 32 + lane] = st;
    }
    if (part == kLvlScanParts - 1 && threadIdx.x == 255) start[k * (kLvlCount + 1) + kLvlCount] = base + carry;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cost += __shfl_xor_sync(0xffffffffu, cost, o);
        big = max(big, __shfl_xor_sync(0xffffffffu, big, o));
    }
    if (lane == 0) {
        bool over = big > kLvlMaxBucket;
        if (cost) over = over || atomicAdd(&cost_acc[k], cost) + cost > (unsigned long long)kLvlBudgetPerRow * (unsigned long long)n;
        if (over) atomicExch(fallback, 1);
    }
}

__device__ __forceinline__ uint32_t level_field(const uint32_t *__restrict__ q, int npad, int k, int p) {
    return (q[(int64_t)(k >> 1) * npad + p] >> (16 * (k & 1))) & 0x7fffu;
}

// Positions grouped by level, mixed levels only.  grid = (npad / 256, K)
__global__ void __launch_bounds__(256)
level_scatter_kernel(const float *__restrict__ xt, const uint32_t *__restrict__ q, int npad, int n,
                     const uint32_t *__restrict__ kmin, const uint32_t *__restrict__ kmax, uint32_t *__restrict__ cursor,
                     int2 *__restrict__ items, const int *__restrict__ fallback) {
    if (*fallback) return;
    const int p = blockIdx.x * 256 + threadIdx.x, k = blockIdx.y;
    if (p >= n) return;
    const float x = xt[(int64_t)(k + 1) * npad + p];
    if (x != x) return;
    const int i = k * kLvlCount + (int)level_field(q, npad, k, p);
    if (kmin[i] == kmax[i]) return;
    items[(int64_t)k * npad + atomicAdd(&cursor[i], 1u)] = make_int2(p, __float_as_int(x));
}

// Finds the pairs (row i, column j in i's chunk or a later one) that share a level in objective k + 1 but have
// x_ik > x_jk, and whose bit the level kernel will set (all other levels <=, uid test): their bits must be cleared.
// Runs BEFORE the build kernel, while levels and items are still in the L2 (the build's 3 GB of stores evict
// them), and only records (word index << 5 | bit); level_fixup_apply_kernel clears them after the build.
// grid = (npad / 256, K); one thread per (position, objective) walks its level, four members per round so that
// their loads are in flight together.  A full list raises the fall-back flag (the build has not started yet).
// ---- packed variant for W <= 2 (M <= 5): an item carries everything the test needs -------------------------------
// items4[k][slot] = (position, float bits, level word 0, level word 1).  The find kernel then runs one thread per
// SLOT: the members of a level are the neighbouring slots, so its loads are coalesced 16-byte reads that the threads
// of a level share, instead of three scattered 4-byte gathers per member (0.10 -> 0.03 ms at N = 2^18, M = 5).
// Slots of pure levels stay at the -1 the buffer is set to before every call.
__global__ void __launch_bounds__(256)
level_scatter_packed_kernel(const float *__restrict__ xt, const uint32_t *__restrict__ q, int npad, int n, int W,
                            const uint32_t *__restrict__ kmin, const uint32_t *__restrict__ kmax,
                            uint32_t *__restrict__ cursor, int4 *__restrict__ items4, const int *__restrict__ fallback) {
    if (*fallback) return;
    const int p = blockIdx.x * 256 + threadIdx.x, k = blockIdx.y;
    if (p >= n) return;
    const float x = xt[(int64_t)(k + 1) * npad + p];
    if (x != x) return;
    const uint32_t q0 = q[p], q1 = W > 1 ? q[(int64_t)npad + p] : 0u;
    const int i = k * kLvlCount + (int)(((k >> 1 ? q1 : q0) >> (16 * (k & 1))) & 0x7fffu);
    if (kmin[i] == kmax[i]) return;
    items4[(int64_t)k * npad + atomicAdd(&cursor[i], 1u)] = make_int4(p, __float_as_int(x), (int)q0, (int)q1);
}

constexpr int kLvlFindBuf = 1024;  // entries a CTA collects in shared memory before it reserves room in the list
template <int W>
__global__ void __launch_bounds__(256)
level_fixup_find_kernel(const float *__restrict__ xt, const uint32_t *__restrict__ q, int npad, int n,
                        const uint32_t *__restrict__ kmin, const uint32_t *__restrict__ kmax,
                        const uint32_t *__restrict__ start, const int2 *__restrict__ items, const int *__restrict__ uid_row,
                        const int *__restrict__ uid_col, const float *__restrict__ tile_min1,
                        const float *__restrict__ tile_max1, TriRows rows, ChunkMap cm,
                        unsigned long long *__restrict__ list, int *__restrict__ list_col,
                        uint32_t *__restrict__ list_count, uint32_t list_cap, int *__restrict__ fallback) {
    __shared__ unsigned long long sh_list[kLvlFindBuf];
    __shared__ int sh_col[kLvlFindBuf];
    __shared__ uint32_t sh_count, sh_base;
    if (*fallback) return;
    if (threadIdx.x == 0) sh_count = 0u;
    __syncthreads();
    const int pi = blockIdx.x * 256 + threadIdx.x, k = blockIdx.y;
    bool active = pi < n;
    float xi = 0.0f;
    int lv = 0;
    if (active) {
        xi = xt[(int64_t)(k + 1) * npad + pi];
        active = xi == xi;
    }
    if (active) {
        lv = (int)level_field(q, npad, k, pi);
        const int i = k * kLvlCount + lv;
        active = kmin[i] != kmax[i];
    }
    if (active) {
        const uint32_t s0 = start[k * (kLvlCount + 1) + lv], s1 = start[k * (kLvlCount + 1) + lv + 1];
        const int2 *it = items + (int64_t)k * npad;
        const int ci = pi >> 12;
        const float r_lo = tile_min1[pi >> 10];
        for (uint32_t t0 = s0; t0 < s1; t0 += 4) {
            int2 mem[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) mem[e] = (t0 + e < s1) ? __ldg(it + t0 + e) : make_int2(pi, 0);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int pj = mem[e].x;
                // i is the row: j must sit in the same chunk (where positions carry no order) or in a later one
                if (pj == pi || (pj >> 12) < ci) continue;
                if (!(xi > __int_as_float(mem[e].y))) continue;  // floats agree with the levels: nothing to clear
                const int off = (pj >> 12) - cm.first;
                if (off < 0 || off % cm.stride != 0) continue;    // column chunk of another GPU
                uint32_t d = kLvlGuard;
#pragma unroll
                for (int w = 0; w < W; ++w) d &= (q[(int64_t)w * npad + pj] | kLvlGuard) - q[(int64_t)w * npad + pi];
                if (d != kLvlGuard || !(uid_row[pi] < uid_col[pj])) continue;  // the level kernel leaves this bit clear
                // a tile pair that the build kernels skip on their x_1 test is never written: its words hold whatever
                // the workspace held before, and must not be taken for bits that were set and counted
                if (r_lo > tile_max1[pj >> 10]) continue;
                const int64_t row_word0 = rows.tri ? rows.vstart(pi) * 4 : (int64_t)pi * rows.row_words;
                const unsigned long long word =
                    (unsigned long long)(row_word0 + (int64_t)(off / cm.stride) * kChunkWords + ((pj & 4095) >> 5));
                const unsigned long long entry = (word << 5) | (unsigned long long)(pj & 31);
                const uint32_t mine = atomicAdd(&sh_count, 1u);
                if (mine < kLvlFindBuf) {
                    sh_list[mine] = entry;
                    sh_col[mine] = pj;
                } else {  // rare: the CTA's buffer is full
                    const uint32_t slot = atomicAdd(list_count, 1u);
                    if (slot < list_cap) {
                        list[slot] = entry;
                        list_col[slot] = pj;
                    } else {
                        atomicExch(fallback, 1);
                    }
                }
            }
        }
    }
    __syncthreads();
    const uint32_t held = min(sh_count, (uint32_t)kLvlFindBuf);
    if (threadIdx.x == 0 && held) sh_base = atomicAdd(list_count, held);
    __syncthreads();
    if (held) {
        if (sh_base + held > list_cap) {
            if (threadIdx.x == 0) atomicExch(fallback, 1);
        } else {
            for (uint32_t i = threadIdx.x; i < held; i += 256) {
                list[sh_base + i] = sh_list[i];
                list_col[sh_base + i] = sh_col[i];
            }
        }
    }
}

__global__ void __launch_bounds__(256)
level_fixup_find_packed_kernel(const int4 *__restrict__ items4, int npad, const uint32_t *__restrict__ start,
                               const int *__restrict__ uid_row, const int *__restrict__ uid_col,
                               const float *__restrict__ tile_min1, const float *__restrict__ tile_max1, TriRows rows,
                               ChunkMap cm, unsigned long long *__restrict__ list, int *__restrict__ list_col,
                               uint32_t *__restrict__ list_count, uint32_t list_cap, int *__restrict__ fallback) {
    __shared__ unsigned long long sh_list[kLvlFindBuf];
    __shared__ int sh_col[kLvlFindBuf];
    __shared__ uint32_t sh_count, sh_base;
    if (*fallback) return;
    if (threadIdx.x == 0) sh_count = 0u;
    __syncthreads();
    const int slot = blockIdx.x * 256 + threadIdx.x, k = blockIdx.y;
    const int4 *it = items4 + (int64_t)k * npad;
    const int4 me = slot < npad ? it[slot] : make_int4(-1, 0, 0, 0);
    if (me.x >= 0) {
        const int pi = me.x;
        const float xi = __int_as_float(me.y);
        const uint32_t a0 = (uint32_t)me.z, a1 = (uint32_t)me.w;
        const int lv = (int)(((k >> 1 ? a1 : a0) >> (16 * (k & 1))) & 0x7fffu);
        const uint32_t s0 = start[k * (kLvlCount + 1) + lv], s1 = start[k * (kLvlCount + 1) + lv + 1];
        const int ci = pi >> 12;
        for (uint32_t t = s0; t < s1; ++t) {
            const int4 o = __ldg(it + t);
            const int pj = o.x;
            // i is the row: j must sit in the same chunk (where positions carry no order) or in a later one
            if (pj == pi || (pj >> 12) < ci) continue;
            if (!(xi > __int_as_float(o.y))) continue;      // floats agree with the levels: nothing to clear
            const uint32_t d = (((uint32_t)o.z | kLvlGuard) - a0) & (((uint32_t)o.w | kLvlGuard) - a1) & kLvlGuard;
            if (d != kLvlGuard) continue;                    // the level kernel leaves this bit clear
            const int off = (pj >> 12) - cm.first;
            if (off < 0 || off % cm.stride != 0) continue;  // column chunk of another GPU
            if (!(uid_row[pi] < uid_col[pj])) continue;
            if (tile_min1[pi >> 10] > tile_max1[pj >> 10]) continue;  // tile pair never written (see the other kernel)
            const int64_t row_word0 = rows.tri ? rows.vstart(pi) * 4 : (int64_t)pi * rows.row_words;
            const unsigned long long word =
                (unsigned long long)(row_word0 + (int64_t)(off / cm.stride) * kChunkWords + ((pj & 4095) >> 5));
            const unsigned long long entry = (word << 5) | (unsigned long long)(pj & 31);
            const uint32_t mine = atomicAdd(&sh_count, 1u);
            if (mine < kLvlFindBuf) {
                sh_list[mine] = entry;
                sh_col[mine] = pj;
            } else {  // rare: the CTA's buffer is full
                const uint32_t g = atomicAdd(list_count, 1u);
                if (g < list_cap) {
                    list[g] = entry;
                    list_col[g] = pj;
                } else {
                    atomicExch(fallback, 1);
                }
            }
        }
    }
    __syncthreads();
    const uint32_t held = min(sh_count, (uint32_t)kLvlFindBuf);
    if (threadIdx.x == 0 && held) sh_base = atomicAdd(list_count, held);
    __syncthreads();
    if (held) {
        if (sh_base + held > list_cap) {
            if (threadIdx.x == 0) atomicExch(fallback, 1);
        } else {
            for (uint32_t i = threadIdx.x; i < held; i += 256) {
                list[sh_base + i] = sh_list[i];
                list_col[sh_base + i] = sh_col[i];
            }
        }
    }
}

// Clears the recorded bits; a bit that was set takes one dominator off its column's count (the level kernel
// counted it).  A pair recorded twice (two objectives the wrong way round) is only counted down once.
__global__ void __launch_bounds__(256)
level_fixup_apply_kernel(const unsigned long long *__restrict__ list, const int *__restrict__ list_col,
                         const uint32_t *__restrict__ list_count, uint32_t list_cap, uint32_t *__restrict__ D,
                         int *__restrict__ count, const int *__restrict__ fallback) {
    if (*fallback) return;
    const uint32_t n_entries = min(*list_count, list_cap);
    for (uint32_t i = blockIdx.x * 256 + threadIdx.x; i < n_entries; i += gridDim.x * 256) {
        const unsigned long long e = list[i];
        const uint32_t bit = 1u << (uint32_t)(e & 31ull);
        if (atomicAnd(D + (e >> 5), ~bit) & bit) atomicSub(count + list_col[i], 1);
    }
}

}  // namespace evoxb
