// K4/K5: crowding distance.
//
// Replaces src/evox/operators/selection/non_dominate.py:116-156 of the reference:
//   per objective k: o = lexsort((cost_k, ~mask)); s = cost_k[o]; range = s[nv-1] - s[0];
//   d_k[o[p]] = (s[p+1] - s[p-1]) / range for interior p; d_k[o[0]] = d_k[o[nv-1]] = +inf;
//   d_k = where(mask, d_k, -inf); result = sum_k d_k.
// Only masked-in rows can end up finite or +inf, and their sorted neighbours are masked-in too
// (they occupy positions 0..nv-1), so the masked-in rows are compacted first (stable, by index)
// and only those nv rows are sorted -- in NSGA-II the mask is the last admitted front.
// The sum over objectives runs k = 0..M-1 in float32, like a sequential reduce.
#pragma once

#include "common.cuh"
#include "sort.cuh"

namespace evoxb {

constexpr int kScanTile = 4096;  // elements per CTA in the compaction kernels (256 threads x 16)

// ---- mask compaction: stage 1, per-tile population count -------------------------------------------
__global__ void __launch_bounds__(256)
mask_count_kernel(const uint8_t *__restrict__ mask, int64_t n, int *__restrict__ tile_count) {
    __shared__ int wsum[8];
    const int64_t base = (int64_t)blockIdx.x * kScanTile;
    int c = 0;
#pragma unroll
    for (int it = 0; it < 16; ++it) {
        const int64_t i = base + it * 256 + threadIdx.x;
        c += (i < n) && (mask == nullptr || mask[i] != 0);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += wsum[w];
        tile_count[blockIdx.x] = t;
    }
}

// ---- mask compaction: stage 2, stable scatter of (key, index) for every objective -------------------
// Thread t of a CTA owns the 16 consecutive elements [base + 16 t, base + 16 t + 16), so the
// compacted order is the index order.  Writes keys[k][pos] = float_to_key(cost[i][k]), vals[k][pos] = i
// and nv (total number of masked-in rows).
__global__ void __launch_bounds__(256)
mask_compact_kernel(const float *__restrict__ cost, const uint8_t *__restrict__ mask, int64_t n, int m,
                    const int *__restrict__ tile_count, int tiles, uint32_t *__restrict__ keys,
                    uint32_t *__restrict__ vals, int64_t cap, int *__restrict__ nv_out) {
    __shared__ int wsum[8];
    __shared__ int tile_base_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // prefix of the tiles before this one (tiles <= 4096 for n <= 2^24)
    int part = 0;
    for (int t = threadIdx.x; t < (int)blockIdx.x; t += 256) part += tile_count[t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) wsum[warp] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < 8; ++w) t += wsum[w];
        tile_base_s = t;
        if (blockIdx.x == (unsigned)tiles - 1) *nv_out = t + tile_count[blockIdx.x];
    }
    __syncthreads();
    const int64_t first = (int64_t)blockIdx.x * kScanTile + threadIdx.x * 16;
    uint32_t bits = 0u;
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        const int64_t i = first + e;
        if ((i < n) && (mask == nullptr || mask[i] != 0)) bits |= 1u << e;
    }
    const int mine = __popc(bits);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    __syncthreads();
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int wbase = 0;
    for (int w = 0; w < warp; ++w) wbase += wsum[w];
    int pos = tile_base_s + wbase + incl - mine;
#pragma unroll
    for (int e = 0; e < 16; ++e) {
        if (bits & (1u << e)) {
            const int64_t i = first + e;
            for (int k = 0; k < m; ++k) {
                keys[(int64_t)k * cap + pos] = float_to_key(cost[i * m + k]);
                vals[(int64_t)k * cap + pos] = (uint32_t)i;
            }
            ++pos;
        }
    }
}

// ---- K5a: neighbour gaps in sorted order, scattered back to the rows (one launch for all objectives) --
// contrib[k][i] = d_k(i) for masked-in rows (others are never read).
__global__ void __launch_bounds__(256)
crowd_gap_kernel(const float *__restrict__ cost, int m, const uint32_t *__restrict__ sorted_idx, int64_t cap,
                 const int *__restrict__ nv_dev, float *__restrict__ contrib, int64_t n) {
    const int nv = __ldcg(nv_dev);
    const int k = blockIdx.y;
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p >= nv) return;
    const uint32_t *o = sorted_idx + (int64_t)k * cap;
    const uint32_t i = o[p];
    float d;
    if (p == 0 || p == nv - 1) {
        d = __int_as_float(0x7f800000);  // +inf at both ends (:150-151)
    } else {
        const float lo = cost[(int64_t)o[0] * m + k];
        const float hi = cost[(int64_t)o[nv - 1] * m + k];
        const float range = hi - lo;                                  // :147
        const float a = cost[(int64_t)o[p + 1] * m + k];
        const float b = cost[(int64_t)o[p - 1] * m + k];
        d = __fdiv_rn(__fsub_rn(a, b), range);                        // :149, IEEE divide
    }
    contrib[(int64_t)k * n + i] = d;
}

// ---- K5b: dist[i] = mask[i] ? sum_k contrib[k][i] : -inf  (:153, :156) ---------------------------------
__global__ void __launch_bounds__(256)
crowd_sum_kernel(const float *__restrict__ contrib, const uint8_t *__restrict__ mask, int64_t n, int m,
                 float *__restrict__ dist) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float acc;
    if (mask == nullptr || mask[i] != 0) {
        acc = 0.0f;
        for (int k = 0; k < m; ++k) acc = __fadd_rn(acc, contrib[(int64_t)k * n + i]);
    } else {
        acc = __int_as_float(0xff800000);  // -inf
    }
    dist[i] = acc;
}

struct CrowdLayout {
    SortBufs sb;
    int *tile_count;
    int *nv;
    float *contrib;
    int scan_tiles;
    size_t bytes;
};

inline CrowdLayout crowd_layout(void *ws, int64_t n, int m) {
    CrowdLayout L{};
    Carver c(ws);
    const int64_t cap = round_up(n > 0 ? n : 1, 4);
    L.scan_tiles = (int)div_up(n > 0 ? n : 1, kScanTile);
    L.nv = c.take<int>(64);
    L.tile_count = c.take<int>((size_t)L.scan_tiles);
    L.sb.cap = cap;
    L.sb.tiles = (int)div_up(cap, kSortTile);
    for (int b = 0; b < 2; ++b) {
        L.sb.keys[b] = c.take<uint32_t>((size_t)cap * m);
        L.sb.vals[b] = c.take<uint32_t>((size_t)cap * m);
    }
    L.sb.hist = c.take<uint32_t>(sort_hist_words(cap, m));
    L.contrib = c.take<float>((size_t)n * m);
    L.bytes = c.total();
    return L;
}

// mask may be nullptr (all rows valid).  dist gets n floats.
inline int launch_crowding(const float *cost, const uint8_t *mask, int64_t n, int m, float *dist, const CrowdLayout &L,
                           cudaStream_t s) {
    mask_count_kernel<<<L.scan_tiles, 256, 0, s>>>(mask, n, L.tile_count);
    EVOXB_LAUNCH_CHECK("mask_count_kernel");
    mask_compact_kernel<<<L.scan_tiles, 256, 0, s>>>(cost, mask, n, m, L.tile_count, L.scan_tiles, L.sb.keys[0],
                                                     L.sb.vals[0], L.sb.cap, L.nv);
    EVOXB_LAUNCH_CHECK("mask_compact_kernel");
    int res = 0;
    if (int rc = radix_sort_pairs(L.sb, 0, m, n, L.nv, 0, 32, s, &res)) return rc;
    dim3 grid((unsigned)div_up(n, 256), (unsigned)m);
    crowd_gap_kernel<<<grid, 256, 0, s>>>(cost, m, L.sb.vals[res], L.sb.cap, L.nv, L.contrib, n);
    EVOXB_LAUNCH_CHECK("crowd_gap_kernel");
    crowd_sum_kernel<<<(unsigned)div_up(n, 256), 256, 0, s>>>(L.contrib, mask, n, m, dist);
    EVOXB_LAUNCH_CHECK("crowd_sum_kernel");
    return 0;
}

}  // namespace evoxb
