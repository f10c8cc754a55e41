// Batched stable LSD radix sort of (u32 key, u32 value) pairs -- hand-written, no CUB.
//
// Serves `jnp.lexsort` / `jnp.argsort` of the reference path
// (src/evox/operators/selection/non_dominate.py:145, src/evox/algorithms/mo/nsga2.py:91,96).
// `lax.sort` is stable, orders NaN last and treats -0 == +0; float_to_key() below gives keys with
// exactly that order and the sort is stable, so tie order (by original index) is identical.
//
// B independent segments (one per objective) of the same length are sorted by the same launches
// (blockIdx.y = segment).  The segment length may live on the device (`n_dev`): it is only known
// there when a mask has been compacted, and reading it back would cost a host sync.
// Per 8-bit pass: tile histogram -> scan over (digit, tile) -> stable scatter.
#pragma once

#include "common.cuh"

namespace evoxb {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 keys per CTA
constexpr int kRadix = 256;

// Order-preserving map float -> u32 with lax.sort semantics (NaN last and all equal, -0 == +0).
__host__ __device__ __forceinline__ uint32_t float_to_key(float f) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    union { float f; uint32_t u; } cv; cv.f = f; uint32_t u = cv.u;
#endif
    if ((u & 0x7fffffffu) > 0x7f800000u) return 0xffffffffu;  // NaN of either sign
    if (u == 0x80000000u) u = 0u;                              // -0 -> +0
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct SortBufs {
    uint32_t *keys[2];  // [B][cap] each
    uint32_t *vals[2];
    uint32_t *hist;     // [B][kRadix][tiles]
    int64_t cap;        // segment stride
    int tiles;          // ceil(cap / kSortTile)
};

__device__ __forceinline__ int64_t seg_len(int64_t n_host, const int *n_dev) {
    return n_dev ? (int64_t)__ldcg(n_dev) : n_host;
}

// ---- A: per-tile digit histogram ---------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
radix_hist_kernel(const uint32_t *__restrict__ keys, int64_t cap, int64_t n_host, const int *__restrict__ n_dev,
                  int shift, uint32_t *__restrict__ hist, int tiles) {
    __shared__ uint32_t h[kRadix];
    const int64_t n = seg_len(n_host, n_dev);
    const int seg = blockIdx.y;
    const int64_t tile0 = (int64_t)blockIdx.x * kSortTile;
    h[threadIdx.x] = 0u;
    __syncthreads();
    if (tile0 < n) {
        const uint32_t *k = keys + (int64_t)seg * cap;
#pragma unroll
        for (int it = 0; it < kSortItems; ++it) {
            const int64_t i = tile0 + it * kSortThreads + threadIdx.x;
            if (i < n) atomicAdd(&h[(k[i] >> shift) & 0xffu], 1u);
        }
    }
    __syncthreads();
    hist[((int64_t)seg * kRadix + threadIdx.x) * tiles + blockIdx.x] = h[threadIdx.x];
}

// ---- B: exclusive scan over (digit-major, tile-minor), one CTA per segment ---------------------
__global__ void __launch_bounds__(1024)
radix_scan_kernel(uint32_t *__restrict__ hist, int tiles) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry_s;
    uint32_t *h = hist + (int64_t)blockIdx.x * kRadix * tiles;
    const int total = kRadix * tiles;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0u;
    __syncthreads();
    for (int base = 0; base < total; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = (i < total) ? h[i] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_tot[lane];
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            warp_tot[lane] = wi - w;  // exclusive
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        if (i < total) h[i] = carry + warp_tot[warp] + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + warp_tot[warp] + incl;
        __syncthreads();
    }
}

// ---- C: stable scatter --------------------------------------------------------------------------
// Warp w of the CTA owns the contiguous 512 keys [w*512, w*512+512) of the tile and walks them in
// 16 rounds of 32 consecutive keys, so "earlier in the input" == (earlier warp, earlier round,
// lower lane).  __match_any_sync groups equal digits inside a round.
__global__ void __launch_bounds__(kSortThreads)
radix_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                     uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, int64_t cap, int64_t n_host,
                     const int *__restrict__ n_dev, int shift, const uint32_t *__restrict__ hist, int tiles) {
    __shared__ uint32_t wc[kSortThreads / 32][kRadix];  // per-warp running digit counts
    const int64_t n = seg_len(n_host, n_dev);
    const int seg = blockIdx.y;
    const int64_t tile0 = (int64_t)blockIdx.x * kSortTile;
    if (tile0 >= n) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t *kin = keys_in + (int64_t)seg * cap;
    const uint32_t *vin = vals_in + (int64_t)seg * cap;
    for (int t = threadIdx.x; t < (kSortThreads / 32) * kRadix; t += kSortThreads) (&wc[0][0])[t] = 0u;
    __syncthreads();

    uint32_t key[kSortItems], val[kSortItems], off[kSortItems];
    const int64_t wbase = tile0 + (int64_t)warp * (32 * kSortItems);
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int it = 0; it < kSortItems; ++it) {
        const int64_t i = wbase + it * 32 + lane;
        const bool valid = i < n;
        key[it] = valid ? kin[i] : 0xffffffffu;
        val[it] = valid ? vin[i] : 0u;
        const uint32_t d = (key[it] >> shift) & 0xffu;
        // invalid lanes get digit 256 + something unique-ish: keep them out of every valid group
        const uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : 0x100u);
        const uint32_t before = wc[warp][d];
        __syncwarp();
        if (valid && (peers & lt_mask) == 0u) wc[warp][d] = before + __popc(peers);
        __syncwarp();
        off[it] = before + __popc(peers & lt_mask);
    }
    __syncthreads();
    // exclusive prefix over warps for every digit; thread t handles digit t
    {
        uint32_t run = hist[((int64_t)seg * kRadix + threadIdx.x) * tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortThreads / 32; ++w) {
            const uint32_t c = wc[w][threadIdx.x];
            wc[w][threadIdx.x] = run;
            run += c;
        }
    }
    __syncthreads();
    uint32_t *kout = keys_out + (int64_t)seg * cap;
    uint32_t *vout = vals_out + (int64_t)seg * cap;
#pragma unroll
    for (int it = 0; it < kSortItems; ++it) {
        const int64_t i = wbase + it * 32 + lane;
        if (i < n) {
            const uint32_t d = (key[it] >> shift) & 0xffu;
            const uint32_t pos = wc[warp][d] + off[it];
            kout[pos] = key[it];
            vout[pos] = val[it];
        }
    }
}

inline size_t sort_hist_words(int64_t cap, int segments) {
    return (size_t)segments * kRadix * (size_t)div_up(cap > 0 ? cap : 1, kSortTile);
}

// Sorts bits [begin_bit, end_bit) with 8-bit passes.  Data starts in buffers [src]; returns the
// index (0/1) of the buffers holding the result.
inline int radix_sort_pairs(const SortBufs &b, int src, int segments, int64_t n_host, const int *n_dev, int begin_bit,
                            int end_bit, cudaStream_t s, int *result_idx) {
    const int64_t n_grid = n_host;  // upper bound of the segment length
    if (n_grid <= 0 || segments <= 0) {
        *result_idx = src;
        return 0;
    }
    const int tiles_used = (int)div_up(n_grid, kSortTile);
    dim3 grid((unsigned)tiles_used, (unsigned)segments);
    for (int shift = begin_bit; shift < end_bit; shift += 8) {
        radix_hist_kernel<<<dim3((unsigned)b.tiles, (unsigned)segments), kSortThreads, 0, s>>>(
            b.keys[src], b.cap, n_host, n_dev, shift, b.hist, b.tiles);
        EVOXB_LAUNCH_CHECK("radix_hist_kernel");
        radix_scan_kernel<<<segments, 1024, 0, s>>>(b.hist, b.tiles);
        EVOXB_LAUNCH_CHECK("radix_scan_kernel");
        radix_scatter_kernel<<<grid, kSortThreads, 0, s>>>(b.keys[src], b.vals[src], b.keys[src ^ 1], b.vals[src ^ 1],
                                                          b.cap, n_host, n_dev, shift, b.hist, b.tiles);
        EVOXB_LAUNCH_CHECK("radix_scatter_kernel");
        src ^= 1;
    }
    *result_idx = src;
    return 0;
}

}  // namespace evoxb
