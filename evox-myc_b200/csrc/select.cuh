// NSGA-II survivor selection tail.
//
// Replaces src/evox/algorithms/mo/nsga2.py:91-98 of the reference (same block in gde3.py:88-97,
// im_moea.py:131-141):
//   order = argsort(rank); worst_rank = rank[order[pop_size]]; mask = rank == worst_rank
//   crowding_dis = crowding_distance(merged_fitness, mask)
//   combined_order = lexsort((-crowding_dis, rank))[:pop_size]; survivor = merged_pop[combined_order]
// worst_rank only needs the value of the pop_size-th smallest rank, so a rank histogram replaces
// the argsort.  The lexsort is an LSD radix sort: first by the key of -crowding (NaN last), then
// by rank; both stable, so ties resolve by index exactly like lax.sort.
#pragma once

#include "common.cuh"
#include "sort.cuh"

namespace evoxb {

__global__ void rank_hist_kernel(const int *__restrict__ rank, int64_t n, int *__restrict__ hist) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) {
        const int r = rank[i];
        if ((unsigned)r < (unsigned)n) atomicAdd(hist + r, 1);  // ranks outside [0, n) (a failed sharded call: -1) are skipped
    }
}

// One CTA: smallest r with hist[0] + ... + hist[r] > k  (the value of the k-th smallest rank, 0-based).
__global__ void __launch_bounds__(1024)
kth_rank_kernel(const int *__restrict__ hist, int64_t n, int64_t k, int *__restrict__ out) {
    __shared__ long long warp_tot[32];
    __shared__ long long carry_s;
    __shared__ int found_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { carry_s = 0; found_s = -1; }
    __syncthreads();
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const long long v = (i < n) ? hist[i] : 0;
        long long incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            long long w = warp_tot[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            warp_tot[lane] = wi - w;
        }
        __syncthreads();
        const long long pre = carry_s + warp_tot[warp] + incl;  // inclusive prefix at i
        if (i < n && pre > k && pre - v <= k) found_s = (int)i;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = pre;
        __syncthreads();
        if (found_s >= 0) break;
    }
    if (threadIdx.x == 0) *out = found_s;
}

__global__ void rank_mask_kernel(const int *__restrict__ rank, int64_t n, const int *__restrict__ worst,
                                 uint8_t *__restrict__ mask) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) mask[i] = rank[i] == __ldcg(worst);
}

__global__ void crowd_keys_kernel(const float *__restrict__ x, int64_t n, uint32_t *__restrict__ keys,
                                  uint32_t *__restrict__ vals) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) {
        keys[i] = float_to_key(x[i]);
        vals[i] = (uint32_t)i;
    }
}

// keys[i] = order of (-cd[i]) for the int array cd = argsort(...): ascending -cd == ascending (n-1-cd)
__global__ void neg_index_keys_kernel(const uint32_t *__restrict__ cd, int64_t n, uint32_t *__restrict__ keys,
                                      uint32_t *__restrict__ vals) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) {
        keys[i] = (uint32_t)(n - 1) - cd[i];
        vals[i] = (uint32_t)i;
    }
}

__global__ void gather_rank_keys_kernel(const int *__restrict__ rank, const uint32_t *__restrict__ vals, int64_t n,
                                        uint32_t *__restrict__ keys) {
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p < n) keys[p] = (uint32_t)rank[vals[p]];
}

__global__ void take_prefix_kernel(const uint32_t *__restrict__ vals, int64_t k, int *__restrict__ order) {
    const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (p < k) order[p] = (int)vals[p];
}

// dst[r, :] = src[idx[r], :]   (survivor = merged_pop[combined_order], nsga2.py:97-98)
__global__ void gather_rows_kernel(const float *__restrict__ src, int64_t d, const int *__restrict__ idx, int64_t k,
                                   float *__restrict__ dst) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= k * d) return;
    const int64_t r = t / d, c = t - r * d;
    dst[t] = src[(int64_t)idx[r] * d + c];
}

inline int bits_for(int64_t n) {  // number of bits needed to represent values in [0, n)
    int b = 1;
    while (((int64_t)1 << b) < n) ++b;
    return b;
}

}  // namespace evoxb
