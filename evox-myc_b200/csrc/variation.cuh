// NSGA-II variation and evaluation kernels ("next" row 2 of SURVEY.md section 8f): simulated binary
// crossover + polynomial mutation + clip, and DTLZ2 objectives -- what NSGA2.ask and the problem's
// evaluate do between two environmental selections.
//
// Replaces (formulas only; JAX's threefry streams cannot be matched, so parity is statistical and,
// given the same uniforms, exact up to powf rounding -- see oracle/variation_oracle.py):
//   src/evox/operators/crossover/sbx.py:8-41        simulated_binary(type = 1)
//   src/evox/operators/mutation/pm_mutation.py:6-54 polynomial
//   src/evox/algorithms/mo/nsga2.py:76-84           ask(): crossover -> mutation -> clip
//   src/evox/problems/numerical/dtlz.py:84-103      DTLZ2.evaluate
// Random numbers: Philox-4x32-10 (Salmon et al., SC'11), counter = (element, stream, 0, 0),
// key = (seed, generation); uniforms are the top 24 bits * 2^-24, i.e. in [0, 1).
#pragma once

#include "common.cuh"

namespace evoxb {

__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned long long p0 = 0xD2511F53ull * c0, p1 = 0xCD9E8D57ull * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__host__ __device__ inline float u01(uint32_t bits) { return (float)(bits >> 8) * (1.0f / 16777216.0f); }

// streams (second counter word)
constexpr uint32_t kStreamSbx = 1u;      // element = pair * d + j : {mu, sign bit, "keep" uniform, -}
constexpr uint32_t kStreamSbxPair = 2u;  // element = pair         : {crossover-probability uniform, -, -, -}
constexpr uint32_t kStreamPm = 3u;       // element = row * d + j  : {site uniform, mu, -, -}

__device__ __forceinline__ float pm_mutate(float v, float lo, float up, float site_u, float mu, float pro_m_over_d,
                                           float dis_m) {
    // pm_mutation.py:13-50
    v = fmaxf(fminf(v, up), lo);
    const bool site = site_u < pro_m_over_d;
    const float span = up - lo;
    if (site && mu <= 0.5f) {
        const float norm = (v - lo) / span;
        v = v + span * (powf(2.0f * mu + (1.0f - 2.0f * mu) * powf(1.0f - norm, dis_m + 1.0f), 1.0f / (dis_m + 1.0f)) - 1.0f);
    } else if (site) {
        const float norm = (up - v) / span;
        v = v + span * (1.0f - powf(2.0f * (1.0f - mu) + 2.0f * (mu - 0.5f) * powf(1.0f - norm, dis_m + 1.0f),
                                    1.0f / (dis_m + 1.0f)));
    }
    return v;
}

// One thread per (pair, dimension): both children, mutated and clipped.  pop is (n, d); rows
// [0, n/2) are first parents, rows [n/2, 2*(n/2)) second parents (sbx.py:12-13); an odd last row
// is copied through both operators' "odd" branches (sbx.py:35-36, pm_mutation.py:51-52), then clipped.
__global__ void sbx_pm_clip_kernel(const float *__restrict__ pop, int64_t n, int d, const float *__restrict__ lb,
                                   const float *__restrict__ ub, float pro_c, float dis_c, float pro_m, float dis_m,
                                   uint32_t seed, uint32_t gen, float *__restrict__ out) {
    const int64_t half = n / 2;
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t < half * d) {
        const int64_t p = t / d;
        const int j = (int)(t - p * d);
        uint32_t r[4], rp[4], m1[4], m2[4];
        philox4x32_10((uint32_t)t, kStreamSbx, (uint32_t)(t >> 32), 0u, seed, gen, r);
        philox4x32_10((uint32_t)p, kStreamSbxPair, (uint32_t)(p >> 32), 0u, seed, gen, rp);
        const float mu = u01(r[0]);
        float beta = (mu <= 0.5f) ? powf(2.0f * mu, 1.0f / (dis_c + 1.0f)) : powf(2.0f - 2.0f * mu, -1.0f / (dis_c + 1.0f));
        if (r[1] & 1u) beta = -beta;               // (-1) ** randint(0, 2)            sbx.py:19-21
        if (u01(r[2]) < 0.5f) beta = 1.0f;         //                                   sbx.py:22
        if (u01(rp[0]) > pro_c) beta = 1.0f;       // whole pair skips crossover        sbx.py:23-25
        const float a = pop[p * d + j], b = pop[(half + p) * d + j];
        float c1 = (a + b) / 2.0f + beta * (a - b) / 2.0f;  // sbx.py:27-33
        float c2 = (a + b) / 2.0f - beta * (a - b) / 2.0f;
        const float lo = lb[j], up = ub[j];
        const float pmd = pro_m / (float)d;
        const int64_t e1 = p * d + j, e2 = (half + p) * d + j;
        philox4x32_10((uint32_t)e1, kStreamPm, (uint32_t)(e1 >> 32), 0u, seed, gen, m1);
        philox4x32_10((uint32_t)e2, kStreamPm, (uint32_t)(e2 >> 32), 0u, seed, gen, m2);
        c1 = pm_mutate(c1, lo, up, u01(m1[0]), u01(m1[1]), pmd, dis_m);
        c2 = pm_mutate(c2, lo, up, u01(m2[0]), u01(m2[1]), pmd, dis_m);
        out[e1] = fmaxf(fminf(c1, up), lo);  // nsga2.py:82 clip
        out[e2] = fmaxf(fminf(c2, up), lo);
    } else if ((n & 1) && t < half * d + d) {
        const int j = (int)(t - half * d);
        out[(n - 1) * d + j] = fmaxf(fminf(pop[(n - 1) * d + j], ub[j]), lb[j]);
    }
}

// DTLZ2 (dtlz.py:84-103): g = sum_{j >= m-1} (x_j - 0.5)^2;
// f_k = (1 + g) * prod_{j < m-1-k} max(cos(x_j pi/2), 0) * (k > 0 ? sin(x_{m-1-k} pi/2) : 1).
__global__ void dtlz2_kernel(const float *__restrict__ x, int64_t n, int d, int m, float *__restrict__ f) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *xi = x + i * d;
    float g = 0.0f;
    for (int j = m - 1; j < d; ++j) g += (xi[j] - 0.5f) * (xi[j] - 0.5f);
    const float half_pi = 1.57079632679489661923f;
    const float prod = 1.0f + g;
    // forward form: cp[q] = prod_{j<q} max(cos(x_j pi/2), 0)
    float cp = 1.0f;
    for (int q = 0; q <= m - 1; ++q) {
        const int k = m - 1 - q;  // objective whose cosine product has q factors
        const float s = (k > 0) ? sinf(xi[q] * half_pi) : 1.0f;  // k > 0 multiplies sin(x_{m-1-k}) = sin(x_q)
        f[i * m + k] = prod * cp * s;
        if (q < m - 1) cp *= fmaxf(cosf(xi[q] * half_pi), 0.0f);
    }
}

// LSMOP1 (lsmop.py:127-141): linear linkage x'_j = (1 + (j + 1) / d) x_j - 10 x_0 for j >= m - 1, then for
// objective i the sphere function over its nk = 5 consecutive sub-blocks (lsmop.py:88-111; the blocks of a
// group are adjacent, so a group is one run of nk * sublen[i] variables starting at start[i] + m - 1),
// g_i = sum / sublen[i] / nk, f_k = (1 + g_k) * prod_{t < m-1-k} x_t * (k > 0 ? 1 - x_{m-1-k} : 1).
// One warp per row (a row is d floats, 4 KB at d = 1000): lanes stride over a group, shuffle-reduce per group.
constexpr int kLsmopNk = 5;
struct LsmopGroups {
    int sublen[EVOXB_MAX_OBJECTIVES];  // variables per sub-block of objective i
    int start[EVOXB_MAX_OBJECTIVES];   // lsmop.py:52 `len[i]` (offset of the group among the distance variables)
};

__global__ void __launch_bounds__(256)
lsmop1_kernel(const float *__restrict__ x, int64_t n, int d, int m, LsmopGroups gr, float *__restrict__ f) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (row >= n) return;  // warp-uniform
    const float *xi = x + row * d;
    const float x0_10 = xi[0] * 10.0f;
    const float inv_d = (float)d;
    float g_mine = 0.0f;  // lane i keeps g_i
    for (int i = 0; i < m; ++i) {
        const int first = gr.start[i] + m - 1, len = kLsmopNk * gr.sublen[i];
        float acc = 0.0f;
        for (int j = first + lane; j < first + len; j += 32) {
            const float scale = 1.0f + (float)(j + 1) / inv_d;
            const float v = scale * xi[j] - x0_10;
            acc += v * v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        const float g = acc / (float)gr.sublen[i] / (float)kLsmopNk;
        if (lane == i) g_mine = g;
    }
    // lane k writes objective k: cumulative product of the first m - 1 - k position variables
    if (lane < m) {
        const int k = lane;
        float cp = 1.0f;
        for (int t = 0; t < m - 1 - k; ++t) cp *= xi[t];
        const float tail = (k > 0) ? 1.0f - xi[m - 1 - k] : 1.0f;
        f[row * m + k] = (1.0f + g_mine) * cp * tail;
    }
}

}  // namespace evoxb
