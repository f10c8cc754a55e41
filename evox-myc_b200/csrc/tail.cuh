// The NSGA-II selection tail in ONE cooperative launch: cut-off rank, mask, crowding distance of the
// cut-off front, survivor order.
//
// Replaces src/evox/algorithms/mo/nsga2.py:91-96 of the reference
//     order = argsort(rank); worst_rank = rank[order[pop_size]]; mask = rank == worst_rank
//     crowding_dis = crowding_distance(merged_fitness, mask)      (non_dominate.py:116-156)
//     combined_order = lexsort((-crowding_dis, rank))[:pop_size]
// with the same results as the multi-kernel path (select_impl variant 0), which it supersedes for
// speed: ~45 small launches become one kernel whose phases are separated by grid-wide syncs.
//
// How the lexsort is decomposed (exact, including tie order):
//   * rows of rank < worst all carry crowding -inf, so inside a rank they keep index order:
//     part A = rows with rank < worst, stably sorted by rank (radix passes over the bits of `worst`);
//   * rows of rank == worst are ordered by (-crowding, index), NaN last:
//     part B = the compacted cut-off front, stably sorted by the key of -crowding;
//   * rows of rank > worst can never reach the first pop_size positions (rank[order[pop_size]] ==
//     worst means at least pop_size + 1 rows have rank <= worst);
//   combined_order[:pop_size] = (A ++ B)[:pop_size].
// Crowding: the cut-off front is compacted in index order, each objective is sorted by its key
// (stable, so ties keep index order like lexsort((cost, ~mask))), gaps use the original floats.
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"
#include "coop_sort.cuh"

namespace evoxb {

struct TailArgs {
    const float *fitness;  // (n, m)
    const int *rank;       // (n)
    int n, m;
    int kth;               // index into the sorted ranks that defines the cut-off (pop_size for NSGA2.tell)
    int pop;               // number of survivors to emit
    int *order;            // out (pop)
    float *crowd;          // out (n)
    // workspace
    int *hist;             // (n) zeroed by the caller before the launch
    int *tile_a;           // (tiles) rows of rank < worst per tile
    int *tile_b;           // (tiles) rows of rank == worst per tile
    uint32_t *list_a;      // (n) rows of rank < worst, index order
    uint32_t *list_b;      // (n) rows of rank == worst, index order (the compacted mask)
    uint32_t *buf0;        // (max(m,1) * cap) index ping-pong
    uint32_t *buf1;
    uint32_t *hist_g;      // (max(m,1) * tiles * 256)
    float *contrib;        // (m * n)
    int64_t cap;           // segment stride of buf0/buf1 (>= n)
    int *info;             // out: [0] worst rank, [1] rows before it, [2] rows in it
};

__global__ void __launch_bounds__(256)
select_tail_kernel(TailArgs a) {
    namespace cgx = cooperative_groups;
    cgx::grid_group grid = cgx::this_grid();
    __shared__ CoopSortSmem csm;
    __shared__ long long sh_scan[8];
    __shared__ long long sh_carry;
    __shared__ int sh_worst, sh_before, sh_nv;
    __shared__ int sh_wa[8], sh_wb[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = a.n, m = a.m;
    const int tiles = (n + kCoopTile - 1) / kCoopTile;

    // ---- 1. rank histogram (one atomic per distinct rank per warp) ----
    for (int i = blockIdx.x * 256 + threadIdx.x; i < tiles * kCoopTile; i += gridDim.x * 256) {
        const int r = (i < n) ? a.rank[i] : -1;
        const uint32_t peers = __match_any_sync(0xffffffffu, r);
        if (r >= 0 && (peers & ((1u << lane) - 1u)) == 0u) atomicAdd(a.hist + r, __popc(peers));
    }
    grid.sync();

    // ---- 2. every CTA finds the cut-off rank itself: smallest r with hist[0..r] > kth ----
    {
        if (threadIdx.x == 0) { sh_carry = 0; sh_worst = -1; }
        __syncthreads();
        for (int base = 0; base < n; base += 256) {
            const int i = base + threadIdx.x;
            const long long v = (i < n) ? __ldcg(a.hist + i) : 0;
            long long incl = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            if (lane == 31) sh_scan[warp] = incl;
            __syncthreads();
            long long wb = sh_carry;
            for (int q = 0; q < warp; ++q) wb += sh_scan[q];
            const long long pre = wb + incl;  // inclusive prefix at i
            if (i < n && pre > a.kth && pre - v <= a.kth) {
                sh_worst = i;
                sh_before = (int)(pre - v);
                sh_nv = (int)v;
            }
            __syncthreads();
            if (threadIdx.x == 255) sh_carry = pre;
            __syncthreads();
            if (sh_worst >= 0) break;
        }
    }
    const int worst = sh_worst, before = sh_before, nv = sh_nv;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        a.info[0] = worst;
        a.info[1] = before;
        a.info[2] = nv;
    }

    // ---- 3. per-tile counts of part A (rank < worst) and part B (rank == worst) ----
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int ca = 0, cb = 0;
#pragma unroll
        for (int it = 0; it < kCoopItems; ++it) {
            const int i = tile * kCoopTile + it * 256 + threadIdx.x;
            const int r = (i < n) ? a.rank[i] : 0x7fffffff;
            ca += r < worst;
            cb += r == worst;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            ca += __shfl_xor_sync(0xffffffffu, ca, o);
            cb += __shfl_xor_sync(0xffffffffu, cb, o);
        }
        __syncthreads();
        if (lane == 0) { sh_wa[warp] = ca; sh_wb[warp] = cb; }
        __syncthreads();
        if (threadIdx.x == 0) {
            int ta = 0, tb = 0;
            for (int q = 0; q < 8; ++q) { ta += sh_wa[q]; tb += sh_wb[q]; }
            a.tile_a[tile] = ta;
            a.tile_b[tile] = tb;
        }
    }
    grid.sync();

    // ---- 4. stable compaction of both parts (thread t owns 8 consecutive rows of its tile) ----
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int pa = 0, pb = 0;
        for (int t = threadIdx.x; t < tile; t += 256) { pa += __ldcg(a.tile_a + t); pb += __ldcg(a.tile_b + t); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            pa += __shfl_xor_sync(0xffffffffu, pa, o);
            pb += __shfl_xor_sync(0xffffffffu, pb, o);
        }
        __syncthreads();
        if (lane == 0) { sh_wa[warp] = pa; sh_wb[warp] = pb; }
        __syncthreads();
        int base_a = 0, base_b = 0;
        for (int q = 0; q < 8; ++q) { base_a += sh_wa[q]; base_b += sh_wb[q]; }
        __syncthreads();
        const int first = tile * kCoopTile + threadIdx.x * kCoopItems;
        uint32_t ma = 0u, mb = 0u;
#pragma unroll
        for (int e = 0; e < kCoopItems; ++e) {
            const int i = first + e;
            const int r = (i < n) ? a.rank[i] : 0x7fffffff;
            if (r < worst) ma |= 1u << e;
            if (r == worst) mb |= 1u << e;
        }
        int ia = __popc(ma), ib = __popc(mb);
        const int mine_a = ia, mine_b = ib;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int ta = __shfl_up_sync(0xffffffffu, ia, o), tb = __shfl_up_sync(0xffffffffu, ib, o);
            if (lane >= o) { ia += ta; ib += tb; }
        }
        if (lane == 31) { sh_wa[warp] = ia; sh_wb[warp] = ib; }
        __syncthreads();
        for (int q = 0; q < warp; ++q) { base_a += sh_wa[q]; base_b += sh_wb[q]; }
        int pos_a = base_a + ia - mine_a, pos_b = base_b + ib - mine_b;
#pragma unroll
        for (int e = 0; e < kCoopItems; ++e) {
            if (ma & (1u << e)) a.list_a[pos_a++] = (uint32_t)(first + e);
            if (mb & (1u << e)) a.list_b[pos_b++] = (uint32_t)(first + e);
        }
        __syncthreads();
    }
    grid.sync();

    // ---- 5. crowding: every objective sorts the cut-off front by its key (4 passes, m segments) ----
    const float *fit = a.fitness;
    const uint32_t *src = a.list_b;
    bool shared_src = true;
    uint32_t *dst = a.buf0;
    if (nv > 0) {
        for (int shift = 0; shift < 32; shift += 8) {
            auto key = [=](int seg, uint32_t row) -> uint32_t { return float_to_key(fit[(int64_t)row * m + seg]); };
            coop_radix_pass(grid, csm, key, m, nv, a.cap, src, shared_src, dst, a.hist_g, shift);
            src = dst;
            shared_src = false;
            dst = (dst == a.buf0) ? a.buf1 : a.buf0;
        }
        // ---- 6. neighbour gaps (non_dominate.py:147-151) ----
        for (int64_t t = blockIdx.x * 256 + threadIdx.x; t < (int64_t)m * nv; t += (int64_t)gridDim.x * 256) {
            const int k = (int)(t / nv), p = (int)(t - (int64_t)k * nv);
            const uint32_t *o = src + (int64_t)k * a.cap;
            const uint32_t i = __ldcg(o + p);
            float d;
            if (p == 0 || p == nv - 1) {
                d = __int_as_float(0x7f800000);
            } else {
                const float lo = fit[(int64_t)__ldcg(o) * m + k], hi = fit[(int64_t)__ldcg(o + nv - 1) * m + k];
                const float up = fit[(int64_t)__ldcg(o + p + 1) * m + k], dn = fit[(int64_t)__ldcg(o + p - 1) * m + k];
                d = __fdiv_rn(__fsub_rn(up, dn), __fsub_rn(hi, lo));
            }
            a.contrib[(int64_t)k * n + i] = d;
        }
    }
    grid.sync();
    // ---- 7. crowding distance of every row (sum over objectives in order; -inf outside the mask) ----
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
        float acc;
        if (a.rank[i] == worst) {
            acc = 0.0f;
            for (int k = 0; k < m; ++k) acc = __fadd_rn(acc, __ldcg(a.contrib + (int64_t)k * n + i));
        } else {
            acc = __int_as_float(0xff800000);
        }
        a.crowd[i] = acc;
    }
    grid.sync();

    // ---- 8. part B: the cut-off front by (-crowding, index) ----
    const uint32_t *src_b = a.list_b;
    if (nv > 0 && before < a.pop) {
        uint32_t *db = a.buf0;
        const float *crowd = a.crowd;
        for (int shift = 0; shift < 32; shift += 8) {
            auto key = [=](int, uint32_t row) -> uint32_t { return float_to_key(-__ldcg(crowd + row)); };
            coop_radix_pass(grid, csm, key, 1, nv, a.cap, src_b, false, db, a.hist_g, shift);
            src_b = db;
            db = (db == a.buf0) ? a.buf1 : a.buf0;
        }
    }
    // ---- 9. part A: rows of rank < worst by (rank, index); index buffers of segment 1 are free now ----
    const uint32_t *src_a = a.list_a;
    if (before > 0 && worst > 1) {
        uint32_t *a0 = a.buf0 + a.cap, *a1 = a.buf1 + a.cap;  // segment 1 (the caller provides >= 2 segments)
        uint32_t *da = a0;
        const int *rank = a.rank;
        for (int shift = 0; (worst - 1) >> shift; shift += 8) {
            auto key = [=](int, uint32_t row) -> uint32_t { return (uint32_t)rank[row]; };
            coop_radix_pass(grid, csm, key, 1, before, a.cap, src_a, false, da, a.hist_g, shift);
            src_a = da;
            da = (da == a0) ? a1 : a0;
        }
    }
    // ---- 10. combined_order[:pop] = (A ++ B)[:pop] ----
    for (int p = blockIdx.x * 256 + threadIdx.x; p < a.pop; p += gridDim.x * 256)
        a.order[p] = (int)(p < before ? __ldcg(src_a + p) : __ldcg(src_b + (p - before)));
}

}  // namespace evoxb
