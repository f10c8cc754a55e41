// K2/K3: dominator counts and the front-peeling loop over the bit-packed dominance matrix.
//
// Replaces src/evox/operators/selection/non_dominate.py:61 (dominate_count = sum(D, axis=0)) and
// the while_loop at :72-111 (same loop as host_rank_from_domination_matrix, :11-27) of the
// reference.  The reference's GPU formulation re-reads the whole N x N byte matrix once per front;
// here the packed row of every individual is streamed exactly once, when it enters a front.
//
// Column sums over a set of packed rows are a "positional popcount": every lane keeps bit-sliced
// (vertical) counters for the words it owns and feeds them through a carry-save-adder tree, so
// counting costs a few LOP3 per 32 pairs instead of one add per pair.  The eight warps of a CTA
// then add their bit-sliced counters in shared memory, expand them once to integers and apply
// them to count[] with one atomicAdd per touched column.
#pragma once

#include <cooperative_groups.h>

#include "common.cuh"

namespace evoxb {
namespace cg = cooperative_groups;

constexpr int kPeelThreads = 256;
constexpr int kPeelWarps = kPeelThreads / 32;
constexpr int kChunkWords = 128;          // one warp-load: 32 lanes x uint4 = 512 B = 4096 columns
constexpr int kChunkCols = kChunkWords * 32;
constexpr int kBatch = 8;                 // rows in flight per warp (8 x 512 B)
constexpr int kPlanesWarp = 12;           // a warp accumulates at most 2^12 - 8 rows between flushes
constexpr int kPlanes = kPlanesWarp + 3;  // after adding the 8 warps of a CTA
constexpr int kWarpRowsCap = (1 << kPlanesWarp) - kBatch;


// sum (low) and carry (high) of three bit-vectors
__device__ __forceinline__ void csa(uint32_t &hi, uint32_t &lo, uint32_t a, uint32_t b, uint32_t c) {
    lo = lop3_xor3(a, b, c);
    hi = lop3_maj(a, b, c);
}

// Adds eight one-bit-per-position words into a bit-sliced counter (Harley-Seal step).
__device__ __forceinline__ void vc_add8(uint32_t (&pl)[kPlanes], const uint32_t (&w)[kBatch]) {
    uint32_t c0, s0, c1, s1, c2, s2, c3, d0, t0, d1, e0;
    csa(c0, s0, w[0], w[1], w[2]);
    csa(c1, s1, w[3], w[4], w[5]);
    csa(c2, s2, w[6], w[7], pl[0]);
    csa(c3, pl[0], s0, s1, s2);
    csa(d0, t0, c0, c1, c2);
    csa(d1, pl[1], t0, c3, pl[1]);
    csa(e0, pl[2], d0, d1, pl[2]);
    uint32_t carry = e0;
#pragma unroll
    for (int p = 3; p < kPlanesWarp; ++p) {
        const uint32_t t = pl[p] & carry;
        pl[p] ^= carry;
        carry = t;
    }
}

// Shared memory used by one CTA for the cross-warp reduction and the expansion.
constexpr int kStageRows = 1024;  // member rows resolved per staging round

// Band of objective 1 of every 1024-position tile (presort.cuh).  A (row, 1024-column tile) segment of
// the matrix is all zero when the row tile's smallest x_1 exceeds the column tile's largest -- the
// build kernels leave it UNWRITTEN on that very test -- so count and peel must not, and do not, fetch those
// 128-byte quarters of a row chunk (predicated loads: no branch, the batch stays in flight).  Every load of
// the presorted matrix has to carry this predicate.
struct TileBands {
    const float *tmin;  // nullptr: no band information (unsorted full-matrix mode)
    const float *tmax;
};

struct PeelSmem {
    uint32_t xchg[kPeelWarps / 2][4 * kPlanes][32];  // up to 4 senders per round
    union {
        uint32_t fin[kPlanes][kChunkWords];          // CTA totals, plane-major (expansion)
        struct {
            int rows[kStageRows];                    // where the members' packed rows start (TriRows::vstart, in uint4)
            float rowmin[kStageRows];                // smallest x_1 of each such row's 1024-tile
        } st;
    };
};

// ---- work decomposition -----------------------------------------------------------------------
// Positions are grouped in regions of 4096 (= one column chunk).  A front's members are stored
// region-major: region r keeps its members (any order) in list[r * 4096 ...] and their number in
// cnt[r].  For column chunk c only members that can have bits there are streamed:
//   triangular matrix (presorted population): regions 0..c   (row i has no bits left of column i)
//   full matrix:                               all regions
// pc[r]  = number of members in regions < r          (prefix of cnt)
// S[c]   = number of (chunk, member) segments in chunks < c
// The segment space [0, S[nchunks]) is split evenly over the CTAs, chunk-major, so a CTA touches
// at most two chunks when the grid has at least nchunks CTAs.
constexpr int kMaxChunks = 512;  // N <= 2M rows; beyond that the matrix would not fit in HBM anyway

struct PrefixSmem {
    int pc[kMaxChunks + 1];
    long long S[kMaxChunks + 1];
};

// Order in which the local chunks are laid out in the segment space.  With a triangular matrix chunk
// c carries (c + 1) / nchunks of the members, so small and large chunks alternate (0, last, 1,
// last-1, ...): a CTA's share then never spans a run of tiny chunks (each chunk costs a flush).
__device__ __forceinline__ int chunk_order(int i, int nlocal, bool tri) {
    return tri ? ((i & 1) ? nlocal - 1 - (i >> 1) : (i >> 1)) : i;
}

// cnt == nullptr: every row 0..n-1 is a member (dominator count, K2).
// pc[] is indexed by region; S[] by position in chunk_order() over the LOCAL chunks of `cm`.
// kPer = entries per lane (compile-time: a run-time bound here makes ptxas keep ~100 extra registers
// alive across the persistent front loop).
template <int kPer>
__device__ __forceinline__ void build_prefix_t(const int *__restrict__ cnt, int nchunks, int64_t n, bool tri, ChunkMap cm,
                                               PrefixSmem &ps) {
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        constexpr int per = kPer;
        int v[kPer];
        int run = 0;
#pragma unroll
        for (int e = 0; e < kPer; ++e) {
            if (e < per) {
                const int r = lane * per + e;
                int c = 0;
                if (r < nchunks) {
                    if (cnt) c = __ldcg(cnt + r);
                    else c = (int)max((int64_t)0, min((int64_t)kChunkCols, n - (int64_t)r * kChunkCols));
                }
                run += c;
                v[e] = run;
            }
        }
        int incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        const int before = incl - run;
        if (lane == 0) ps.pc[0] = 0;
#pragma unroll
        for (int e = 0; e < kPer; ++e)
            if (e < per) ps.pc[lane * per + e + 1] = before + v[e];
        __syncwarp();
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        constexpr int lper = kPer;
        long long w[kPer];
        long long lrun = 0;
#pragma unroll
        for (int e = 0; e < kPer; ++e) {
            if (e < lper) {
                const int i = lane * lper + e;
                long long len = 0;
                if (i < cm.nlocal)
                    len = tri ? (long long)ps.pc[min(cm.global(chunk_order(i, cm.nlocal, true)), nchunks - 1) + 1]
                              : (long long)total;
                lrun += len;
                w[e] = lrun;
            }
        }
        long long lincl = lrun;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, lincl, o);
            if (lane >= o) lincl += t;
        }
        const long long lbefore = lincl - lrun;
        if (lane == 0) ps.S[0] = 0;
#pragma unroll
        for (int e = 0; e < kPer; ++e)
            if (e < lper) ps.S[lane * lper + e + 1] = lbefore + w[e];
    }
    __syncthreads();
}

__device__ __forceinline__ void build_prefix(const int *__restrict__ cnt, int nchunks, int64_t n, bool tri, ChunkMap cm,
                                             PrefixSmem &ps) {
    if (nchunks <= 32) build_prefix_t<1>(cnt, nchunks, n, tri, cm, ps);        // N <= 131072
    else if (nchunks <= 128) build_prefix_t<4>(cnt, nchunks, n, tri, cm, ps);  // N <= 524288
    else build_prefix_t<kMaxChunks / 32>(cnt, nchunks, n, tri, cm, ps);
}

// Row index of member number t (in region-major order) -- binary search in pc[], then the list.
__device__ __forceinline__ int member_row(const int *__restrict__ list, const PrefixSmem &ps, int nregions, int64_t t) {
    if (!list) return (int)t;
    int lo = 0, hi = nregions;  // pc[lo] <= t < pc[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if ((int64_t)ps.pc[mid] <= t) lo = mid; else hi = mid;
    }
    return __ldcg(list + (int64_t)lo * kChunkCols + (t - ps.pc[lo]));
}

// Streams members t0..t1 (region-major numbering) restricted to the 4096-column chunk `chunk` and
// adds sign * (column sums) to count[].
__device__ void accumulate_chunk(const uint32_t *__restrict__ D, TriRows rows, const int *__restrict__ list,
                                 const PrefixSmem &ps, int nregions, int64_t t0, int64_t t1, int local_chunk, int chunk,
                                 int *__restrict__ count, int sign, PeelSmem &sm, TileBands tb) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const uint4 *base = reinterpret_cast<const uint4 *>(D + (int64_t)local_chunk * kChunkWords) + lane;
    // this lane's 128 columns lie in column tile chunk * 4 + lane / 8; +inf = "always fetch"
    const float col_max = tb.tmax ? tb.tmax[chunk * 4 + (lane >> 3)] : __int_as_float(0x7f800000);

    for (int64_t f0 = t0; f0 < t1; f0 += (int64_t)kWarpRowsCap * kPeelWarps) {
        const int64_t f1 = min(t1, f0 + (int64_t)kWarpRowsCap * kPeelWarps);
        uint32_t pl[4][kPlanes];
#pragma unroll
        for (int s = 0; s < 4; ++s)
#pragma unroll
            for (int p = 0; p < kPlanes; ++p) pl[s][p] = 0u;

        for (int64_t g0 = f0; g0 < f1; g0 += kStageRows) {
            const int gn = (int)min((int64_t)kStageRows, f1 - g0);
            // all threads resolve member -> row in parallel (one list-load latency per round)
            for (int q = threadIdx.x; q < gn; q += kPeelThreads) {
                const int row = member_row(list, ps, nregions, g0 + q);
                sm.st.rows[q] = (int)rows.vstart(row);  // < 2^31 uint4 for every matrix that fits in HBM
                sm.st.rowmin[q] = tb.tmin ? tb.tmin[row >> 10] : 0.0f;
            }
            __syncthreads();
            // warp `warp` takes staged members warp, warp + 8, ... in batches of kBatch
            for (int q = warp; q < gn; q += kPeelWarps * kBatch) {
                uint4 v[kBatch];
#pragma unroll
                for (int b = 0; b < kBatch; ++b) {
                    const int qq = q + b * kPeelWarps;
                    // a NaN band (a tile of NaN rows only) compares false and is fetched
                    const int qc = min(qq, gn - 1);
                    const bool fetch = qq < gn && !(sm.st.rowmin[qc] > col_max);
                    v[b] = ldg_stream_v4_if(base + (int64_t)sm.st.rows[qc], fetch);
                }
                uint32_t w[kBatch];
#pragma unroll
                for (int b = 0; b < kBatch; ++b) w[b] = v[b].x;
                vc_add8(pl[0], w);
#pragma unroll
                for (int b = 0; b < kBatch; ++b) w[b] = v[b].y;
                vc_add8(pl[1], w);
#pragma unroll
                for (int b = 0; b < kBatch; ++b) w[b] = v[b].z;
                vc_add8(pl[2], w);
#pragma unroll
                for (int b = 0; b < kBatch; ++b) w[b] = v[b].w;
                vc_add8(pl[3], w);
            }
            __syncthreads();  // rows[] is reused by the next round and aliases fin[]
        }

        // ---- add the eight warps' bit-sliced counters (tree: 8 -> 4 -> 2 -> 1) ----
#pragma unroll
        for (int stride = 1; stride < kPeelWarps; stride <<= 1) {
            const bool sender = (warp & (2 * stride - 1)) == stride;
            const bool receiver = (warp & (2 * stride - 1)) == 0;
            const int slot = warp / (2 * stride);
            __syncthreads();
            if (sender) {
#pragma unroll
                for (int s = 0; s < 4; ++s)
#pragma unroll
                    for (int p = 0; p < kPlanes; ++p)
                        sm.xchg[slot][s * kPlanes + p][lane] = pl[s][p];
            }
            __syncthreads();
            if (receiver) {
#pragma unroll
                for (int s = 0; s < 4; ++s) {
                    uint32_t carry = 0u;
#pragma unroll
                    for (int p = 0; p < kPlanes; ++p) {
                        const uint32_t b = sm.xchg[slot][s * kPlanes + p][lane];
                        const uint32_t a = pl[s][p];
                        pl[s][p] = lop3_xor3(a, b, carry);
                        carry = lop3_maj(a, b, carry);
                    }
                }
            }
        }
        if (warp == 0) {
#pragma unroll
            for (int s = 0; s < 4; ++s)
#pragma unroll
                for (int p = 0; p < kPlanes; ++p) sm.fin[p][lane * 4 + s] = pl[s][p];
        }
        __syncthreads();

        // ---- expand: thread -> (word, 16-bit half), one atomic per non-zero column ----
        {
            const int wd = threadIdx.x & (kChunkWords - 1);
            const int half = threadIdx.x >> 7;  // 256 threads = 128 words x 2 halves
            uint32_t any = 0u;
            uint32_t pw[kPlanes];
#pragma unroll
            for (int p = 0; p < kPlanes; ++p) {
                pw[p] = sm.fin[p][wd] >> (16 * half);
                any |= pw[p];
            }
            any &= 0xffffu;
            int *cbase = count + (int64_t)chunk * kChunkCols + wd * 32 + 16 * half;
            while (any) {
                const int bit = __ffs(any) - 1;
                any &= any - 1;
                int val = 0;
#pragma unroll
                for (int p = 0; p < kPlanes; ++p) val |= (int)((pw[p] >> bit) & 1u) << p;
                atomicAdd(cbase + bit, sign * val);
            }
        }
        __syncthreads();
    }
}

// One front (or, with list == nullptr, all rows): the caller has run build_prefix().
__device__ void accumulate_front(const uint32_t *__restrict__ D, TriRows rows, int nchunks, bool tri, ChunkMap cm,
                                 const int *__restrict__ list, const PrefixSmem &ps, int *__restrict__ count, int sign,
                                 PeelSmem &sm, TileBands tb) {
    const long long total = ps.S[cm.nlocal];
    const long long s0 = total * blockIdx.x / gridDim.x;
    const long long s1 = total * (blockIdx.x + 1) / gridDim.x;
    if (s0 >= s1) return;
    int lo = 0, hi = cm.nlocal;  // S[lo] <= s0 < S[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ps.S[mid] <= s0) lo = mid; else hi = mid;
    }
    long long s = s0;
    int i = lo;  // position in chunk_order()
    while (s < s1) {
        const long long len = ps.S[i + 1] - ps.S[i];
        const long long t0 = s - ps.S[i];
        const long long t1 = min(len, t0 + (s1 - s));
        const int lc = chunk_order(i, cm.nlocal, tri);
        const int chunk = cm.global(lc);
        if (t1 > t0)
            accumulate_chunk(D, rows, list, ps, tri ? chunk + 1 : nchunks, t0, t1, lc, chunk, count, sign, sm, tb);
        s += t1 - t0;
        ++i;
    }
}

// K2: count[j] += number of rows in [0, n_rows) that dominate j   (non_dominate.py:61)
__global__ void __launch_bounds__(kPeelThreads)
dominator_count_kernel(const uint32_t *__restrict__ D, TriRows rows, int nchunks, int64_t n_rows, int tri,
                       ChunkMap cm, int *__restrict__ count, TileBands tb, const int *__restrict__ go) {
    __shared__ PeelSmem sm;
    __shared__ PrefixSmem ps;
    // go != nullptr: the level build has already counted the dominators while it built the matrix (levels.cuh);
    // this pass only runs when that build was abandoned on the device for the float kernel
    if (go && *go == 0) return;
    build_prefix(nullptr, nchunks, n_rows, tri != 0, cm, ps);
    accumulate_front(D, rows, nchunks, tri != 0, cm, nullptr, ps, count, +1, sm, tb);
}

// Front detection: every j < n with count[j] == 0 joins the front, gets its rank, and drops to -1
// for good (the reference's "count -= front" trick, non_dominate.py:103).  Members are appended to
// their region's slice of the list; a warp covers 32 consecutive positions, i.e. one region.
__device__ void collect_front(int *__restrict__ count, int n, int r, int *__restrict__ rank,
                              const int *__restrict__ perm, int *__restrict__ list, int *__restrict__ cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n32 = round_up(n, 32);
    constexpr int kU = 4;  // counts fetched together: one L2 latency for kU positions
    for (int64_t j0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j0 < n32; j0 += kU * stride) {
        int v[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const int64_t j = j0 + u * stride;
            v[u] = (j < n) ? __ldcg(count + j) : -1;  // count is updated by L2 atomics
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const int64_t j = j0 + u * stride;
            if (j >= n32) break;  // warp-uniform: stride and n32 are multiples of 32
            const bool in = v[u] == 0;
            const uint32_t m = __ballot_sync(0xffffffffu, in);
            if (m) {
                const int region = (int)(j >> 12);
                int base = 0;
                if (lane == 0) base = atomicAdd(cnt + region, __popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (in) {
                    list[(int64_t)region * kChunkCols + base + __popc(m & ((1u << lane) - 1u))] = (int)j;
                    rank[perm ? perm[j] : j] = r;
                    count[j] = -1;
                }
            }
        }
    }
}

struct PeelCtrl {
    int cnt[2][kMaxChunks];  // members per region of the current / next front (double-buffered)
    int fronts_done;         // number of fronts peeled (written at exit)
    int front0;              // size of front 0 (front0_size_kernel); only set when both peel kernels are enqueued
    int presort_slow;        // raised by the presort when a run of equal key_0 is too long for its fast path
    int level_fallback;      // raised by level_scan_kernel: the level build is skipped, the float kernel runs (levels.cuh)
};

// K3: the whole peel loop in one cooperative launch (grid-wide sync twice per front).
template <int kPer>
__global__ void __launch_bounds__(kPeelThreads, 2)
peel_persistent_kernel(const uint32_t *__restrict__ D, TriRows rows, int nchunks, int n, int tri,
                       int *__restrict__ count, int *__restrict__ rank, const int *__restrict__ perm,
                       int *__restrict__ list, PeelCtrl *__restrict__ ctrl, TileBands tb, int front0_limit) {
    __shared__ PeelSmem sm;
    __shared__ PrefixSmem ps;
    cg::grid_group grid = cg::this_grid();
    // front0_limit >= 0: the data-flow kernel (peel_flow.cuh) was enqueued in front of this one and did the work
    // unless front 0 is larger than the limit (uniform over the grid, before any grid-wide sync)
    if (front0_limit >= 0 && __ldcg(&ctrl->front0) <= front0_limit) return;
    const ChunkMap cm{0, 1, nchunks};
    int r = 0;
    for (;;) {
        collect_front(count, n, r, rank, perm, list, ctrl->cnt[r & 1]);
        grid.sync();
        build_prefix_t<kPer>(ctrl->cnt[r & 1], nchunks, n, tri != 0, cm, ps);
        if (ps.pc[nchunks] == 0) break;  // non_dominate.py:76-78: while any(front)
        if (blockIdx.x == 0)
            for (int c = threadIdx.x; c < nchunks; c += blockDim.x) ctrl->cnt[(r + 1) & 1][c] = 0;
        accumulate_front(D, rows, nchunks, tri != 0, cm, list, ps, count, -1, sm, tb);
        grid.sync();
        ++r;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) ctrl->fronts_done = r;
}

// ---- multi-GPU pieces (host drives the front loop; one small all-gather per front) -------------
// Ownership is by column chunk (the dominated index): a GPU keeps count[] for its own columns, so
// dominator counts never cross GPUs; what every GPU needs from the others is only WHO is in the
// current front -- one bit per individual.

// Marks the front among this GPU's own columns: bit j of `bits` (region-major, 128 words per local
// chunk) is set iff count[j] == 0; those counts drop to -1 (non_dominate.py:103).
__global__ void __launch_bounds__(256)
front_detect_owned_kernel(int *__restrict__ count, int n, ChunkMap cm, uint32_t *__restrict__ bits) {
    const int lane = threadIdx.x & 31;
    const int64_t words = (int64_t)cm.nlocal * kChunkWords;
    const int64_t gw = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;  // one warp per word
    if (gw >= words) return;
    const int lc = (int)(gw / kChunkWords);
    const int64_t j = (int64_t)cm.global(lc) * kChunkCols + (gw % kChunkWords) * 32 + lane;
    const bool in = (j < n) && (count[j] == 0);
    if (in) count[j] = -1;
    const uint32_t m = __ballot_sync(0xffffffffu, in);
    if (lane == 0) bits[gw] = m;
}

// Turns the gathered bitmask (P slices of `slice_chunks` local chunks each) into region-major member
// lists, writes the members' rank and counts the front.  One CTA of 128 threads per global chunk.
__global__ void __launch_bounds__(kChunkWords)
front_expand_kernel(const uint32_t *__restrict__ gathered, int nranks, int slice_chunks, int nchunks, int r,
                    int *__restrict__ rank, const int *__restrict__ perm, int *__restrict__ list,
                    int *__restrict__ cnt, int *__restrict__ total) {
    __shared__ int wsum[kChunkWords / 32];
    const int c = blockIdx.x;
    if (c >= nchunks) return;
    const int owner = c % nranks, lc = c / nranks;
    const uint32_t w = gathered[((int64_t)owner * slice_chunks + lc) * kChunkWords + threadIdx.x];
    const int mine = __popc(w);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int before = incl - mine;
    for (int q = 0; q < warp; ++q) before += wsum[q];
    uint32_t rest = w;
    int pos = before;
    while (rest) {
        const int b = __ffs(rest) - 1;
        rest &= rest - 1;
        const int j = c * kChunkCols + threadIdx.x * 32 + b;
        list[(int64_t)c * kChunkCols + pos++] = j;
        rank[perm ? perm[j] : j] = r;
    }
    if (threadIdx.x == kChunkWords - 1) {
        cnt[c] = before + mine;
        if (before + mine) atomicAdd(total, before + mine);
    }
}

// Subtracts the rows of the current front from this GPU's columns.
__global__ void __launch_bounds__(kPeelThreads)
front_subtract_kernel(const uint32_t *__restrict__ D, TriRows rows, int nchunks, int n, int tri, ChunkMap cm,
                      const int *__restrict__ list, const int *__restrict__ cnt, int *__restrict__ count, TileBands tb) {
    __shared__ PeelSmem sm;
    __shared__ PrefixSmem ps;
    build_prefix(cnt, nchunks, n, tri != 0, cm, ps);
    accumulate_front(D, rows, nchunks, tri != 0, cm, list, ps, count, -1, sm, tb);
}

// ---- multi-GPU, fused: the front loop and its exchange in ONE persistent kernel per GPU ---------------
// Instead of three launches and an ncclAllGather per front, every GPU runs one cooperative kernel for
// the whole peel.  Per front it marks the front among its own columns and CTA q stores that slice of
// the bitmask straight into rank q's exchange buffer over NVLink (peer pointers obtained through
// CUDA IPC at communicator creation).  Every 32-bit word of the bitmask travels as one 8-byte store
// {word, epoch}: 8-byte stores arrive whole, so the reader spins on the epoch tag of exactly the word
// it needs and no fence, flag or separate wait phase is required (the low-latency protocol NCCL uses
// for small messages).  The buffer is double-buffered by epoch parity: a sender can be at most one
// front ahead of the slowest reader, because it needs that reader's slice of the next front to go
// further -- every rank waits for a per-sender heartbeat of the epoch even when it needs none of that
// sender's words (a rank that owns no chunk would otherwise let the others run ahead and overwrite
// words it has not read yet).  Epochs keep counting across calls (stored in the header), so no barrier
// is needed between calls.  A wait that exceeds ~4 s raises `error` and ends the kernel instead of hanging the GPU.
constexpr int kXchgMaxRanks = 16;
constexpr int kXchgSliceWordsCap = kMaxChunks * kChunkWords;  // words per parity buffer (all ranks together)

struct XchgHeader {            // at the start of every rank's exchange buffer
    unsigned int epoch;        // last epoch this rank completed (persists across calls)
    unsigned int error;        // 1 = a wait timed out
    unsigned int pad0[62];
    unsigned int reserved[64];
    uint2 hb[2][16];           // hb[parity][q] = {0, epoch}: sender q has started delivering that epoch
};
static_assert(sizeof(XchgHeader) == 768, "header layout");
static_assert(kXchgMaxRanks <= 16, "hb[] holds 16 senders");

__host__ __device__ inline size_t xchg_buffer_bytes() {
    return sizeof(XchgHeader) + sizeof(uint2) * 2 * (size_t)kXchgSliceWordsCap;
}
__device__ __forceinline__ uint2 *xchg_slices(unsigned char *base, int parity) {
    return reinterpret_cast<uint2 *>(base + sizeof(XchgHeader)) + (size_t)parity * kXchgSliceWordsCap;
}

struct XchgCtx {
    unsigned char *const *peers;  // device array: peers[q] = rank q's exchange buffer (peers[me] is local)
    int me, nranks, slice_chunks;
};

// single 8-byte transactions (never torn): the {word, epoch} pairs of the exchange protocol
__device__ __forceinline__ uint2 ld_volatile_u2(const uint2 *p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_volatile_u2(uint2 *p, uint2 v) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

template <int kPer>
__global__ void __launch_bounds__(kPeelThreads, 2)
peel_p2p_kernel(const uint32_t *__restrict__ D, TriRows rows, int nchunks, int n, ChunkMap cm,
                int *__restrict__ count, int *__restrict__ rank, const int *__restrict__ perm, int *__restrict__ list,
                int *__restrict__ cnt, XchgCtx x, int *__restrict__ fronts_done,
                TileBands tb, unsigned int *host_error, unsigned long long timeout_ns) {
    __shared__ PeelSmem sm;
    __shared__ PrefixSmem ps;
    __shared__ int sh_wsum[kPeelThreads / 32];
    __shared__ int sh_fail;
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *self = x.peers[x.me];
    XchgHeader *hdr = reinterpret_cast<XchgHeader *>(self);
    unsigned int e = *reinterpret_cast<volatile unsigned int *>(&hdr->epoch);
    const int slice_words = x.slice_chunks * kChunkWords;
    int r = 0;
    bool failed = false;
    for (;;) {
        ++e;
        // 1+2. front among my own columns; the warp that finds a 32-column word sends it, tagged with the
        //      epoch, straight to every rank (my own buffer included) as one 8-byte store each -- no fence,
        //      no staging.  Padding words of my slice (ranks owning fewer chunks) are sent as zeros.
        {
            const int64_t words = (int64_t)slice_words;
            const int64_t own_words = (int64_t)cm.nlocal * kChunkWords;
            const int64_t nwarps = (int64_t)gridDim.x * (kPeelThreads / 32);
            for (int64_t gw = (int64_t)blockIdx.x * (kPeelThreads / 32) + warp; gw < words; gw += nwarps) {
                uint32_t m = 0u;
                if (gw < own_words) {
                    const int lc = (int)(gw / kChunkWords);
                    const int64_t j = (int64_t)cm.global(lc) * kChunkCols + (gw % kChunkWords) * 32 + lane;
                    const bool in = (j < n) && (__ldcg(count + j) == 0);
                    if (in) count[j] = -1;
                    m = __ballot_sync(0xffffffffu, in);
                }
                if (lane < x.nranks)
                    st_volatile_u2(xchg_slices(x.peers[lane], e & 1) + (size_t)x.me * slice_words + gw, make_uint2(m, e));
            }
            if (blockIdx.x == 0 && threadIdx.x < x.nranks)
                st_volatile_u2(&reinterpret_cast<XchgHeader *>(x.peers[threadIdx.x])->hb[e & 1][x.me], make_uint2(0u, e));
        }
        // 3+4. bitmask of all ranks -> region-major member lists, ranks, region sizes; a reader spins on
        //      the epoch tag of the word it needs
        if (threadIdx.x == 0) sh_fail = 0;
        __syncthreads();
        if (blockIdx.x == 0 && threadIdx.x < x.nranks) {  // heartbeats of all senders (see the header comment)
            const uint2 *src = &hdr->hb[e & 1][threadIdx.x];
            uint2 v = ld_volatile_u2(src);
            unsigned long long t0 = 0ull;
            unsigned int spins = 0u;
            while (v.y != e) {
                if ((++spins & 1023u) == 0u) {
                    const unsigned long long now = global_ns();
                    if (t0 == 0ull) t0 = now;
                    if (now - t0 > timeout_ns || *reinterpret_cast<volatile unsigned int *>(&hdr->error)) {
                        sh_fail = 1;
                        break;
                    }
                }
                v = ld_volatile_u2(src);
            }
        }
        {
            const uint2 *gathered = xchg_slices(self, e & 1);
            for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
                const int owner = c % x.nranks, lc = c / x.nranks;
                uint32_t w = 0u;
                if (threadIdx.x < kChunkWords) {
                    const uint2 *src = gathered + ((size_t)owner * x.slice_chunks + lc) * kChunkWords + threadIdx.x;
                    uint2 v = ld_volatile_u2(src);
                    unsigned long long t0 = 0ull;
                    unsigned int spins = 0u;
                    while (v.y != e) {
                        if ((++spins & 1023u) == 0u) {
                            const unsigned long long now = global_ns();
                            if (t0 == 0ull) t0 = now;
                            if (now - t0 > timeout_ns || *reinterpret_cast<volatile unsigned int *>(&hdr->error)) {
                                sh_fail = 1;
                                break;
                            }
                        }
                        v = ld_volatile_u2(src);
                    }
                    w = v.x;
                }
                const int mine = __popc(w);
                int incl = mine;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += t;
                }
                __syncthreads();
                if (lane == 31) sh_wsum[warp] = incl;
                __syncthreads();
                if (sh_fail) break;
                int before = incl - mine;
                for (int q = 0; q < warp; ++q) before += sh_wsum[q];
                int pos = before;
                while (w) {
                    const int b = __ffs(w) - 1;
                    w &= w - 1;
                    const int j = c * kChunkCols + threadIdx.x * 32 + b;
                    list[(int64_t)c * kChunkCols + pos++] = j;
                    rank[perm ? perm[j] : j] = r;
                }
                if (threadIdx.x == kPeelThreads - 1) cnt[c] = before + mine;
            }
        }
        if (sh_fail) {  // a peer never delivered: flag it (all CTAs see hdr->error and leave too)
            if (threadIdx.x == 0) {
                hdr->error = 1;
                if (host_error) {  // mapped host word: the library refuses further calls on this communicator
                    *host_error = 1u;
                    __threadfence_system();
                }
            }
            failed = true;
        }
        grid.sync();
        if (*reinterpret_cast<volatile unsigned int *>(&hdr->error)) {  // uniform across the grid after the sync
            failed = true;
            break;
        }
        // 5. subtract the front's rows from my columns
        build_prefix_t<kPer>(cnt, nchunks, n, true, cm, ps);
        if (ps.pc[nchunks] == 0) break;  // the same bitmask everywhere: every rank stops at the same front
        accumulate_front(D, rows, nchunks, true, cm, list, ps, count, -1, sm, tb);
        grid.sync();
        ++r;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (!failed) hdr->epoch = e;
        *fronts_done = failed ? -1 : r;
    }
}

// After a peer time-out (fronts_done < 0) the ranks written so far are meaningless: make that visible.
__global__ void __launch_bounds__(256)
poison_ranks_kernel(int *__restrict__ rank, int n, const int *__restrict__ fronts_done) {
    if (*fronts_done >= 0) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rank[i] = -1;
}

}  // namespace evoxb
