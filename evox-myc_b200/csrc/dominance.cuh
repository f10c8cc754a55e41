// K0/K1: objective transpose and the tiled pairwise-dominance build.
//
// Replaces src/evox/utils/common.py:95-104 (`_dominate`, `dominate_relation`) as called at
// src/evox/operators/selection/non_dominate.py:60 of the reference.  The reference materialises
// an (N, N) byte matrix; here row i of the matrix is a run of 32-bit words, bit (j & 31) of word
// (j >> 5) = "i dominates j".
#pragma once

#include "common.cuh"

namespace evoxb {

// ---- K0: (n, m) row-major -> (m, npad) objective-major, padded with NaN --------------------
// NaN padding makes every compare against a padding column false, so padding bits are 0.
__global__ void transpose_pad_kernel(const float *__restrict__ x, int64_t n, int m, int64_t npad,
                                     float *__restrict__ xt) {
    const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (j >= npad) return;
    const float nanv = __int_as_float(0x7fc00000);
    for (int k = 0; k < m; ++k) xt[(int64_t)k * npad + j] = (j < n) ? x[j * m + k] : nanv;
}

// ---- K1: dominance tile -----------------------------------------------------------------------
// CTA = 256 threads = 8 warps.  The CTA owns kTileRows rows x 1024 columns.  Warp w owns the 128
// columns [w*128, w*128+128): lane l keeps objective vectors of columns l, l+32, l+64, l+96 in
// registers (kC = 4 ballot words per row).  Rows are read from shared memory as warp-uniform
// broadcasts.  For row i and column word c the 32 lanes evaluate the predicate and
// __ballot_sync packs them; lane (i & 31) keeps the word, so after 32 rows every lane holds the
// four words of one row and stores them as a single 16-byte vector.
constexpr int kBuildThreads = 256;
constexpr int kTileRows = 128;
constexpr int kC = 4;
constexpr int kTileCols = (kBuildThreads / 32) * kC * 32;  // 1024

template <int M>
__device__ __forceinline__ bool dominates(const float (&a)[M], const float (&b)[M]) {
    // all(a <= b) & any(a < b) -- utils/common.py:98.  IEEE compares: NaN => false.
    bool lt = false;
#pragma unroll
    for (int k = 0; k < M; ++k) lt = lt || (a[k] < b[k]);
    bool p = lt;
#pragma unroll
    for (int k = 0; k < M; ++k) p = p && (a[k] <= b[k]);
    return p;
}

template <int M>
__global__ void __launch_bounds__(kBuildThreads)
dominance_build_kernel(const float *__restrict__ xt, int64_t npad, int64_t n_rows, int64_t row0,
                       uint32_t *__restrict__ D, int64_t row_words) {
    // xs[r][k]: objective vectors of the tile's rows
    __shared__ float xs[kTileRows][M];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int64_t col_base = (int64_t)blockIdx.x * kTileCols + warp * (kC * 32);
    const int64_t tile_row0 = (int64_t)blockIdx.y * kTileRows;  // local row index (0-based in this shard)

    for (int t = threadIdx.x; t < kTileRows * M; t += kBuildThreads) {
        const int r = t / M, k = t - r * M;
        const int64_t gi = row0 + tile_row0 + r;  // global row index into xt
        xs[r][k] = (tile_row0 + r < n_rows) ? xt[(int64_t)k * npad + gi] : __int_as_float(0x7fc00000);
    }
    float xj[kC][M];
#pragma unroll
    for (int c = 0; c < kC; ++c)
#pragma unroll
        for (int k = 0; k < M; ++k) xj[c][k] = xt[(int64_t)k * npad + col_base + c * 32 + lane];
    __syncthreads();

    for (int rb = 0; rb < kTileRows; rb += 32) {
        uint32_t w[kC];
#pragma unroll
        for (int c = 0; c < kC; ++c) w[c] = 0u;
#pragma unroll
        for (int ii = 0; ii < 32; ++ii) {
            float xi[M];
#pragma unroll
            for (int k = 0; k < M; ++k) xi[k] = xs[rb + ii][k];
            const bool mine = (lane == ii);
#pragma unroll
            for (int c = 0; c < kC; ++c) {
                const uint32_t b = __ballot_sync(0xffffffffu, dominates<M>(xi, xj[c]));
                if (mine) w[c] = b;
            }
        }
        const int64_t r = tile_row0 + rb + lane;
        if (r < n_rows) {
            uint4 v = make_uint4(w[0], w[1], w[2], w[3]);
            *reinterpret_cast<uint4 *>(D + r * row_words + (col_base >> 5)) = v;
        }
    }
}

// Generic-M fallback (M > 12): same tiling, objectives looped at run time from shared memory.
__global__ void __launch_bounds__(kBuildThreads)
dominance_build_generic_kernel(const float *__restrict__ xt, int64_t npad, int64_t n_rows, int64_t row0,
                               int m, uint32_t *__restrict__ D, int64_t row_words) {
    extern __shared__ float sm[];  // rows: kTileRows*m, then cols: kTileCols*m (objective-major)
    float *xs = sm;
    float *cs = sm + kTileRows * m;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int64_t tile_col0 = (int64_t)blockIdx.x * kTileCols;
    const int64_t tile_row0 = (int64_t)blockIdx.y * kTileRows;
    for (int t = threadIdx.x; t < kTileRows * m; t += kBuildThreads) {
        const int r = t / m, k = t - r * m;
        xs[t] = (tile_row0 + r < n_rows) ? xt[(int64_t)k * npad + row0 + tile_row0 + r] : __int_as_float(0x7fc00000);
    }
    for (int t = threadIdx.x; t < kTileCols * m; t += kBuildThreads) {
        const int k = t / kTileCols, j = t - k * kTileCols;
        cs[t] = xt[(int64_t)k * npad + tile_col0 + j];
    }
    __syncthreads();
    for (int rb = 0; rb < kTileRows; rb += 32) {
        uint32_t w[kC] = {0u, 0u, 0u, 0u};
        for (int ii = 0; ii < 32; ++ii) {
            const float *xi = xs + (rb + ii) * m;
#pragma unroll
            for (int c = 0; c < kC; ++c) {
                const int j = warp * (kC * 32) + c * 32 + lane;
                bool le = true, lt = false;
                for (int k = 0; k < m; ++k) {
                    const float a = xi[k], b = cs[k * kTileCols + j];
                    le = le && (a <= b);
                    lt = lt || (a < b);
                }
                const uint32_t b = __ballot_sync(0xffffffffu, le && lt);
                if (lane == ii) w[c] = b;
            }
        }
        const int64_t r = tile_row0 + rb + lane;
        if (r < n_rows)
            *reinterpret_cast<uint4 *>(D + r * row_words + ((tile_col0 + warp * (kC * 32)) >> 5)) =
                make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// Launches K1 for rows [row0, row0 + n_rows) of the population against all npad columns.
inline int launch_dominance_build(const float *xt, int64_t npad, int64_t row0, int64_t n_rows, int m,
                                  uint32_t *D, int64_t row_words, cudaStream_t s) {
    if (n_rows <= 0) return 0;
    dim3 grid((unsigned)(npad / kTileCols), (unsigned)div_up(n_rows, kTileRows));
    if (grid.y > 65535u) return fail(EVOXB_E_UNSUPPORTED, "too many row tiles (%u)", grid.y);
#define EVOXB_BUILD_CASE(MM)                                                                           \
    case MM:                                                                                           \
        dominance_build_kernel<MM><<<grid, kBuildThreads, 0, s>>>(xt, npad, n_rows, row0, D, row_words); \
        break;
    switch (m) {
        EVOXB_BUILD_CASE(1)
        EVOXB_BUILD_CASE(2)
        EVOXB_BUILD_CASE(3)
        EVOXB_BUILD_CASE(4)
        EVOXB_BUILD_CASE(5)
        EVOXB_BUILD_CASE(6)
        EVOXB_BUILD_CASE(7)
        EVOXB_BUILD_CASE(8)
        EVOXB_BUILD_CASE(9)
        EVOXB_BUILD_CASE(10)
        EVOXB_BUILD_CASE(11)
        EVOXB_BUILD_CASE(12)
        default: {
            const size_t smem = sizeof(float) * (size_t)(kTileRows + kTileCols) * (size_t)m;
            // the attribute is per device and the call is cheap: set it whenever more than the default is needed
            if (smem > 48 * 1024)
                EVOXB_CUDA(cudaFuncSetAttribute(dominance_build_generic_kernel,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            dominance_build_generic_kernel<<<grid, kBuildThreads, smem, s>>>(xt, npad, n_rows, row0, m, D, row_words);
        }
    }
#undef EVOXB_BUILD_CASE
    EVOXB_LAUNCH_CHECK("dominance_build_kernel");
    return 0;
}

// ---- K1 (presorted population): upper-triangular build ------------------------------------------
// Rows are in lexicographic order (presort.cuh), so D[i][j] = (uid_i < uid_j) & all_{k>=1}(x_ik <= x_jk):
// M-1 float compares + 1 integer compare per pair instead of 2 M, and only tiles that reach above
// the diagonal are computed.  Row tiles (128 rows) are grouped in bands of 32 (= one 4096-column
// chunk); band b needs column tiles 4 b .. ntc-1 (inside the diagonal chunk positions carry no order, so
// all of its tiles are computed).
struct TriGrid {
    int ntc;     // column tiles (npad / 1024)
    int nrt;     // row tiles of this launch
    int nbands;  // ceil(nrt / 32)
    long long total;
    int sharded;  // 0: 1-D grid over the needed tiles of the whole matrix; 1: 2-D grid (local column tile, row tile)
    ChunkMap cm;  // column chunks owned by this GPU (sharded only)
};

__host__ __device__ inline long long tri_tiles_before_band(int b, int ntc) {
    return 32ll * ((long long)b * ntc - 2ll * b * (b - 1));
}

inline TriGrid make_tri_grid(int64_t npad, int64_t n_rows) {
    TriGrid g;
    g.ntc = (int)(npad / 1024);
    g.nrt = (int)div_up(n_rows, 128);
    g.nbands = (g.nrt + 31) / 32;
    const int last = g.nbands - 1;
    g.total = tri_tiles_before_band(last, g.ntc) + (long long)(g.nrt - 32 * last) * (g.ntc - 4 * last);
    g.sharded = 0;
    g.cm = ChunkMap{0, 1, g.ntc / 4};
    return g;
}

// Tile id -> (row tile, column tile, word offset of the tile's first column inside a (local) matrix row).
// (bx, by) are the block indices of the one-tile-per-CTA launch: bx = tile id (one GPU) or local column tile
// (sharded), by = row tile (sharded).  Returns false for a sharded tile left of its rows' chunk (never read).
__device__ __forceinline__ bool tri_tile_coords(const TriGrid &g, long long bx, int by, int &rt, int &ct, int64_t &word0) {
    if (g.sharded) {
        const int lct = (int)bx, lc = lct >> 2, c = g.cm.global(lc);
        rt = by;
        if (c < (rt >> 5)) return false;
        ct = c * 4 + (lct & 3);
        word0 = (int64_t)lct * (kTileCols / 32);
    } else {
        const long long t = bx;
        int lo = 0, hi = g.nbands;  // before(lo) <= t < before(hi)
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (tri_tiles_before_band(mid, g.ntc) <= t) lo = mid; else hi = mid;
        }
        const int band = lo;
        const int width = g.ntc - 4 * band;
        const long long local = t - tri_tiles_before_band(band, g.ntc);
        rt = 32 * band + (int)(local / width);
        ct = 4 * band + (int)(local % width);
        word0 = (int64_t)ct * (kTileCols / 32);
    }
    return true;
}

// One 128 x 1024 tile of the presorted matrix (body: dominance_tri_body.cuh).  kPacked: triangular storage
// (TriRows::vstart); the rectangular instance keeps the plain row * stride addressing.
template <int M, bool kPacked>
__global__ void __launch_bounds__(kBuildThreads)
dominance_build_tri_kernel(const float *__restrict__ xt, int64_t npad, int64_t n_rows, const int *__restrict__ uid_row,
                           const int *__restrict__ uid_col, const float *__restrict__ tile_min1,
                           const float *__restrict__ tile_max1, uint32_t *__restrict__ D, TriRows rows, TriGrid g) {
#define EVOXB_BX blockIdx.x
#define EVOXB_BY blockIdx.y
#include "dominance_tri_body.cuh"
#undef EVOXB_BX
#undef EVOXB_BY
}

template <int M, bool kPacked>
__device__ __forceinline__ void build_tri_tile(const float *__restrict__ xt, int64_t npad, int64_t n_rows,
                                               const int *__restrict__ uid_row, const int *__restrict__ uid_col,
                                               const float *__restrict__ tile_min1, const float *__restrict__ tile_max1,
                                               uint32_t *__restrict__ D, const TriRows &rows, const TriGrid &g, long long bx,
                                               int by) {
#define EVOXB_BX bx
#define EVOXB_BY by
#include "dominance_tri_body.cuh"
#undef EVOXB_BX
#undef EVOXB_BY
}

// cm == nullptr: the whole matrix on one GPU.  Otherwise only the column chunks of *cm are built and a
// matrix row holds cm->nlocal * 128 words.
inline int launch_dominance_build_tri(const float *xt, int64_t npad, int64_t n_rows, int m, const int *uid_row,
                                      const int *uid_col, const float *tile_min1, const float *tile_max1, uint32_t *D,
                                      TriRows rows, cudaStream_t s, const ChunkMap *cm = nullptr) {
    if (n_rows <= 0) return 0;
    TriGrid g = make_tri_grid(npad, n_rows);
    if (g.total > 0x7fffffffll) return fail(EVOXB_E_UNSUPPORTED, "too many tiles (%lld)", g.total);
    dim3 grid((unsigned)g.total);
    if (cm) {
        g.sharded = 1;
        g.cm = *cm;
        if (cm->nlocal <= 0) return 0;
        if (g.nrt > 65535) return fail(EVOXB_E_UNSUPPORTED, "too many row tiles (%d)", g.nrt);
        grid = dim3((unsigned)(cm->nlocal * 4), (unsigned)g.nrt);
    }
#define EVOXB_TRI_CASE(MM)                                                                                         \
    case MM:                                                                                                       \
        if (rows.tri)                                                                                              \
            dominance_build_tri_kernel<MM, true><<<grid, kBuildThreads, 0, s>>>(xt, npad, n_rows, uid_row, uid_col, \
                                                                                tile_min1, tile_max1, D, rows, g);  \
        else                                                                                                       \
            dominance_build_tri_kernel<MM, false><<<grid, kBuildThreads, 0, s>>>(xt, npad, n_rows, uid_row, uid_col, \
                                                                                 tile_min1, tile_max1, D, rows, g); \
        break;
    switch (m) {
        EVOXB_TRI_CASE(1) EVOXB_TRI_CASE(2) EVOXB_TRI_CASE(3) EVOXB_TRI_CASE(4) EVOXB_TRI_CASE(5) EVOXB_TRI_CASE(6)
        EVOXB_TRI_CASE(7) EVOXB_TRI_CASE(8) EVOXB_TRI_CASE(9) EVOXB_TRI_CASE(10) EVOXB_TRI_CASE(11) EVOXB_TRI_CASE(12)
        EVOXB_TRI_CASE(13) EVOXB_TRI_CASE(14) EVOXB_TRI_CASE(15) EVOXB_TRI_CASE(16)
        default:
            return fail(EVOXB_E_UNSUPPORTED, "presorted build supports m <= 16 (got %d)", m);
    }
#undef EVOXB_TRI_CASE
    EVOXB_LAUNCH_CHECK("dominance_build_tri_kernel");
    return 0;
}

// The same tiles from a small persistent grid, run only when *go is set: the exact float kernel as the fall-back of
// the level build below (both are enqueued; the flag is decided on the device).
template <int M, bool kPacked>
__global__ void __launch_bounds__(kBuildThreads)
dominance_build_tri_fallback_kernel(const float *__restrict__ xt, int64_t npad, int64_t n_rows,
                                    const int *__restrict__ uid_row, const int *__restrict__ uid_col,
                                    const float *__restrict__ tile_min1, const float *__restrict__ tile_max1,
                                    uint32_t *__restrict__ D, TriRows rows, TriGrid g, long long gx, int gy,
                                    const int *__restrict__ go) {
    if (*go == 0) return;
    const long long total = gx * gy;
    for (long long t = blockIdx.x; t < total; t += gridDim.x) {
        build_tri_tile<M, kPacked>(xt, npad, n_rows, uid_row, uid_col, tile_min1, tile_max1, D, rows, g, t % gx, (int)(t / gx));
        __syncthreads();  // the tile's shared arrays are reused by the next one
    }
}

inline int launch_dominance_build_tri_fallback(const float *xt, int64_t npad, int64_t n_rows, int m, const int *uid_row,
                                               const int *uid_col, const float *tile_min1, const float *tile_max1,
                                               uint32_t *D, TriRows rows, TriGrid g, long long gx, int gy, const int *go,
                                               cudaStream_t s) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned grid = (unsigned)(sms * 5);  // five CTAs of 256 threads per SM, as the one-tile-per-CTA launch reaches
#define EVOXB_TRI_CASE(MM)                                                                                          \
    case MM:                                                                                                        \
        if (rows.tri)                                                                                               \
            dominance_build_tri_fallback_kernel<MM, true><<<grid, kBuildThreads, 0, s>>>(                           \
                xt, npad, n_rows, uid_row, uid_col, tile_min1, tile_max1, D, rows, g, gx, gy, go);                  \
        else                                                                                                        \
            dominance_build_tri_fallback_kernel<MM, false><<<grid, kBuildThreads, 0, s>>>(                          \
                xt, npad, n_rows, uid_row, uid_col, tile_min1, tile_max1, D, rows, g, gx, gy, go);                  \
        break;
    switch (m) {
        EVOXB_TRI_CASE(2) EVOXB_TRI_CASE(3) EVOXB_TRI_CASE(4) EVOXB_TRI_CASE(5) EVOXB_TRI_CASE(6)
        EVOXB_TRI_CASE(7) EVOXB_TRI_CASE(8) EVOXB_TRI_CASE(9) EVOXB_TRI_CASE(10) EVOXB_TRI_CASE(11) EVOXB_TRI_CASE(12)
        EVOXB_TRI_CASE(13) EVOXB_TRI_CASE(14) EVOXB_TRI_CASE(15) EVOXB_TRI_CASE(16)
        default:
            return fail(EVOXB_E_UNSUPPORTED, "presorted build supports 2 <= m <= 16 here (got %d)", m);
    }
#undef EVOXB_TRI_CASE
    EVOXB_LAUNCH_CHECK("dominance_build_tri_fallback_kernel");
    return 0;
}

// ---- K1, level form ------------------------------------------------------------------------------
// Objectives 1..M-1 are replaced by 15-bit LEVELS (levels.cuh): a monotone map of the float order, so
// x_ik <= x_jk implies level_ik <= level_jk.  Two levels share a 32-bit word, bit 15 of each half is a guard bit
// that the column side sets: (column word - row word) keeps a guard bit iff that half's row level <= the column's
// (a borrow out of a failing half can only make a higher half fail too, and the product is an AND anyway).  Per 32
// pairs the inner loop is W = ceil((M-1)/2) IMAD (FMA pipe) + the ANDs + one LOP3 with predicate output + VOTE
// instead of M-1 FSETP + VOTE on the ALU pipe alone: twice the pairs per clock at M = 5 (measured).  The result is
// a superset of the dominance relation; the pairs that share a level in some objective but compare the other way
// as floats are cleared afterwards by level_fixup_kernel.  Rows / columns with a NaN are rejected by the uid test
// exactly as in the float kernel.
constexpr uint32_t kLvlGuard = 0x80008000u;

// bad = some guard bit of (t & dl) is missing; returns the ballot of the pairs that pass and adds this lane's
// pair to cnt (the lane's column gains a dominator)
__device__ __forceinline__ uint32_t lvl_ballot(uint32_t t, uint32_t dl, uint32_t &cnt) {
    uint32_t b, dummy;
    asm("{ .reg .pred p, q;\n"
        "setp.eq.u32 q, 0, 0;\n"
        "lop3.and.b32 %1|p, %3, %4, 0x80008000, 0x2a, q;\n"
        "@!p add.u32 %2, %2, 1;\n"
        "vote.sync.ballot.b32 %0, !p, 0xffffffff; }"
        : "=r"(b), "=r"(dummy), "+r"(cnt)
        : "r"(t), "r"(dl));
    return b;
}
__device__ __forceinline__ uint32_t lvl_ballot_uid(uint32_t t, uint32_t dl, int ui, int uj, uint32_t &cnt) {
    uint32_t b, dummy;
    asm("{ .reg .pred p, q, r;\n"
        "setp.eq.u32 q, 0, 0;\n"
        "lop3.and.b32 %1|p, %3, %4, 0x80008000, 0x2a, q;\n"
        "setp.lt.and.s32 r, %5, %6, !p;\n"
        "@r add.u32 %2, %2, 1;\n"
        "vote.sync.ballot.b32 %0, r, 0xffffffff; }"
        : "=r"(b), "=r"(dummy), "+r"(cnt)
        : "r"(t), "r"(dl), "r"(ui), "r"(uj));
    return b;
}

template <int W>
struct LvlCfg {
    // row words padded to a power of two so that a row is one vector load from shared memory
    static constexpr int kPad = W <= 1 ? 1 : W <= 2 ? 2 : W <= 4 ? 4 : 8;
    // rows per CTA: as many as keep rows + columns + staging under 48 KB of shared memory
    static constexpr int kRows = W <= 2 ? 1024 : W <= 4 ? 512 : W <= 6 ? 256 : 128;
    static constexpr int kRowTiles = kRows / kTileRows;  // 128-row tiles per CTA
    static constexpr int kBlocks = kRows / 32;
};

__host__ __device__ inline long long lvl_groups_before_band(int b, int ntc, int row_tiles) {
    return tri_tiles_before_band(b, ntc) / row_tiles;  // a band holds 32 row tiles: divisible by 8, 4, 2, 1
}

// CTAs of the one-GPU launch
inline long long lvl_total_groups(const TriGrid &g, int row_tiles) {
    const int last = g.nbands - 1;
    return lvl_groups_before_band(last, g.ntc, row_tiles) +
           (long long)((g.nrt - 32 * last + row_tiles - 1) / row_tiles) * (g.ntc - 4 * last);
}

// A CTA takes kRows consecutive rows (inside one 1024-tile, hence one band of x_1) against one 1024-column tile.
// Rows and columns are staged in shared memory once; the work is then cut into items of 32 rows x 128 columns
// (one x_2 band of rows against one x_2 band of columns) that the eight warps draw from a counter in shared
// memory.  No barrier after the prologue: the items differ a lot in cost (dead, one word dropped, full, with uid
// test) and a static split by columns left the warps of the low-x_2 columns idle (ncu: 27 % of the issue slots
// lost, the largest stall reason the barrier).
template <int W, bool kPacked>
__global__ void __launch_bounds__(kBuildThreads)
dominance_build_lvl_kernel(const uint32_t *__restrict__ q, int64_t npad, int64_t n_rows, const int *__restrict__ uid_row,
                           const int *__restrict__ uid_col, const float *__restrict__ tile_min1,
                           const float *__restrict__ tile_max1, uint32_t *__restrict__ D, TriRows rows, TriGrid g,
                           int *__restrict__ count, const int *__restrict__ fallback) {
    using Cfg = LvlCfg<W>;
    constexpr int WP = Cfg::kPad, R = Cfg::kRows, NB = Cfg::kBlocks;
    __shared__ __align__(16) uint32_t qs[R][WP];          // row level words
    __shared__ int us[R];                                 // row uids
    __shared__ uint32_t qc[W][kTileCols];                 // column level words, guard bits set
    __shared__ int uc[kTileCols];                         // column uids
    __shared__ __align__(16) uint32_t stage[kBuildThreads / 32][32][kC];
    __shared__ int rb_min2[NB], rb_max2[NB], rb_umax[NB];  // per 32-row block: band of objective 2's level, largest uid
    __shared__ int oc_min2[8], oc_max2[8], oc_umin[8];     // per 128 columns
    constexpr bool kSmemCnt = W <= 7;                      // W = 8 has no shared memory left: counts go straight to global
    __shared__ int ccnt[kSmemCnt ? kTileCols : 1];         // dominators found by this CTA, per column
    __shared__ int next_item;
    if (*fallback) return;  // the levels are too coarse for this population: the float kernel runs instead
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;

    int rt0, ct;
    int64_t word0;  // word offset of the tile's first column inside a (local) matrix row
    if (g.sharded) {
        const int lct = blockIdx.x, lc = lct >> 2, c = g.cm.global(lc);
        rt0 = blockIdx.y * Cfg::kRowTiles;
        if (c < (rt0 >> 5)) return;  // chunk entirely left of the rows' own chunk: never read by the peel
        ct = c * 4 + (lct & 3);
        word0 = (int64_t)lct * (kTileCols / 32);
    } else {
        const long long t = blockIdx.x;
        int lo = 0, hi = g.nbands;  // before(lo) <= t < before(hi)
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (lvl_groups_before_band(mid, g.ntc, Cfg::kRowTiles) <= t) lo = mid; else hi = mid;
        }
        const int band = lo;
        const int width = g.ntc - 4 * band;
        const long long local = t - lvl_groups_before_band(band, g.ntc, Cfg::kRowTiles);
        rt0 = 32 * band + Cfg::kRowTiles * (int)(local / width);
        ct = 4 * band + (int)(local % width);
        word0 = (int64_t)ct * (kTileCols / 32);
    }
    const int64_t row0 = (int64_t)rt0 * kTileRows;
    const int64_t col0 = (int64_t)ct * kTileCols;
    // tile-level test on objective 1 (floats, as in the float kernel): the CTA's rows lie in one 1024-tile
    const int rt1024 = (int)(row0 / kTileCols);
    const float r_lo = tile_min1[rt1024], r_hi = tile_max1[rt1024];
    const float c_lo = tile_min1[ct], c_hi = tile_max1[ct];
    if (r_lo > c_hi) return;  // never fetched by the readers (see dominance_tri_body.cuh): nothing to write
    const bool sure1 = r_hi <= c_lo;
    const int nb = (int)min((int64_t)NB, (n_rows - row0 + 31) / 32);  // 32-row blocks that hold a row

    for (int t = threadIdx.x; t < R * W; t += kBuildThreads) {
        const int w = t / R, r = t - w * R;
        qs[r][w] = (row0 + r < n_rows) ? q[(int64_t)w * npad + row0 + r] : 0u;
    }
    for (int r = threadIdx.x; r < R; r += kBuildThreads) us[r] = (row0 + r < n_rows) ? uid_row[row0 + r] : 0x7fffffff;
    for (int t = threadIdx.x; t < kTileCols * W; t += kBuildThreads) {
        const int w = t / kTileCols, j = t - w * kTileCols;
        qc[w][j] = q[(int64_t)w * npad + col0 + j] | kLvlGuard;
    }
    for (int j = threadIdx.x; j < kTileCols; j += kBuildThreads) {
        uc[j] = uid_col[col0 + j];
        if (kSmemCnt) ccnt[j] = 0;
    }
    if (threadIdx.x == 0) next_item = 0;
    __syncthreads();
    // bands: warp w reduces row blocks w, w + 8, ... and the columns of octile w.  The band of objective 2 is the
    // high half of word 0 (only meaningful when M >= 3; a padding half is 0 on both sides).
    for (int b = warp; b < NB; b += kBuildThreads / 32) {
        int lo = (int)(qs[b * 32 + lane][0] >> 16), hi = lo, um = us[b * 32 + lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            um = max(um, __shfl_xor_sync(0xffffffffu, um, o));
        }
        if (lane == 0) {
            rb_min2[b] = lo;
            rb_max2[b] = hi;
            rb_umax[b] = um;
        }
    }
    {
        int lo = 0x7fff, hi = 0, um = 0x7fffffff;
#pragma unroll
        for (int c = 0; c < kC; ++c) {
            const int v = (int)(qc[0][warp * 128 + c * 32 + lane] >> 16) & 0x7fff;
            lo = min(lo, v);
            hi = max(hi, v);
            um = min(um, uc[warp * 128 + c * 32 + lane]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            um = min(um, __shfl_xor_sync(0xffffffffu, um, o));
        }
        if (lane == 0) {
            oc_min2[warp] = lo;
            oc_max2[warp] = hi;
            oc_umin[warp] = um;
        }
    }
    __syncthreads();
    uint32_t(*st)[kC] = stage[warp];

    // One 32-row block against 128 columns.  FIRST = 1: word 0 (objectives 1 and 2) is known to pass for every
    // pair of the block and is left out.
#define EVOXB_LVL_BLOCK(WITH_UID, FIRST)                                                                   \
    {                                                                                                       \
        _Pragma("unroll 8") for (int ii = 0; ii < 32; ++ii) {                                               \
            uint32_t a[WP];                                                                                 \
            if (WP == 1) a[0] = qs[rb + ii][0];                                                             \
            else if (WP == 2) { const uint2 v2 = *reinterpret_cast<const uint2 *>(qs[rb + ii]); a[0] = v2.x; a[1] = v2.y; } \
            else {                                                                                          \
                _Pragma("unroll") for (int h = 0; h < WP / 4; ++h) {                                        \
                    const uint4 v4 = *reinterpret_cast<const uint4 *>(&qs[rb + ii][4 * h]);                 \
                    a[4 * h] = v4.x; a[4 * h + 1] = v4.y; a[4 * h + 2] = v4.z; a[4 * h + 3] = v4.w;         \
                }                                                                                           \
            }                                                                                               \
            const int ui = WITH_UID ? us[rb + ii] : 0;                                                      \
            uint32_t b[kC];                                                                                 \
            _Pragma("unroll") for (int c = 0; c < kC; ++c) {                                                \
                constexpr int F = (FIRST) && W > 1 ? 1 : 0;                                                 \
                uint32_t t = xj[c][F] - a[F];                                                               \
                _Pragma("unroll") for (int w = F + 1; w < W - 1; ++w) t &= xj[c][w] - a[w];                 \
                const uint32_t dl = (W - 1 > F) ? xj[c][W - 1] - a[W - 1] : t;                              \
                b[c] = WITH_UID ? lvl_ballot_uid(t, dl, ui, uj[c], cn[c]) : lvl_ballot(t, dl, cn[c]);       \
            }                                                                                               \
            if (lane == 0) *reinterpret_cast<uint4 *>(st[ii]) = make_uint4(b[0], b[1], b[2], b[3]);         \
        }                                                                                                   \
        __syncwarp();                                                                                       \
        v = *reinterpret_cast<const uint4 *>(st[lane]);                                                     \
        __syncwarp();                                                                                       \
    }

    // An item is one 32-row block against 256 columns (two octiles, done one after the other): a lane then owns a
    // whole 32-byte sector of its row, so no sector is written in two halves at different times (which costs a
    // read-modify-write in DRAM once the first half has left the L2).  Column pair 3 first: the high-x_2 columns
    // have the most live row blocks.
    uint32_t xj[kC][W];
    int uj[kC];
    const int total = 4 * nb;
    for (;;) {
        int item = 0;
        if (lane == 0) item = atomicAdd(&next_item, 1);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= total) break;
        const int pair = 3 - item / nb, blk = item - (3 - pair) * nb;
        const int rb = blk * 32;
        uint4 out0 = make_uint4(0u, 0u, 0u, 0u), out1 = out0;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            const int oct = pair * 2 + h;
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            const int c_min2 = oc_min2[oct], c_max2 = oc_max2[oct];
            const bool dead = rb_min2[blk] > c_max2;    // no pair of the block passes objective 2
            const bool sure2 = rb_max2[blk] <= c_min2;  // every pair passes it
            if (!dead) {
                uint32_t cn[kC];
#pragma unroll
                for (int c = 0; c < kC; ++c) {
                    cn[c] = 0u;
                    uj[c] = uc[oct * 128 + c * 32 + lane];
#pragma unroll
                    for (int w = 0; w < W; ++w) xj[c][w] = qc[w][oct * 128 + c * 32 + lane];
                }
                if (!(rb_umax[blk] < oc_umin[oct])) {
                    EVOXB_LVL_BLOCK(true, 0)
                } else if (W > 1 && sure1 && sure2) {
                    EVOXB_LVL_BLOCK(false, 1)
                } else {
                    EVOXB_LVL_BLOCK(false, 0)
                }
                // K2 for free: the lane has counted the rows of the block that dominate each of its 4 columns
#pragma unroll
                for (int c = 0; c < kC; ++c)
                    if (cn[c]) {
                        if (kSmemCnt) atomicAdd(&ccnt[oct * 128 + c * 32 + lane], (int)cn[c]);
                        else atomicAdd(count + col0 + oct * 128 + c * 32 + lane, (int)cn[c]);
                    }
            }
            if (h == 0) out0 = v; else out1 = v;
        }
        const int64_t r = row0 + rb + lane;
        const int64_t wb = word0 + pair * 2 * kC;
        if (r < n_rows) {
            uint4 *dst = reinterpret_cast<uint4 *>((kPacked ? D + rows.vstart(r) * 4 + wb : D + r * rows.row_words + wb));
            dst[0] = out0;
            dst[1] = out1;
        }
    }
#undef EVOXB_LVL_BLOCK
    if (kSmemCnt) {
        __syncthreads();
        for (int j = threadIdx.x; j < kTileCols; j += kBuildThreads)
            if (ccnt[j]) atomicAdd(count + col0 + j, ccnt[j]);
    }
}

// ---- a2: dense boolean relation between two different sets (utils/common.py:101-104) ----------
// One thread per output byte, x row kept in registers via shared memory; small helper, not on the
// N^2 hot path (the hot path never materialises bytes).
__global__ void dominate_relation_dense_kernel(const float *__restrict__ x, int64_t nx, const float *__restrict__ y,
                                               int64_t ny, int m, uint8_t *__restrict__ out) {
    const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t i = blockIdx.y;
    if (j >= ny || i >= nx) return;
    bool le = true, lt = false;
    for (int k = 0; k < m; ++k) {
        const float a = x[i * m + k], b = y[j * m + k];
        le = le && (a <= b);
        lt = lt || (a < b);
    }
    out[i * ny + j] = (le && lt) ? 1 : 0;
}

}  // namespace evoxb
