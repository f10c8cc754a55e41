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
// chunk); band b needs column tiles 4 b .. ntc-1.  Tiles of the diagonal chunk that lie entirely
// left of their rows are written as zeros because the peel streams whole 512-byte chunks.
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

// kPacked: triangular storage (TriRows::vstart); the rectangular instance keeps the plain row * stride addressing
template <int M, bool kPacked>
__global__ void __launch_bounds__(kBuildThreads)
dominance_build_tri_kernel(const float *__restrict__ xt, int64_t npad, int64_t n_rows, const int *__restrict__ uid_row,
                           const int *__restrict__ uid_col, const float *__restrict__ tile_min1,
                           const float *__restrict__ tile_max1, uint32_t *__restrict__ D, TriRows rows, TriGrid g) {
    constexpr int K = (M > 1) ? (M - 1) : 1;  // objectives 1..M-1 are compared
    __shared__ float xs[kTileRows][K];
    __shared__ int us[kTileRows];
    __shared__ __align__(16) uint32_t stage[kBuildThreads / 32][32][kC];  // per warp: 32 rows x 4 ballot words
    __shared__ float rb_min1[kTileRows / 32], rb_max1[kTileRows / 32];     // band of objective 1 per 32-row block
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;

    int rt, ct;
    int64_t word0;  // word offset of the tile's first column inside a (local) matrix row
    if (g.sharded) {
        const int lct = blockIdx.x, lc = lct >> 2, c = g.cm.global(lc);
        rt = blockIdx.y;
        if (c < (rt >> 5)) return;  // chunk entirely left of the rows' own chunk: never read by the peel
        ct = c * 4 + (lct & 3);
        word0 = (int64_t)lct * (kTileCols / 32);
    } else {
        // tile id -> (band, row tile, column tile)
        const long long t = blockIdx.x;
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

    const int64_t tile_row0 = (int64_t)rt * kTileRows;
    const int64_t tile_col0 = (int64_t)ct * kTileCols;
    const int64_t col_base = tile_col0 + warp * (kC * 32);
    const int64_t word_base = word0 + warp * kC;
    // Tile-level test on objective 1 (rows inside a chunk are ordered by it, so a 1024-tile is a band of
    // x_1; presort.cuh): if every row's x_1 exceeds every column's, nothing in the tile can dominate.
    // (A column tile of the rows' own chunk may lie "left" of the rows: inside a chunk positions carry no
    // order, so such tiles are computed like any other.)
    bool sure1 = false;  // every row's x_1 <= every column's x_1: that compare is dropped for the tile
    if (M > 1) {
        const int rt1024 = (int)(tile_row0 / kTileCols);
        const float r_lo = tile_min1[rt1024], r_hi = tile_max1[rt1024];
        const float c_lo = tile_min1[ct], c_hi = tile_max1[ct];
        if (r_lo > c_hi) {
            for (int rb = 0; rb < kTileRows; rb += 32) {
                const int64_t r = tile_row0 + rb + lane;
                if (r < n_rows) *reinterpret_cast<uint4 *>((kPacked ? D + rows.vstart(r) * 4 + word_base : D + r * rows.row_words + word_base)) = make_uint4(0u, 0u, 0u, 0u);
            }
            return;
        }
        sure1 = r_hi <= c_lo;
    }

    for (int q = threadIdx.x; q < kTileRows * K; q += kBuildThreads) {
        const int r = q / K, k = q - r * K;
        xs[r][k] = (M > 1 && tile_row0 + r < n_rows) ? xt[(int64_t)(k + 1) * npad + tile_row0 + r]
                                                      : __int_as_float(0x7fc00000);
    }
    for (int r = threadIdx.x; r < kTileRows; r += kBuildThreads)
        us[r] = (tile_row0 + r < n_rows) ? uid_row[tile_row0 + r] : 0x7fffffff;
    float xj[kC][K];
    int uj[kC];
#pragma unroll
    for (int c = 0; c < kC; ++c) {
        uj[c] = uid_col[col_base + c * 32 + lane];
#pragma unroll
        for (int k = 0; k < K; ++k)
            xj[c][k] = (M > 1) ? xt[(int64_t)(k + 1) * npad + col_base + c * 32 + lane] : 0.0f;
    }
    __syncthreads();

    // If every row uid of the tile is below every column uid of this warp, the uid test is true for
    // all of the warp's pairs and is dropped from the inner loop (the common case away from the
    // diagonal; NaN rows / columns carry INT_MAX / 0 and force the full predicate).
    int row_umax = max(max(us[lane], us[lane + 32]), max(us[lane + 64], us[lane + 96]));
    int col_umin = min(min(uj[0], uj[1]), min(uj[2], uj[3]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        row_umax = max(row_umax, __shfl_xor_sync(0xffffffffu, row_umax, o));
        col_umin = min(col_umin, __shfl_xor_sync(0xffffffffu, col_umin, o));
    }
    const bool need_uid = !(row_umax < col_umin);
    uint32_t(*st)[kC] = stage[warp];

    // Bands of objective 2 (rows inside a 1024-tile are ordered by it, presort.cuh): smallest / largest
    // value of every 32-row block and of this warp's 128 columns.  fminf / fmaxf skip NaN; rows and
    // columns holding a NaN are rejected by the uid test, which is always active when one is present.
    constexpr int kB = (K > 1) ? 1 : 0;  // index of objective 2 in xs / xj
    float c_min1 = 0.0f, c_max1 = 0.0f;
    if (K > 1) {
        c_min1 = fminf(fminf(xj[0][kB], xj[1][kB]), fminf(xj[2][kB], xj[3][kB]));
        c_max1 = fmaxf(fmaxf(xj[0][kB], xj[1][kB]), fmaxf(xj[2][kB], xj[3][kB]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            c_min1 = fminf(c_min1, __shfl_xor_sync(0xffffffffu, c_min1, o));
            c_max1 = fmaxf(c_max1, __shfl_xor_sync(0xffffffffu, c_max1, o));
        }
        if (warp < kTileRows / 32) {
            float lo = xs[warp * 32 + lane][kB], hi = lo;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            }
            if (lane == 0) {
                rb_min1[warp] = lo;
                rb_max1[warp] = hi;
            }
        }
        __syncthreads();
    }

    // One 32-row block against this warp's 128 columns.  WITH_UID: keep the uid compare.  DROP: bit k set =
    // the compare of objective k+1 is known to hold for every pair of the block and is dropped.
#define EVOXB_TRI_BLOCK(WITH_UID, DROP)                                                                    \
    {                                                                                                       \
        _Pragma("unroll 8") for (int ii = 0; ii < 32; ++ii) {                                               \
            float xi[K];                                                                                    \
            _Pragma("unroll") for (int k = 0; k < K; ++k) xi[k] = xs[rb + ii][k];                           \
            const int ui = WITH_UID ? us[rb + ii] : 0;                                                      \
            uint32_t b[kC];                                                                                 \
            _Pragma("unroll") for (int c = 0; c < kC; ++c) {                                                \
                bool p = WITH_UID ? (ui < uj[c]) : true;                                                    \
                if (M > 1) {                                                                                \
                    _Pragma("unroll") for (int k = 0; k < K; ++k)                                           \
                        if (!((DROP) >> k & 1)) p = p && (xi[k] <= xj[c][k]);                               \
                }                                                                                           \
                b[c] = __ballot_sync(0xffffffffu, p);                                                       \
            }                                                                                               \
            /* the four words of row ii are warp-uniform: one lane parks them in shared memory */          \
            if (lane == 0) *reinterpret_cast<uint4 *>(st[ii]) = make_uint4(b[0], b[1], b[2], b[3]);         \
        }                                                                                                   \
        __syncwarp();                                                                                       \
        v = *reinterpret_cast<const uint4 *>(st[lane]);                                                     \
        __syncwarp();                                                                                       \
    }

    for (int rb = 0; rb < kTileRows; rb += 32) {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        // no row of the block can satisfy x_i2 <= x_j2 for any column of the warp: the words stay zero
        const bool dead = (K > 1) && (rb_min1[rb >> 5] > c_max1);
        // every row of the block satisfies it for every column: the compare is dropped
        const bool sure2 = (K > 1) && (rb_max1[rb >> 5] <= c_min1);
        if (!dead) {
            // Dropping a compare needs its own copy of the unrolled loop; copies are only worth their
            // instruction-cache footprint where they save a sizeable share of the work: not near the
            // diagonal (uid test active, rare) and not for many objectives.
            if (need_uid) {
                EVOXB_TRI_BLOCK(true, 0)
            } else if (K > 5) {
                EVOXB_TRI_BLOCK(false, 0)
            } else {
                if (sure1) { if (sure2) EVOXB_TRI_BLOCK(false, 3) else EVOXB_TRI_BLOCK(false, 1) }
                else       { if (sure2) EVOXB_TRI_BLOCK(false, 2) else EVOXB_TRI_BLOCK(false, 0) }
            }
        }
        const int64_t r = tile_row0 + rb + lane;
        if (r < n_rows) *reinterpret_cast<uint4 *>((kPacked ? D + rows.vstart(r) * 4 + word_base : D + r * rows.row_words + word_base)) = v;
    }
#undef EVOXB_TRI_BLOCK
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
