// (a fragment, not a header: no include guard on purpose)
// Body of one 128 x 1024 tile of the presorted float build, included twice by dominance.cuh: in
// dominance_build_tri_kernel (one tile per CTA, EVOXB_BX / EVOXB_BY = blockIdx) and in the tile function of the
// persistent fall-back kernel.  Expects the kernel's parameters (xt, npad, n_rows, uid_row, uid_col, tile_min1,
// tile_max1, D, rows, g) and the template parameters M, kPacked in scope; leaves a tile early with `return`.
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
        const int lct = EVOXB_BX, lc = lct >> 2, c = g.cm.global(lc);
        rt = EVOXB_BY;
        if (c < (rt >> 5)) return;  // chunk entirely left of the rows' own chunk: never read by the peel
        ct = c * 4 + (lct & 3);
        word0 = (int64_t)lct * (kTileCols / 32);
    } else {
        // tile id -> (band, row tile, column tile)
        const long long t = EVOXB_BX;
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
        // Nothing in the tile can dominate, and nothing is written: every reader of the matrix (accumulate_chunk in
        // peel.cuh, the staging loops of peel_flow.cuh) evaluates this same test on the same two arrays and does not
        // fetch the tile's 128-byte row quarters.
        if (r_lo > c_hi) return;
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
