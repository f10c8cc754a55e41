// K3 (data-flow form): the front-peeling loop without grid-wide barriers.
//
// Same integers as the while_loop of src/evox/operators/selection/non_dominate.py:72-111 (and the
// NumPy loop at :11-27) of the reference; different schedule.  peel_persistent_kernel (peel.cuh)
// spreads every front evenly over the grid and pays two grid-wide syncs, a prefix and an atomic flush
// per front (about 16 us); for populations whose fronts are small that fixed cost IS the peel time.
// Here every CTA OWNS a fixed set of matrix columns:
//   * the dominator counts of its columns live in its shared memory (one int per column, -1 once the column has
//     left), so subtracting a front needs no global atomics, and the warp ballot of "count reached zero" IS the
//     column word of the next front;
//   * a front travels between CTAs as tagged words: every 32-column word of the front bitmask is one
//     8-byte {word, epoch} store (arrives whole) into a double buffer, and a word is valid when its tag is the
//     epoch being read.  No fence and no barrier.  Every CTA reads the complete bitmask of a front before it
//     publishes its share of the next one, which bounds run-ahead to one front and makes two buffers (epoch
//     parity) enough; every CTA sees the same bitmask, so all of them stop at the same (empty) front.  A progress
//     flag per CTA (polled by one warp) only tells WHEN reading the words is worthwhile;
//   * the same tagged stores and flags go to the peer GPUs' buffers over NVLink in the multi-GPU use of the
//     kernel (peers[] holds one buffer per rank; a single GPU has one entry in its own workspace).
// Per front and CTA: flags -> the front's words -> staged member rows -> the piece's bytes of those rows (through
// a shared-memory tile, or streamed straight into per-thread bit planes when the front is large) -> Harley-Seal
// adds -> one thread per column subtracts and detects -> tagged stores of the next front.
//
// Column ownership: a matrix row is cut in units of 256 columns (32 bytes).  A chunk (4096 columns)
// is split in 1..16 pieces of equal width, narrower where more rows have bits (chunk c of the
// presorted, triangular matrix only takes rows of chunks <= c, weight c + 1); pieces are dealt to the
// CTAs round-robin in order of decreasing chunk, one piece per CTA whenever there are enough columns.
#pragma once

#include "peel.cuh"

namespace evoxb {

constexpr int kFlowLgThreads = 9;
constexpr int kFlowThreads = 1 << kFlowLgThreads;
constexpr int kFlowWarps = kFlowThreads / 32;
constexpr int kUnitWords = 8;                                // 256 columns
constexpr int kUnitsPerChunk = kChunkWords / kUnitWords;     // 16
constexpr int kFlowMaxPieces = 32;                           // per CTA
constexpr int kFlowStageCap = 16384;                         // members staged per round
constexpr int kFlowMaxChunks = 128;                          // N <= 524288 (front words are kept in shared memory)
constexpr int kFlowMaxRanks = 16;
constexpr int kFlowMaxCtas = 192;                            // flag slots per rank (one CTA per SM)
constexpr int kFlowTileWords = 8192;                         // shared-memory tile of the small path / gather scratch of the large one
constexpr int kFlowStreamPlanes = 9;                         // streaming mode: a stream adds at most 2^9 - 8 rows
constexpr int kFlowAuxWords = kFlowStreamPlanes * 4 * kFlowThreads;  // per-thread row planes parked for the merge

#ifndef EVOXB_FLOW_TIMING
#define EVOXB_FLOW_TIMING 0
#endif

// ---- column ownership plan (host and device agree by construction) ------------------------------
__host__ __device__ inline int flow_unit_weight(int chunk, int nchunks, bool tri) { return tri ? chunk + 1 : nchunks; }

__host__ __device__ inline long long flow_total_weight(int nchunks, ChunkMap cm, bool tri) {
    long long w = 0;
    for (int lc = 0; lc < cm.nlocal; ++lc) w += (long long)kUnitsPerChunk * flow_unit_weight(cm.global(lc), nchunks, tri);
    return w;
}

// units per piece of a chunk whose units weigh `w` each: the widest power of two that keeps a piece at
// or below the target weight (at least one unit)
__host__ __device__ inline int flow_piece_units(int w, long long target) {
    int nu = kUnitsPerChunk;
    while (nu > 1 && (long long)nu * w > target) nu >>= 1;
    return nu;
}

struct FlowPlan {
    int grid;       // CTAs to launch; every one of them owns at least one piece
    int npieces;
    int max_words;  // most 32-column words owned by one CTA
    int max_pieces;
    long long target;
};

inline int flow_count_pieces(int nchunks, ChunkMap cm, bool tri, long long target) {
    int np = 0;
    for (int lc = 0; lc < cm.nlocal; ++lc) np += kUnitsPerChunk / flow_piece_units(flow_unit_weight(cm.global(lc), nchunks, tri), target);
    return np;
}

// A pass over a piece costs about the same whether it streams ten rows or a thousand (it is latency),
// so a CTA should own ONE piece whenever there are enough columns: the target weight grows until the
// pieces fit the grid (pieces weigh between target / 2 and target, except single units above it).
inline FlowPlan flow_make_plan(int nchunks, ChunkMap cm, bool tri, int max_ctas) {
    FlowPlan p{};
    const long long total = flow_total_weight(nchunks, cm, tri);
    p.target = total / max_ctas > 1 ? total / max_ctas : 1;
    for (int it = 0; it < 64 && flow_count_pieces(nchunks, cm, tri, p.target) > max_ctas; ++it) {
        const long long grown = p.target + (p.target + 7) / 8;
        if ((long long)kUnitsPerChunk * flow_unit_weight(cm.global(cm.nlocal - 1), nchunks, tri) <= p.target) break;  // whole chunks already
        p.target = grown;
    }
    p.npieces = flow_count_pieces(nchunks, cm, tri, p.target);
    p.grid = p.npieces < max_ctas ? p.npieces : max_ctas;
    if (p.grid < 1) return p;
    std::vector<int> words((size_t)p.grid, 0), cnt((size_t)p.grid, 0);
    int idx = 0;
    for (int lc = cm.nlocal - 1; lc >= 0; --lc) {  // pieces in order of decreasing chunk, dealt round-robin
        const int nu = flow_piece_units(flow_unit_weight(cm.global(lc), nchunks, tri), p.target);
        for (int i = 0; i < kUnitsPerChunk / nu; ++i, ++idx) {
            words[(size_t)(idx % p.grid)] += nu * kUnitWords;
            cnt[(size_t)(idx % p.grid)] += 1;
        }
    }
    for (int c = 0; c < p.grid; ++c) {
        if (words[(size_t)c] > p.max_words) p.max_words = words[(size_t)c];
        if (cnt[(size_t)c] > p.max_pieces) p.max_pieces = cnt[(size_t)c];
    }
    return p;
}

struct FlowPiece {
    int lc;         // local chunk (index into this GPU's stored columns)
    int chunk;      // global chunk
    int unit0;      // first unit inside the chunk
    int nunits;     // power of two, 1..16
    int word_base;  // offset of the piece in the CTA's count array (in 32-column words)
};

// ---- the exchange: tagged front words + one progress flag per CTA ------------------------------------
struct FlowHeader {         // at the start of every exchange buffer
    unsigned int epoch;     // tag of the last front published by the previous call (persists across calls)
    unsigned int error;     // 1 = a wait timed out (sticky: the communicator is unusable afterwards)
    unsigned int pad[62];
};
static_assert(sizeof(FlowHeader) == 256, "header layout");

constexpr size_t kFlowFlagBytes = sizeof(unsigned int) * kFlowMaxRanks * kFlowMaxCtas;

__host__ __device__ inline size_t flow_ring_bytes(int64_t slot_words) {
    return sizeof(FlowHeader) + kFlowFlagBytes + 2 * sizeof(uint2) * (size_t)slot_words;
}

struct FlowPeers {
    unsigned char *base[kFlowMaxRanks];  // base[q] = rank q's exchange buffer (base[me] is local memory)
};

struct FlowArgs {
    const uint32_t *D;
    TriRows rows;         // where packed rows start (rectangular or triangular storage)
    int nchunks, n, tri;
    ChunkMap cm;
    const int *count;     // dominator counts of this GPU's columns, indexed by (sorted) position
    int *rank_sorted;     // out: rank by (sorted) position, every position of every GPU
    TileBands tb;
    FlowPeers peers;
    int me, nranks;
    int grids[kFlowMaxRanks];  // CTAs of every rank's kernel (they all publish, so they all are waited for)
    int64_t slot_words;   // words per parity buffer (>= npad / 32)
    int max_words;        // FlowPlan::max_words (shared-memory layout)
    long long target;     // FlowPlan::target
    int *fronts_done;     // out (device): number of fronts, -1 after a time-out
    unsigned int *host_error;  // optional mapped host word raised together with FlowHeader::error
    unsigned long long timeout_ns;
    const int *front0;    // optional (device): size of front 0; the kernel leaves at once when it exceeds front0_limit
    int front0_limit;     //   (a few huge fronts are bandwidth work: the grid-wide kernel of peel.cuh does those)
    long long *timing;    // development builds (EVOXB_FLOW_TIMING): clocks per phase of CTA 0, else nullptr
};

__device__ __forceinline__ unsigned int *flow_flags(unsigned char *base) {
    return reinterpret_cast<unsigned int *>(base + sizeof(FlowHeader));
}
__device__ __forceinline__ uint2 *flow_slot(unsigned char *base, int64_t slot_words, unsigned int e) {
    return reinterpret_cast<uint2 *>(base + sizeof(FlowHeader) + kFlowFlagBytes) + (size_t)(e & 1u) * (size_t)slot_words;
}
__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
// polling loads: strong at GPU scope (served by this GPU's L2, where local and peer stores to this memory land)
__device__ __forceinline__ unsigned int ld_relaxed_u32(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint2 ld_relaxed_u2(const uint2 *p) {
    uint2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(unsigned int *p, unsigned int v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// shared-memory layout (dynamic): see flow_smem_bytes()
struct FlowSmemLayout {
    int *rows;         // staged member rows [kFlowStageCap]
    uint32_t *tile;    // [kFlowTileWords]: row tile of the small path; cross-warp scratch of the large path
    uint32_t *aux;     // [kFlowAuxWords]: per-thread partial planes of the small path
    int *counts;       // [max_words * 32] dominators left per owned column; -1 once the column has left (or is padding)
    float *tmin;       // [ntiles] smallest x_1 of every 1024-tile (copy of TileBands::tmin)
    float *tmax;       // [ntiles] largest (copy of TileBands::tmax; +inf without band information)
};

__host__ __device__ inline size_t flow_smem_bytes(int nchunks, int max_words) {
    size_t b = (size_t)kFlowStageCap * 4;               // rows
    b += (size_t)kFlowTileWords * 4;                    // tile
    b += (size_t)kFlowAuxWords * 4;                     // aux
    b += (size_t)max_words * 32 * 4;                    // counts
    b += (size_t)nchunks * 4 * 4 * 2;                   // tmin, tmax
    return b;
}

// adds eight one-bit-per-column words into the low PL planes of a bit-sliced counter (Harley-Seal step)
template <int PL, int N>
__device__ __forceinline__ void flow_add8(uint32_t (&pl)[N], const uint32_t (&w)[kBatch]) {
    static_assert(PL <= N && PL >= 4, "plane count");
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
    for (int p = 3; p < PL; ++p) {
        const uint32_t t = pl[p] & carry;
        pl[p] ^= carry;
        carry = t;
    }
}

// eight one-bit inputs -> 4-plane count (0..8)
__device__ __forceinline__ void flow_count8(const uint32_t (&in)[8], uint32_t (&val)[4]) {
    uint32_t c0, s0, c1, s1, c2, s2, c3, d0, t0;
    csa(c0, s0, in[0], in[1], in[2]);
    csa(c1, s1, in[3], in[4], in[5]);
    s2 = in[6] ^ in[7];
    c2 = in[6] & in[7];
    csa(c3, val[0], s0, s1, s2);
    csa(d0, t0, c0, c1, c2);
    val[1] = t0 ^ c3;
    const uint32_t d1 = t0 & c3;
    val[2] = d0 ^ d1;
    val[3] = d0 & d1;
}

// acc += val << p   (bit-sliced; val has NV planes, acc NA planes; p is a compile-time constant after unrolling)
template <int NV, int NA>
__device__ __forceinline__ void flow_add_shifted(uint32_t (&acc)[NA], const uint32_t (&val)[NV], int p) {
    uint32_t carry = 0u;
#pragma unroll
    for (int q = 0; q < NA; ++q) {
        if (q >= p) {
            const uint32_t y = (q - p < NV) ? val[(q - p < NV) ? q - p : 0] : 0u;
            const uint32_t x = acc[q];
            acc[q] = lop3_xor3(x, y, carry);
            carry = lop3_maj(x, y, carry);
        }
    }
}

struct FlowShared {
    FlowPiece pieces[kFlowMaxPieces];
    int npieces;
    int owned_words;
    int fail;
    int staged;                    // members of the current front that can touch this CTA's columns
    int kk[kFlowMaxChunks];        // pieces per chunk, in order of decreasing local chunk
    int pref[kFlowMaxChunks + 1];  // prefix of kk
    int warp_any[kFlowWarps];
};

#if EVOXB_FLOW_TIMING
#define FLOW_T(i)                                  \
    if (t == 0) {                                  \
        const long long now__ = clock64();         \
        tacc[i] += now__ - tlast;                  \
        tlast = now__;                             \
    }
#else
#define FLOW_T(i)
#endif

// ---- one piece against the staged members ------------------------------------------------------------
// 1. the piece's bytes of the member rows go to a shared-memory tile (four 16-byte loads in flight per thread);
// 2. thread (word w, group g) adds rows g, g + ng, ... of the tile into bit planes (Harley-Seal, registers);
// 3. with more than 8 groups, thread (word, plane) first counts that plane's bits over the groups;
// 4. one thread per COLUMN sums its bit of every (plane, group) word -- independent terms, no carry chain --
//    subtracts the sum from its count, and the warp's ballot of "count reached zero" is the column word of
//    the next front, published on the spot (tagged, to every rank).
// Latency matters here, not work: a front is a few dozen rows per CTA, so every serial chain was kept short.
constexpr int kFlowRowPlanes = 12;   // a thread adds at most 8192 / 4 rows
constexpr int kFlowGroupPlanes = 9;  // at most 256 groups / streams

// nm == 0: nothing to subtract, only detect and publish.  `publish` is false while more rounds of the same
// front are still to be subtracted (fronts above kFlowStageCap members).
__device__ __forceinline__ void flow_piece_pass(const FlowArgs &a, const FlowPiece &pc, const FlowSmemLayout &sm, int nm, bool publish,
                                                unsigned int e_next, long long (&tacc)[8], long long &tlast) {
    const int t = threadIdx.x, lane = t & 31;
    const int nparts = pc.nunits * 2;  // 16-byte parts of the piece (2..32)
    const int lg_parts = 31 - __clz(nparts);
    const int W = nparts * 4;          // words of the piece
    const int lg_w = lg_parts + 2;
    int np = 0, ng = 1, lg_ng = 0, kg = 0;
    bool merged = false;
    if (nm > 0) {
        // groups: about 32 rows per thread, at least 8 groups, at most 512 / W
        const int max_ng = kFlowThreads >> lg_w;
        ng = 8;
        while (ng < max_ng && ng * 32 < nm) ng <<= 1;
        if (ng > max_ng) ng = max_ng;
        lg_ng = 31 - __clz(ng);
        const int rows_per_tile = kFlowTileWords >> lg_w;
        const bool tri = a.tri != 0;
        const int part = t & (nparts - 1);
        const uint4 *base = reinterpret_cast<const uint4 *>(a.D + (int64_t)pc.lc * kChunkWords + pc.unit0 * kUnitWords) + part;
        const float col_max = sm.tmax[pc.chunk * 4 + ((pc.unit0 * kUnitWords + part * 4) >> 5)];
        uint4 *tile4 = reinterpret_cast<uint4 *>(sm.tile);
        const int w = t & (W - 1), g = t >> lg_w;

        // Two ways to add the rows.  A front of more than one tile is bandwidth work: every thread then streams
        // rows straight from global memory, eight 16-byte loads in flight, into its own bit planes (no tile, no
        // barrier per tile) and the 512 / parts streams are merged like the groups below.  Smaller fronts go
        // through the shared-memory tile, where threads are (word, group) and the merge is short.
        const int nstreams = kFlowThreads >> lg_parts;
        const bool streaming = nm > rows_per_tile && nm <= ((1 << kFlowStreamPlanes) - 2 * kBatch) * nstreams;
        if (streaming) {
            const int strm = t >> lg_parts;
            ng = nstreams;
            lg_ng = kFlowLgThreads - lg_parts;
            uint32_t p4[4][kFlowStreamPlanes];
#pragma unroll
            for (int s4 = 0; s4 < 4; ++s4)
#pragma unroll
                for (int p = 0; p < kFlowStreamPlanes; ++p) p4[s4][p] = 0u;
#pragma unroll 1
            for (int q0 = strm * kBatch; q0 < nm; q0 += nstreams * kBatch) {
                uint4 vv[kBatch];
#pragma unroll
                for (int b = 0; b < kBatch; ++b) {
                    const int q = q0 + b;
                    const int row = sm.rows[min(q, nm - 1)];
                    const bool fetch = q < nm && (!tri || (row >> 12) <= pc.chunk) && !(sm.tmin[row >> 10] > col_max);
                    vv[b] = ldg_stream_v4_if(base + a.rows.vstart(row), fetch);
                }
                uint32_t in[kBatch];
#pragma unroll
                for (int b = 0; b < kBatch; ++b) in[b] = vv[b].x;
                flow_add8<kFlowStreamPlanes>(p4[0], in);
#pragma unroll
                for (int b = 0; b < kBatch; ++b) in[b] = vv[b].y;
                flow_add8<kFlowStreamPlanes>(p4[1], in);
#pragma unroll
                for (int b = 0; b < kBatch; ++b) in[b] = vv[b].z;
                flow_add8<kFlowStreamPlanes>(p4[2], in);
#pragma unroll
                for (int b = 0; b < kBatch; ++b) in[b] = vv[b].w;
                flow_add8<kFlowStreamPlanes>(p4[3], in);
            }
            // a stream takes whole batches of kBatch rows: at most kBatch * ceil(nm / (kBatch * streams)) of them
            np = 32 - __clz(kBatch * ((nm + kBatch * ng - 1) / (kBatch * ng)));
#pragma unroll
            for (int p = 0; p < kFlowStreamPlanes; ++p)
                if (p < np) {
#pragma unroll
                    for (int s4 = 0; s4 < 4; ++s4) sm.aux[(((p << lg_ng) + strm) << lg_w) + part * 4 + s4] = p4[s4][p];
                }
            __syncthreads();
            FLOW_T(5)
        } else {
        uint32_t pl[kFlowRowPlanes];
#pragma unroll
        for (int p = 0; p < kFlowRowPlanes; ++p) pl[p] = 0u;
        // Tiles of kFlowTileWords: every thread owns up to U 16-byte slots of a tile (tile == U * 512 slots).  The
        // loads of tile k + 1 are issued before tile k is added, so DRAM latency overlaps the adds.
        constexpr int U = kFlowTileWords / 4 / kFlowThreads;  // 4
        uint4 v[U];
        auto issue = [&](int qb) {
            const int items = min(rows_per_tile, nm - qb) << lg_parts;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = t + u * kFlowThreads;
                const int row = sm.rows[qb + (min(i, items - 1) >> lg_parts)];
                // rows of later chunks have nothing (and nothing defined) in this chunk's columns; a row tile whose
                // smallest x_1 exceeds the column tile's largest is all zero and was never written by the build kernel
                const bool fetch = i < items && (!tri || (row >> 12) <= pc.chunk) && !(sm.tmin[row >> 10] > col_max);
                v[u] = ldg_stream_v4_if(base + a.rows.vstart(row), fetch);
            }
        };
        issue(0);
#pragma unroll 1
        for (int qb = 0; qb < nm; qb += rows_per_tile) {
            const int nt = min(rows_per_tile, nm - qb);
            const int items = nt << lg_parts;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = t + u * kFlowThreads;
                if (i < items) tile4[i] = v[u];
            }
            __syncthreads();
            if (qb + rows_per_tile < nm) issue(qb + rows_per_tile);
            FLOW_T(5)
            if (g < ng) {
#pragma unroll 1
                for (int q0 = g; q0 < nt; q0 += ng * kBatch) {
                    uint32_t in[kBatch];
#pragma unroll
                    for (int b = 0; b < kBatch; ++b) {
                        const int q = q0 + (b << lg_ng);
                        in[b] = (q < nt) ? sm.tile[(q << lg_w) + w] : 0u;
                    }
                    flow_add8<kFlowRowPlanes>(pl, in);
                }
            }
            __syncthreads();
        }
        // planes that can be non-zero: a thread added at most ceil(nm / ng) rows
        np = 32 - __clz((nm + ng - 1) >> lg_ng);
        if (g < ng) {
#pragma unroll
            for (int p = 0; p < kFlowRowPlanes; ++p)
                if (p < np) sm.aux[(((p << lg_ng) + g) << lg_w) + w] = pl[p];  // [plane][group][word]
        }
        __syncthreads();
        }
        FLOW_T(6)
        if (ng > 8) {  // thread (word, plane): count of that plane's bits over the groups, kg planes, parked in the tile
            merged = true;
            kg = lg_ng + 1;
#pragma unroll 1
            for (int idx = t; idx < (np << lg_w); idx += kFlowThreads) {
                const int w2 = idx & (W - 1), p = idx >> lg_w;
                uint32_t cnt[kFlowGroupPlanes];
#pragma unroll
                for (int k = 0; k < kFlowGroupPlanes; ++k) cnt[k] = 0u;
#pragma unroll 1
                for (int g0 = 0; g0 < ng; g0 += 8) {
                    uint32_t in[8], val[4];
#pragma unroll
                    for (int b = 0; b < 8; ++b) in[b] = sm.aux[(((p << lg_ng) + g0 + b) << lg_w) + w2];
                    flow_count8(in, val);
                    flow_add_shifted<4, kFlowGroupPlanes>(cnt, val, 0);
                }
#pragma unroll
                for (int k = 0; k < kFlowGroupPlanes; ++k)
                    if (k < kg) sm.tile[((p * kg + k) << lg_w) + w2] = cnt[k];
            }
            __syncthreads();
        }
    }
    // one thread per column of the piece (a warp = one 32-column word)
#pragma unroll 1
    for (int col = t; col < (W << 5); col += kFlowThreads) {
        const int w = col >> 5;
        int dec = 0;
        if (nm > 0) {
            const uint32_t lanebit = 1u << lane;
            if (merged) {
                const uint32_t *src = sm.tile + w;  // [plane][group plane][word]
#pragma unroll 1
                for (int p = 0; p < np; ++p) {
                    int s = 0;
#pragma unroll 1
                    for (int k = 0; k < kg; ++k, src += W) s += ((*src & lanebit) ? 1 : 0) << k;
                    dec += s << p;
                }
            } else {
                const uint32_t *src = sm.aux + w;   // [plane][group][word]
#pragma unroll 1
                for (int p = 0; p < np; ++p) {
                    int s = 0;
#pragma unroll 4
                    for (int g8 = 0; g8 < ng; ++g8, src += W) s += (*src & lanebit) ? 1 : 0;
                    dec += s << p;
                }
            }
        }
        int *cp = sm.counts + ((pc.word_base + w) << 5) + lane;
        int c = *cp - dec;
        const bool zero = publish && c == 0;
        if (zero) c = -1;  // leaves for good (non_dominate.py:103)
        if (dec != 0 || zero) *cp = c;
        if (publish) {
            const uint32_t z = __ballot_sync(0xffffffffu, zero);
            if (lane < a.nranks) {
                const size_t gw = (size_t)pc.chunk * kChunkWords + (size_t)(pc.unit0 * kUnitWords + w);
                st_volatile_u2(flow_slot(a.peers.base[lane], a.slot_words, e_next) + gw, make_uint2(z, e_next));
            }
        }
    }
    __syncthreads();
    FLOW_T(7)
}

// slow path of reading a front word: the tag is not there yet.  Returns the word, or sets *failed.
__device__ __noinline__ uint32_t flow_wait_word(const uint2 *src, unsigned int e, const FlowHeader *hdr,
                                                unsigned long long timeout_ns, int *failed) {
    unsigned long long t0 = 0ull;
    unsigned int spins = 0u;
    for (;;) {
        const uint2 v = ld_relaxed_u2(src);
        if (v.y == e) return v.x;
        __nanosleep(40);
        if ((++spins & 255u) == 0u) {
            const unsigned long long now = global_ns();
            if (t0 == 0ull) t0 = now;
            if (now - t0 > timeout_ns || *reinterpret_cast<const volatile unsigned int *>(&hdr->error)) {
                *failed = 1;
                return 0u;
            }
        }
    }
}

__device__ __noinline__ void flow_write_ranks(int *rank_sorted, int w, uint32_t word, int r) {
    while (word) {
        const int b = __ffs(word) - 1;
        word &= word - 1;
        rank_sorted[(size_t)w * 32 + b] = r;
    }
}


__global__ void __launch_bounds__(kFlowThreads, 1)
peel_flow_kernel(FlowArgs a) {
    extern __shared__ __align__(16) unsigned char flow_smem_raw[];
    __shared__ FlowShared sh;
    if (a.front0 && __ldcg(a.front0) > a.front0_limit) return;  // uniform over the grid
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const bool tri = a.tri != 0;
    const int nw = a.nchunks * kChunkWords;  // 32-column words of the whole population
    FlowSmemLayout sm;
    {
        unsigned char *p = flow_smem_raw;
        sm.rows = reinterpret_cast<int *>(p);         p += (size_t)kFlowStageCap * 4;
        sm.tile = reinterpret_cast<uint32_t *>(p);    p += (size_t)kFlowTileWords * 4;
        sm.aux = reinterpret_cast<uint32_t *>(p);     p += (size_t)kFlowAuxWords * 4;
        sm.counts = reinterpret_cast<int *>(p);       p += (size_t)a.max_words * 32 * 4;
        sm.tmin = reinterpret_cast<float *>(p);       p += (size_t)a.nchunks * 4 * 4;
        sm.tmax = reinterpret_cast<float *>(p);
    }
    unsigned char *self = a.peers.base[a.me];
    FlowHeader *hdr = reinterpret_cast<FlowHeader *>(self);
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // only touched by development builds (EVOXB_FLOW_TIMING)
    long long tlast = EVOXB_FLOW_TIMING ? clock64() : 0;

    // ---- my pieces (same arithmetic as flow_make_plan) ----
    const int nlocal = a.cm.nlocal;
    if (t < nlocal) {
        const int lc = nlocal - 1 - t;
        sh.kk[t] = kUnitsPerChunk / flow_piece_units(flow_unit_weight(a.cm.global(lc), a.nchunks, tri), a.target);
    }
    for (int i = t; i < a.nchunks * 4; i += kFlowThreads) {
        sm.tmin[i] = a.tb.tmin ? a.tb.tmin[i] : 0.0f;
        sm.tmax[i] = a.tb.tmax ? a.tb.tmax[i] : __int_as_float(0x7f800000);
    }
    if (t == 0) sh.fail = 0;
    __syncthreads();
    if (t <= nlocal) {
        int s = 0;
        for (int o = 0; o < t; ++o) s += sh.kk[o];
        sh.pref[t] = s;
    }
    __syncthreads();
    const int npieces_all = sh.pref[nlocal];
    if (t < kFlowMaxPieces) {
        const int i = blockIdx.x + t * gridDim.x;
        if (i < npieces_all) {
            int lo = 0, hi = nlocal;  // pref[lo] <= i < pref[hi]
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (sh.pref[mid] <= i) lo = mid; else hi = mid;
            }
            FlowPiece pc;
            pc.lc = nlocal - 1 - lo;
            pc.chunk = a.cm.global(pc.lc);
            pc.nunits = kUnitsPerChunk / sh.kk[lo];
            pc.unit0 = (i - sh.pref[lo]) * pc.nunits;
            pc.word_base = 0;
            sh.pieces[t] = pc;
        }
    }
    __syncthreads();
    if (t == 0) {
        int np = 0, wb = 0;
        while (np < kFlowMaxPieces && blockIdx.x + np * (int)gridDim.x < npieces_all) {
            sh.pieces[np].word_base = wb;
            wb += sh.pieces[np].nunits * kUnitWords;
            ++np;
        }
        sh.npieces = np;
        sh.owned_words = wb;
    }
    __syncthreads();
    const int npieces = sh.npieces;
    int cmax = 0;  // members of regions beyond my last chunk never touch my columns
    for (int k = 0; k < npieces; ++k) cmax = max(cmax, sh.pieces[k].chunk);
    const int need_words = tri ? (cmax + 1) * kChunkWords : nw;
    const int rank_per = (nw + (int)gridDim.x - 1) / (int)gridDim.x;
    const int rank_w0 = (int)blockIdx.x * rank_per, rank_w1 = min(nw, rank_w0 + rank_per);

    // ---- counts of my columns (padding columns start at -1: they never join a front) ----
    for (int k = 0; k < npieces; ++k) {
        const FlowPiece pc = sh.pieces[k];
        const int pcols = pc.nunits * kUnitWords * 32;
        for (int c = t; c < pcols; c += kFlowThreads) {
            const int64_t j = (int64_t)pc.chunk * kChunkCols + (int64_t)pc.unit0 * kUnitWords * 32 + c;
            sm.counts[(pc.word_base << 5) + c] = (j < a.n) ? __ldcg(a.count + j) : -1;
        }
    }
    __syncthreads();
    FLOW_T(0)

    unsigned int e = *reinterpret_cast<volatile unsigned int *>(&hdr->epoch);
    int r = 0;
    bool failed = false;
    int front_total = 0, before = 0;  // members of the front being subtracted / this thread's first member number
    for (;;) {
        // ---- subtract the previous front (none before front 0), detect and publish front e ----
        ++e;
        int lo = 0;
        do {
            const int hi = min(front_total, lo + kFlowStageCap);
            if (front_total > kFlowStageCap) {
                // a front above the stage capacity: rounds of kFlowStageCap members; member number `pos` (order fixed
                // by the atomics of the compaction) goes to round pos / cap; only the last round detects and publishes
                const uint2 *prev = flow_slot(self, a.slot_words, e - 1);  // the front being subtracted
                int pos = before;
                for (int w = t; w < need_words; w += kFlowThreads) {
                    uint32_t rest = ld_relaxed_u2(prev + w).x;
                    while (rest) {
                        const int b = __ffs(rest) - 1;
                        rest &= rest - 1;
                        if (pos >= lo && pos < hi) sm.rows[pos - lo] = w * 32 + b;
                        ++pos;
                    }
                }
                __syncthreads();
            }
#pragma unroll 1
            for (int k = 0; k < npieces; ++k) flow_piece_pass(a, sh.pieces[k], sm, hi - lo, hi == front_total, e, tacc, tlast);
            lo = hi;
        } while (lo < front_total);
        // my progress flag, behind the words of front e (no fence: the flag only tells the readers when looking
        // at the words is worthwhile; the tag of every word is what makes it valid)
        if (t < a.nranks) st_volatile_u32(flow_flags(a.peers.base[t]) + a.me * kFlowMaxCtas + blockIdx.x, e);
        FLOW_T(1)
        // ---- a. wait until every CTA of every rank has published this front: warp 0 polls the (few lines of)
        //         flags, the other warps sleep at the barrier -- not every thread hammering the word it needs ----
        if (warp == 0) {
            unsigned long long t0 = 0ull;
            unsigned int spins = 0u;
            const unsigned int *flags = flow_flags(self);
            if (lane == 0) sh.staged = 0;
            for (;;) {
                int late = 0;  // loads are independent (no short-circuit): one L2 round trip per poll
                for (int q = 0; q < a.nranks; ++q) {
#pragma unroll 4
                    for (int i = lane; i < a.grids[q]; i += 32) late |= (int)(ld_relaxed_u32(flags + q * kFlowMaxCtas + i) - e);
                }
                if (__all_sync(0xffffffffu, late >= 0)) break;
                if ((++spins & 127u) == 0u) {
                    const unsigned long long now = global_ns();
                    if (t0 == 0ull) t0 = now;
                    if (now - t0 > a.timeout_ns || *reinterpret_cast<const volatile unsigned int *>(&hdr->error)) {
                        sh.fail = 1;
                        break;
                    }
                }
            }
        }
        __syncthreads();
        FLOW_T(2)
        // ---- b. the whole front, 4 words per thread at a time: tag check, ranks, and the members that can touch
        //         my columns go straight to the staged row list (any order: one shared-memory atomic per warp) ----
        uint32_t any = 0u;
        const uint2 *slot = flow_slot(self, a.slot_words, e);
        if (!sh.fail) {
            constexpr int U = 4;
#pragma unroll 1
            for (int w0 = t; w0 < nw; w0 += kFlowThreads * U) {
                uint2 v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int w = w0 + u * kFlowThreads;
                    v[u] = (w < nw) ? ld_relaxed_u2(slot + w) : make_uint2(0u, e);
                }
                uint32_t word[U];
                int mine = 0;
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int w = w0 + u * kFlowThreads;
                    word[u] = (v[u].y == e) ? v[u].x : flow_wait_word(slot + w, e, hdr, a.timeout_ns, &sh.fail);
                    any |= word[u];
                    // ranks: the words of the bitmask are dealt to the CTAs of this GPU in contiguous runs
                    if (word[u] && w >= rank_w0 && w < rank_w1) flow_write_ranks(a.rank_sorted, w, word[u], r);
                    if (w >= need_words) word[u] = 0u;
                    mine += __popc(word[u]);
                }
                if (w0 - t < need_words) {  // warp-uniform: this batch may hold members for me
                    int incl = mine;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int x = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= o) incl += x;
                    }
                    int wbase = 0;
                    if (lane == 31 && incl > 0) wbase = atomicAdd(&sh.staged, incl);
                    wbase = __shfl_sync(0xffffffffu, wbase, 31);
                    int pos = wbase + incl - mine;
                    if (pos + mine <= kFlowStageCap) {
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            uint32_t rest = word[u];
                            const int wcol = (w0 + u * kFlowThreads) << 5;
                            while (rest) {
                                const int bbit = __ffs(rest) - 1;
                                rest &= rest - 1;
                                sm.rows[pos++] = wcol + bbit;
                            }
                        }
                    }
                }
            }
        }
        const int some = __syncthreads_or(any != 0u);
        FLOW_T(3)
        if (sh.fail) {
            failed = true;
            break;
        }
        if (!some) break;  // non_dominate.py:76-78: while any(front)
        front_total = sh.staged;
        if (front_total > kFlowStageCap) {
            // rare: a front above the stage capacity is subtracted in rounds; the rounds need a reproducible member
            // numbering (thread-major), computed here from a second look at the words (still valid: front e + 1
            // is not published before the last round)
            int my_total = 0;
            for (int w = t; w < need_words; w += kFlowThreads) my_total += __popc(ld_relaxed_u2(slot + w).x);
            int incl = my_total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += x;
            }
            if (lane == 31) sh.warp_any[warp] = incl;
            __syncthreads();
            before = incl - my_total;
            for (int q = 0; q < warp; ++q) before += sh.warp_any[q];
        }
        FLOW_T(4)
        ++r;
    }
    if (failed && t == 0) {
        hdr->error = 1u;
        for (int q = 0; q < a.nranks; ++q) reinterpret_cast<FlowHeader *>(a.peers.base[q])->error = 1u;
        if (a.host_error) {
            *a.host_error = 1u;
            __threadfence_system();
        }
    }
    if (blockIdx.x == 0 && t == 0) {
        if (!failed) hdr->epoch = e;
        *a.fronts_done = failed ? -1 : r;
    }
#if EVOXB_FLOW_TIMING
    if (a.timing && t == 0) {
        for (int i = 0; i < 8; ++i) a.timing[blockIdx.x * 8 + i] = tacc[i];
        if (blockIdx.x == 0) a.timing[192 * 8] = gridDim.x;
    }
#endif
}

// Size of front 0 = number of positions nobody dominates; decides which peel kernel does the work.
__global__ void __launch_bounds__(256)
front0_size_kernel(const int *__restrict__ count, int n, int *__restrict__ front0) {
    int mine = 0;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) mine += (count[j] == 0) ? 1 : 0;
    mine = __reduce_add_sync(0xffffffffu, mine);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(front0, mine);
}

// rank[perm[p]] = rank_sorted[p]: ranks leave the (sorted) position space
__global__ void __launch_bounds__(256)
unpermute_rank_kernel(const int *__restrict__ rank_sorted, const int *__restrict__ perm, int npad, const int *__restrict__ fronts_done,
                      const int *__restrict__ front0, int front0_limit, int *__restrict__ rank) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npad) return;
    if (front0 && *front0 > front0_limit) return;  // the other kernel wrote the ranks
    const int row = perm[p];
    if (row >= 0) rank[row] = (*fronts_done < 0) ? -1 : rank_sorted[p];
}

}  // namespace evoxb
