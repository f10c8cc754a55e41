// C-ABI entry points of libevoxb.so (see include/evoxb.h).  Hand-written sm_100a CUDA only:
// no CPU fallback, no library kernels on the compute path.
#include <dlfcn.h>

#include <mutex>
#include <vector>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "dominance.cuh"
#include "peel.cuh"
#include "peel_flow.cuh"
#include "sort.cuh"
#include "presort.cuh"
#include "crowding.cuh"
#include "select.cuh"
#include "variation.cuh"
#include "tail.cuh"
#include "nsga3.cuh"
#include "levels.cuh"

namespace evoxb {

std::atomic<unsigned long long> g_launches{0};

struct NcclId {
    char internal[128];  // == ncclUniqueId
};

}  // namespace evoxb

// Opaque communicator handed out by evoxb_comm_create (one per process / GPU).
struct evoxb_comm {
    void *nccl = nullptr;  // ncclComm_t
    int rank = 0, nranks = 1;
    int *h_flags = nullptr;  // pinned host ints for the per-batch termination check
    // peer-memory exchange (CUDA IPC); when it cannot be set up the NCCL all-gather path is used
    bool p2p = false;
    unsigned char *xbuf = nullptr;        // this rank's exchange buffer (cudaMalloc, exported to the peers)
    unsigned char *peer_ptr[16] = {};     // peers' exchange buffers mapped into this process
    unsigned char **peers_dev = nullptr;  // the same pointers in device memory, for the kernel
    int *fronts_dev = nullptr;            // fronts peeled by the last call (device int)
    size_t flow_off = 0;                  // offset of the data-flow kernel's exchange region inside every xbuf
    unsigned int *h_err = nullptr;        // pinned, mapped host word the kernels raise on a peer time-out (sticky)
    unsigned long long timeout_ns = 4000000000ull;  // how long a kernel waits for a peer before it gives up
    int last_path = 0;                    // which exchange the last call used: 0 NCCL all-gather, 1 grid-wide peer kernel, 2 data-flow kernel
};

namespace evoxb {

struct NdsLayout {
    int64_t n, npad, row_words;  // row_words: words per row of the rectangular layout (also the triangle's widest row)
    int64_t d_words;             // words of the bit matrix as stored
    bool packed;                 // triangular storage (TriRows): only for populations whose rectangle is too big
    int m, nchunks;
    bool tri;  // presorted population, upper-triangular matrix (m <= 16); otherwise full matrix in input order
    float *xt;
    uint32_t *D;
    int *count;
    int *list;
    PeelCtrl *ctrl;
    // presort state (tri only)
    SortBufs sb;
    int *perm;
    int *hp;
    int *tile_max;
    int *uid_row;
    int *uid_col;
    int *uid_row_q;  // uids indexed by the (key_0, S') position, before the within-chunk reordering
    int *uid_col_q;
    float *tile_min1;  // band of objective 1 of every 1024-tile
    float *tile_max1;
    // multi-GPU (column chunks owned cyclically): nranks == 1 means the whole matrix lives here
    ChunkMap cm;
    int nranks, slice_chunks;
    uint32_t *send_bits;  // [slice_chunks][128]
    uint32_t *recv_bits;  // [nranks][slice_chunks][128]
    int *totals;          // front sizes of the current batch of fronts
    int *cnt;             // members per region of the current front
    unsigned char *flow_ring;  // single-GPU exchange buffer of the data-flow peel (peel_flow.cuh)
    bool levels;               // the level form of K1 is used (levels.cuh)
    LevelBufs lv;
    size_t bytes;
};

static bool nds_use_tri(int m) { return m <= 16; }

// Which form of K1 builds the matrix of a presorted population: "levels" = the level kernel wherever it exists
// (m >= 2), "exact" = the float kernel, unset = by size.  Read on every call (tests switch it in-process).
// The levels pay where the build is long enough to absorb their set-up (about 0.1 ms of small kernels) and
// compares are many per pair: m >= 4 from 65536 rows on (measured on the B200).
static bool nds_use_levels(int64_t n, int m) {
    if (!nds_use_tri(m) || m < 2 || n <= 0) return false;
    if (const char *ev = getenv("EVOXB_BUILD_MODE")) {
        if (!strcmp(ev, "levels")) return true;
        if (!strcmp(ev, "exact")) return false;
    }
    return m >= 4 && n >= 65536;
}


constexpr int kFrontBatch = 4;  // fronts enqueued between two host checks of "front empty?" (multi-GPU)

static NdsLayout nds_layout(void *ws, int64_t n, int m, int64_t n_rows, int nranks = 1, int rank = 0,
                            bool sharded = false) {
    NdsLayout L{};
    L.n = n;
    L.m = m;
    L.tri = nds_use_tri(m);
    L.npad = round_up(n > 0 ? n : 1, kColPad);
    L.nchunks = (int)(L.npad / kChunkCols);
    L.nranks = nranks;
    L.slice_chunks = (int)div_up(L.nchunks, nranks);
    L.cm = ChunkMap{rank, nranks, rank < L.nchunks ? (L.nchunks - rank + nranks - 1) / nranks : 0};
    L.row_words = (nranks == 1) ? L.npad / 32 : (int64_t)(L.cm.nlocal > 0 ? L.cm.nlocal : 1) * kChunkWords;
    Carver c(ws);
    L.ctrl = c.take<PeelCtrl>(1);
    L.xt = c.take<float>((size_t)L.npad * (size_t)m);
    L.count = c.take<int>((size_t)L.npad);
    L.list = c.take<int>((size_t)L.npad);
    if (L.tri) {
        L.sb.cap = L.npad;
        L.sb.tiles = (int)div_up(L.sb.cap, kSortTile);
        for (int b = 0; b < 2; ++b) L.sb.vals[b] = c.take<uint32_t>((size_t)L.npad);  // permutation ping-pong
        L.sb.hist = c.take<uint32_t>((size_t)kRadix * (size_t)div_up(L.npad, kCoopTile));
        L.perm = c.take<int>((size_t)L.npad);
        L.hp = c.take<int>((size_t)L.npad);
        L.tile_max = c.take<int>((size_t)div_up(L.npad, kCoopTile));
        L.uid_row = c.take<int>((size_t)L.npad);
        L.uid_col = c.take<int>((size_t)L.npad);
        L.uid_row_q = c.take<int>((size_t)L.npad);
        L.uid_col_q = c.take<int>((size_t)L.npad);
        L.tile_min1 = c.take<float>((size_t)(L.npad / 1024));
        L.tile_max1 = c.take<float>((size_t)(L.npad / 1024));
    }
    if (sharded) {
        L.send_bits = c.take<uint32_t>((size_t)L.slice_chunks * kChunkWords);
        L.recv_bits = c.take<uint32_t>((size_t)nranks * L.slice_chunks * kChunkWords);
        L.totals = c.take<int>(64);
        L.cnt = c.take<int>(kMaxChunks);
    }
    L.flow_ring = c.take<unsigned char>(flow_ring_bytes(L.npad / 32));
    L.levels = L.tri && nds_use_levels(n, m);
    if (L.levels) {
        LevelBufs &v = L.lv;
        v.K = m - 1;
        v.W = (v.K + 1) / 2;
        v.q = c.take<uint32_t>((size_t)v.W * (size_t)L.npad);
        const size_t head = 2 * 16 + 4;  // cost (K <= 15 x 8 bytes, 8-byte aligned at the block start), list_count
        const size_t zwords = head + (size_t)v.K * (kLvlBins + 2 * (size_t)kLvlCount);
        v.zero0 = c.take<uint32_t>(zwords);
        v.zero_bytes = zwords * sizeof(uint32_t);
        v.cost = reinterpret_cast<unsigned long long *>(v.zero0);
        v.list_count = v.zero0 + 32;
        v.hist1 = v.zero0 + head;
        v.cnt = v.hist1 + (size_t)v.K * kLvlBins;
        v.kmax = v.cnt + (size_t)v.K * kLvlCount;
        v.kmin = c.take<uint32_t>((size_t)v.K * kLvlCount);
        v.bins = c.take<uint2>((size_t)v.K * kLvlBins);
        v.start = c.take<uint32_t>((size_t)v.K * (kLvlCount + 1));
        v.cursor = c.take<uint32_t>((size_t)v.K * kLvlCount);
        v.items4 = c.take<int4>((size_t)v.K * (size_t)L.npad * (v.W <= 2 ? 1 : 0) + (v.W <= 2 ? 0 : 1));
        v.items = v.W <= 2 ? reinterpret_cast<int2 *>(v.items4) : c.take<int2>((size_t)v.K * (size_t)L.npad);
        v.list_cap = (uint32_t)((size_t)v.K * (size_t)L.npad);
        v.list = c.take<unsigned long long>(v.list_cap);
        v.list_col = c.take<int>(v.list_cap);
    }
    // Only the triangle of the presorted matrix is ever written or read.  Storing just that (TriRows) halves
    // the workspace, but measured on the B200 it costs time: the build kernel's stores are 8-10 % slower with
    // the dense rows (c4: 4.38 -> 4.8 ms) and the variable row stride adds address arithmetic to every staged
    // row of count / peel.  So the rectangle (row stride = a power of two) stays the layout up to N = 262144
    // (8 GiB per GPU of 180) and the triangle is used beyond, where memory starts to matter more than 7 %.
    L.packed = L.tri && L.npad > 262144;
    if (L.packed) {
        const TriRows tr{L.row_words >> 2, L.row_words, 1, L.cm.first, L.cm.stride, L.cm.nlocal};
        L.d_words = n_rows > 0 ? tr.chunk_base(L.nchunks) * 4 : 0;
    } else {
        L.d_words = n_rows * L.row_words;
    }
    L.D = c.take<uint32_t>((size_t)L.d_words);
    L.bytes = c.total();
    return L;
}

static int check_common(const void *fitness, int64_t n, int m) {
    if (n < 0) return fail(EVOXB_E_INVALID, "n must be >= 0 (got %lld)", (long long)n);
    if (m < 1 || m > EVOXB_MAX_OBJECTIVES)
        return fail(EVOXB_E_INVALID, "m must be in [1, %d] (got %d)", EVOXB_MAX_OBJECTIVES, m);
    if (n > 0 && !fitness) return fail(EVOXB_E_INVALID, "fitness is null");
    if (n > (int64_t)1 << 24) return fail(EVOXB_E_UNSUPPORTED, "n = %lld exceeds 2^24", (long long)n);
    return 0;
}

// Grid of the persistent peel kernels: all co-resident CTAs (two per SM) when there is bandwidth to
// win, one per SM when the per-front latency dominates (rows_per_gpu = rows this GPU streams from).
static int peel_grid_for(int64_t rows_per_gpu, int max_blocks) {
    const int half = max_blocks / 2;
    static const char *ev = getenv("EVOXB_PEEL_SMALL_ROWS");  // development knob
    const int64_t limit = ev ? atoll(ev) : 98304;
    return (rows_per_gpu <= limit && half >= 1) ? half : max_blocks;
}

// Largest co-resident grid (capped at `cap_per_sm` CTAs per SM) for a cooperative kernel, cached per
// (device, kernel).  Every template instance has its own register count, so each is queried itself.
static int coop_blocks(const void *kern, int threads, int cap_per_sm, int *blocks) {
    struct Entry { int dev; const void *kern; int blocks; };
    static thread_local Entry cache[16];
    static thread_local int used = 0;
    int dev = 0;
    EVOXB_CUDA(cudaGetDevice(&dev));
    for (int i = 0; i < used; ++i)
        if (cache[i].dev == dev && cache[i].kern == kern) {
            *blocks = cache[i].blocks;
            return 0;
        }
    int sms = 0, per_sm = 0;
    EVOXB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    EVOXB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessorWithFlags(&per_sm, kern, threads, 0, cudaOccupancyDefault));
    if (per_sm < 1) return fail(EVOXB_E_CUDA, "a cooperative kernel does not fit on an SM");
    if (per_sm > cap_per_sm) per_sm = cap_per_sm;
    *blocks = sms * per_sm;
    if (used < 16) cache[used++] = Entry{dev, kern, *blocks};
    return 0;
}

}  // namespace evoxb


namespace evoxb {

static TriRows make_rows(const NdsLayout &L) {
    return TriRows{L.row_words >> 2, L.row_words, L.packed ? 1 : 0, L.cm.first, L.cm.stride, L.cm.nlocal};
}

// ---- data-flow peel (peel_flow.cuh): launch configuration ------------------------------------------
// Which peel runs: "flow" = peel_flow_kernel wherever it is applicable, "grid" = the cooperative-grid
// kernels of peel.cuh, unset = by population size (see peel_use_flow).  Experiment knob, read once.
static int peel_mode_env() {
    static const int mode = [] {
        const char *ev = getenv("EVOXB_PEEL_MODE");
        if (!ev) return 0;
        if (!strcmp(ev, "flow")) return 1;
        if (!strcmp(ev, "grid")) return 2;
        return 0;
    }();
    return mode;
}

// Default choice between the two peel kernels on one GPU (measured on the B200).
static bool peel_use_flow(int64_t n, int m) {
    if (m <= 3) return n <= 131072;
    if (m == 4) return n <= 65536;
    return n <= 32768;
}

struct FlowLaunch {
    FlowPlan plan;
    size_t smem;
    bool ok;
};

constexpr size_t kFlowSmemLimit = 200 * 1024;

// Picks two CTAs per SM when they fit, else one; the plan (column ownership) depends on that choice.
// Cached per (thread, device, shape): the plan walks every piece on the host.
static int flow_prepare(int nchunks, ChunkMap cm, bool tri, FlowLaunch *out) {
    struct Entry { int dev, nchunks, first, stride, nlocal, tri; FlowLaunch fl; };
    static thread_local Entry cache[8];
    static thread_local int used = 0, next = 0;
    int dev = 0;
    EVOXB_CUDA(cudaGetDevice(&dev));
    for (int i = 0; i < used; ++i) {
        const Entry &e = cache[i];
        if (e.dev == dev && e.nchunks == nchunks && e.first == cm.first && e.stride == cm.stride && e.nlocal == cm.nlocal &&
            e.tri == (int)tri) {
            *out = e.fl;
            return 0;
        }
    }
    FlowLaunch fl{};
    fl.ok = false;
    if (nchunks <= kFlowMaxChunks && cm.nlocal >= 1) {
        static std::mutex mu;
        static bool attr_set[64] = {};
        if (dev >= 0 && dev < 64) {
            std::lock_guard<std::mutex> lock(mu);
            if (!attr_set[dev]) {
                EVOXB_CUDA(cudaFuncSetAttribute(peel_flow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFlowSmemLimit));
                attr_set[dev] = true;
            }
        }
        int sms = 0;
        EVOXB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        for (int per_sm = 1; per_sm >= 1 && !fl.ok; --per_sm) {
            FlowPlan plan = flow_make_plan(nchunks, cm, tri, sms * per_sm);
            if (plan.grid < 1 || plan.max_pieces > kFlowMaxPieces) continue;
            const size_t smem = flow_smem_bytes(nchunks, plan.max_words);
            if (smem > kFlowSmemLimit) continue;
            int occ = 0;
            EVOXB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, peel_flow_kernel, kFlowThreads, smem));
            if (occ < per_sm) continue;
            fl.plan = plan;
            fl.smem = smem;
            fl.ok = true;
        }
    }
    Entry ne{dev, nchunks, cm.first, cm.stride, cm.nlocal, (int)tri, fl};
    if (used < 8) cache[used++] = ne;
    else { cache[next] = ne; next = (next + 1) % 8; }
    *out = fl;
    return 0;
}

// Development builds only (-DEVOXB_FLOW_TIMING=1): a device buffer the kernel's CTA 0 writes its per-phase
// clock totals to; printed by scripts/dev_flow_timing.py through evoxb_flow_timing_read.
static long long *flow_timing_buffer() {
#if EVOXB_FLOW_TIMING
    static long long *buf = nullptr;
    if (!buf) cudaMalloc((void **)&buf, sizeof(long long) * (192 * 8 + 8));
    return buf;
#else
    return nullptr;
#endif
}

// Data-flow peel on one GPU: ranks by (sorted) position into L.list, then back through the permutation.
static int run_peel_flow(const NdsLayout &L, int64_t n, int32_t *rank, const int *perm, const FlowLaunch &fl, int front0_limit,
                         cudaStream_t s) {
    FlowArgs a{};
    a.D = L.D;
    a.rows = make_rows(L);
    a.nchunks = L.nchunks;
    a.n = (int)n;
    a.tri = L.tri ? 1 : 0;
    a.cm = ChunkMap{0, 1, L.nchunks};
    a.count = L.count;
    a.rank_sorted = perm ? L.list : rank;
    a.tb = TileBands{L.tri ? L.tile_min1 : nullptr, L.tri ? L.tile_max1 : nullptr};
    a.peers.base[0] = L.flow_ring;
    a.me = 0;
    a.nranks = 1;
    a.grids[0] = fl.plan.grid;
    a.slot_words = L.npad / 32;
    a.max_words = fl.plan.max_words;
    a.target = fl.plan.target;
    a.fronts_done = &L.ctrl->fronts_done;
    a.host_error = nullptr;
    a.timeout_ns = 4000000000ull;
    a.timing = flow_timing_buffer();
    a.front0 = front0_limit >= 0 ? &L.ctrl->front0 : nullptr;
    a.front0_limit = front0_limit;
    if (front0_limit >= 0) {
        front0_size_kernel<<<(unsigned)div_up(n, 256 * 8), 256, 0, s>>>(L.count, (int)n, &L.ctrl->front0);
        EVOXB_LAUNCH_CHECK("front0_size_kernel");
    }
    EVOXB_CUDA(cudaMemsetAsync(L.flow_ring, 0, flow_ring_bytes(L.npad / 32), s));
    void *args[] = {&a};
    EVOXB_CUDA(cudaLaunchCooperativeKernel((const void *)peel_flow_kernel, dim3(fl.plan.grid), dim3(kFlowThreads), args, fl.smem, s));
    ++g_launches;
    if (perm) {
        unpermute_rank_kernel<<<(unsigned)(L.npad / 256), 256, 0, s>>>(L.list, perm, (int)L.npad, &L.ctrl->fronts_done, a.front0,
                                                                       front0_limit, rank);
        EVOXB_LAUNCH_CHECK("unpermute_rank_kernel");
    }
    return 0;
}

}  // namespace evoxb


namespace evoxb {

// Presort by (key_0, S') + sorted objective-major copy + uids (presort.cuh).  Deterministic, so
// every GPU of a sharded run computes the same permutation on its own.
static int run_presort(const float *fitness, int64_t n, int m, const NdsLayout &L, cudaStream_t s) {
    // order by (key_0, S'): stable LSD passes over S' (low 32 bits, then the few high bits), then key_0
    PresortPasses pp{};
    if (m > 1) {
        for (int sh = 0; sh < 32; sh += 8) { pp.kind[pp.count] = 0; pp.shift[pp.count++] = sh; }
        if (m > 2) { pp.kind[pp.count] = 1; pp.shift[pp.count++] = 0; }  // S' exceeds 32 bits only with >= 2 terms
    }
    pp.fast_first = pp.count;  // the key_0 passes alone are the fast path (presort.cuh)
    for (int sh = 0; sh < 32; sh += 8) { pp.kind[pp.count] = 2; pp.shift[pp.count++] = sh; }
    int blocks = 0;
    if (int rc = coop_blocks((const void *)presort_coop_kernel, 256, 2, &blocks)) return rc;
    const int tiles = (int)div_up(L.npad, kCoopTile);
    if (blocks > tiles) blocks = tiles;
    const float *x = fitness;
    int nn = (int)n, mm = m, npad = (int)L.npad;
    uint32_t *pa = L.sb.vals[0], *pb = L.sb.vals[1], *hist = L.sb.hist;
    int *tile_max = L.tile_max, *hp = L.hp, *perm = L.perm, *ur = L.uid_row, *uc = L.uid_col;
    int *urq = L.uid_row_q, *ucq = L.uid_col_q;
    float *tmin = L.tile_min1, *tmax = L.tile_max1;
    float *xt = L.xt;
    int *slow_flag = &L.ctrl->presort_slow;  // zeroed with the rest of PeelCtrl before this launch
    void *args[] = {&x, &nn, &mm, &npad, &pp, &pa, &pb, &hist, &tile_max, &hp, &xt, &perm, &ur, &uc, &urq, &ucq, &tmin, &tmax, &slow_flag};
    EVOXB_CUDA(cudaLaunchCooperativeKernel((const void *)presort_coop_kernel, dim3(blocks), dim3(256), args, 0, s));
    ++g_launches;
    return 0;
}

// K1 in level form (levels.cuh): levels, level kernel, float kernel as the on-device fall-back, fix-up.
// cm == nullptr: the whole matrix on this GPU.
template <int W>
static int run_level_build_w(const NdsLayout &L, int64_t n, int m, cudaStream_t s, const ChunkMap *cm) {
    const LevelBufs &v = L.lv;
    const int npad = (int)L.npad, nn = (int)n;
    int *fallback = &L.ctrl->level_fallback;  // zeroed with the rest of PeelCtrl
    EVOXB_CUDA(cudaMemsetAsync(v.zero0, 0, v.zero_bytes, s));
    EVOXB_CUDA(cudaMemsetAsync(v.kmin, 0xff, sizeof(uint32_t) * (size_t)v.K * kLvlCount, s));
    const unsigned pos_blocks = (unsigned)div_up(L.npad, 256);
    unsigned hist_blocks = (unsigned)div_up(n, 2048);
    if (hist_blocks > 296u) hist_blocks = 296u;
    level_hist_kernel<<<dim3(hist_blocks, (unsigned)v.K), 256, 0, s>>>(L.xt, npad, nn, v.hist1);
    EVOXB_LAUNCH_CHECK("level_hist_kernel");
    level_plan_kernel<<<(unsigned)v.K, 1024, 0, s>>>(v.hist1, v.bins);
    EVOXB_LAUNCH_CHECK("level_plan_kernel");
    level_assign_kernel<W><<<pos_blocks, 256, 0, s>>>(L.xt, npad, nn, v.K, v.bins, v.q, v.cnt, v.kmin, v.kmax);
    EVOXB_LAUNCH_CHECK("level_assign_kernel");
    level_scan_kernel<<<dim3(kLvlScanParts, (unsigned)v.K), 256, 0, s>>>(v.cnt, v.kmin, v.kmax, v.start, v.cursor, nn, v.cost,
                                                                       fallback);
    EVOXB_LAUNCH_CHECK("level_scan_kernel");
    if (W <= 2) {
        EVOXB_CUDA(cudaMemsetAsync(v.items4, 0xff, sizeof(int4) * (size_t)v.K * (size_t)L.npad, s));
        level_scatter_packed_kernel<<<dim3(pos_blocks, (unsigned)v.K), 256, 0, s>>>(L.xt, v.q, npad, nn, W, v.kmin, v.kmax,
                                                                                  v.cursor, v.items4, fallback);
        EVOXB_LAUNCH_CHECK("level_scatter_packed_kernel");
    } else {
        level_scatter_kernel<<<dim3(pos_blocks, (unsigned)v.K), 256, 0, s>>>(L.xt, v.q, npad, nn, v.kmin, v.kmax, v.cursor,
                                                                           v.items, fallback);
        EVOXB_LAUNCH_CHECK("level_scatter_kernel");
    }

    const TriRows rows = make_rows(L);
    if (cm && cm->nlocal <= 0) return 0;
    if (W <= 2) {
        level_fixup_find_packed_kernel<<<dim3(pos_blocks, (unsigned)v.K), 256, 0, s>>>(
            v.items4, npad, v.start, L.uid_row, L.uid_col, L.tile_min1, L.tile_max1, rows, cm ? *cm : ChunkMap{0, 1, L.nchunks},
            v.list, v.list_col, v.list_count, v.list_cap, fallback);
    } else {
        level_fixup_find_kernel<W><<<dim3(pos_blocks, (unsigned)v.K), 256, 0, s>>>(
            L.xt, v.q, npad, nn, v.kmin, v.kmax, v.start, v.items, L.uid_row, L.uid_col, L.tile_min1, L.tile_max1, rows,
            cm ? *cm : ChunkMap{0, 1, L.nchunks}, v.list, v.list_col, v.list_count, v.list_cap, fallback);
    }
    EVOXB_LAUNCH_CHECK("level_fixup_find_kernel");
    TriGrid g = make_tri_grid(L.npad, n);
    if (g.total > 0x7fffffffll) return fail(EVOXB_E_UNSUPPORTED, "too many tiles (%lld)", g.total);
    constexpr int kRowTiles = LvlCfg<W>::kRowTiles;
    dim3 grid((unsigned)lvl_total_groups(g, kRowTiles));  // level kernel: kRowTiles 128-row tiles per CTA
    long long gx = g.total;                    // float fall-back: one tile per loop iteration
    int gy = 1;
    if (cm) {
        g.sharded = 1;
        g.cm = *cm;
        if (cm->nlocal <= 0) return 0;
        if (g.nrt > 65535) return fail(EVOXB_E_UNSUPPORTED, "too many row tiles (%d)", g.nrt);
        grid = dim3((unsigned)(cm->nlocal * 4), (unsigned)div_up(g.nrt, kRowTiles));
        gx = cm->nlocal * 4;
        gy = g.nrt;
    }
    if (rows.tri)
        dominance_build_lvl_kernel<W, true><<<grid, kBuildThreads, 0, s>>>(v.q, L.npad, n, L.uid_row, L.uid_col, L.tile_min1,
                                                                          L.tile_max1, L.D, rows, g, L.count, fallback);
    else
        dominance_build_lvl_kernel<W, false><<<grid, kBuildThreads, 0, s>>>(v.q, L.npad, n, L.uid_row, L.uid_col, L.tile_min1,
                                                                           L.tile_max1, L.D, rows, g, L.count, fallback);
    EVOXB_LAUNCH_CHECK("dominance_build_lvl_kernel");
    if (int rc = launch_dominance_build_tri_fallback(L.xt, L.npad, n, m, L.uid_row, L.uid_col, L.tile_min1, L.tile_max1, L.D,
                                                     rows, g, gx, gy, fallback, s))
        return rc;
    level_fixup_apply_kernel<<<296, 256, 0, s>>>(v.list, v.list_col, v.list_count, v.list_cap, L.D, L.count, fallback);
    EVOXB_LAUNCH_CHECK("level_fixup_apply_kernel");
    return 0;
}

static int run_level_build(const NdsLayout &L, int64_t n, int m, cudaStream_t s, const ChunkMap *cm = nullptr) {
    switch (L.lv.W) {
        case 1: return run_level_build_w<1>(L, n, m, s, cm);
        case 2: return run_level_build_w<2>(L, n, m, s, cm);
        case 3: return run_level_build_w<3>(L, n, m, s, cm);
        case 4: return run_level_build_w<4>(L, n, m, s, cm);
        case 5: return run_level_build_w<5>(L, n, m, s, cm);
        case 6: return run_level_build_w<6>(L, n, m, s, cm);
        case 7: return run_level_build_w<7>(L, n, m, s, cm);
        case 8: return run_level_build_w<8>(L, n, m, s, cm);
    }
    return fail(EVOXB_E_UNSUPPORTED, "level build supports m <= 16 (got %d)", m);
}

// Enqueues the whole ranking (K0..K3) on `s`.  `L` must come from nds_layout(ws, n, m, n).
static int run_nds(const float *fitness, int64_t n, int m, int32_t *rank, const NdsLayout &L, cudaStream_t s,
                   Profiler &prof) {
    if (L.nchunks > kMaxChunks)
        return fail(EVOXB_E_UNSUPPORTED, "n = %lld exceeds %d rows", (long long)n, kMaxChunks * kChunkCols);
    prof.mark("prep", s);
    EVOXB_CUDA(cudaMemsetAsync(L.ctrl, 0, sizeof(PeelCtrl), s));
    EVOXB_CUDA(cudaMemsetAsync(L.count, 0, sizeof(int) * (size_t)L.npad, s));
    const int *perm = nullptr;
    if (L.tri) {
        if (int rc = run_presort(fitness, n, m, L, s)) return rc;
        perm = L.perm;
        prof.mark("dominance_build", s);
        if (L.levels) {
            if (int rc = run_level_build(L, n, m, s)) return rc;
        } else if (int rc = launch_dominance_build_tri(L.xt, L.npad, n, m, L.uid_row, L.uid_col, L.tile_min1, L.tile_max1, L.D, make_rows(L), s)) {
            return rc;
        }
    } else {
        transpose_pad_kernel<<<(unsigned)div_up(L.npad, 256), 256, 0, s>>>(fitness, n, m, L.npad, L.xt);
        EVOXB_LAUNCH_CHECK("transpose_pad_kernel");
        prof.mark("dominance_build", s);
        if (int rc = launch_dominance_build(L.xt, L.npad, 0, n, m, L.D, L.row_words, s)) return rc;
    }

    // entries per lane of the region-prefix scan: 1 up to 32 chunks (N <= 131072), 4 up to 128, else 16
    const void *peel_kern = L.nchunks <= 32    ? (const void *)peel_persistent_kernel<1>
                            : L.nchunks <= 128 ? (const void *)peel_persistent_kernel<4>
                                               : (const void *)peel_persistent_kernel<kMaxChunks / 32>;
    int blocks = 0;
    if (int rc = coop_blocks(peel_kern, kPeelThreads, 3, &blocks)) return rc;
    // Small populations are bound by the per-front latency, of which the grid-wide syncs and the
    // per-CTA flush are a part: one CTA per SM instead of two is faster there (measured on the B200:
    // N = 32768: 1.11 -> 0.95 ms, N = 65536: 1.54 -> 1.44 ms, N = 262144: 1.24 -> 1.68 ms).
    const int peel_blocks = peel_grid_for(n, blocks);
    const int tri = L.tri ? 1 : 0;

    prof.mark("dominator_count", s);
    const TileBands tb{L.tri ? L.tile_min1 : nullptr, L.tri ? L.tile_max1 : nullptr};
    // with the level build the counts are already there; the kernel then only runs if that build fell back
    dominator_count_kernel<<<blocks, kPeelThreads, 0, s>>>(L.D, make_rows(L), L.nchunks, n, tri, ChunkMap{0, 1, L.nchunks},
                                                           L.count, tb, L.levels ? &L.ctrl->level_fallback : nullptr);
    EVOXB_LAUNCH_CHECK("dominator_count_kernel");

    prof.mark("peel", s);
    // Which kernel peels: the data-flow kernel wins where fronts are many and small (its per-front cost is a
    // few microseconds instead of ~16), the grid-wide kernel where they are few and huge (bandwidth).  The
    // number of fronts is not known in advance, so for populations in the data-flow kernel's range BOTH are
    // enqueued and the size of front 0 decides on the device which one returns at once (no host round trip).
    FlowLaunch fl{};
    if (peel_mode_env() != 2) {
        if (int rc = flow_prepare(L.nchunks, ChunkMap{0, 1, L.nchunks}, L.tri, &fl)) return rc;
    }
    int front0_limit = -1;  // -1: only the grid-wide kernel runs
    if (fl.ok) {
        if (peel_mode_env() == 1) return run_peel_flow(L, n, rank, perm, fl, -1, s);
        if (peel_use_flow(n, m)) {
            // front 0 is the smallest of the early fronts (uniform M = 3: a tenth of the average front; NSGA-II on
            // DTLZ2: a third).  The data-flow kernel wins up to about 2000 rows per front, which n / 32 separates in the
            // measured cases (uniform M = 3..4 at 32768..131072 rows, NSGA-II on DTLZ2 from random to converged)
            front0_limit = (int)(n / 32);
            if (int rc = run_peel_flow(L, n, rank, perm, fl, front0_limit, s)) return rc;
        }
    }
    {
        const uint32_t *D = L.D;
        TriRows row_words = make_rows(L);
        int nchunks = L.nchunks;
        int nn = (int)n;
        int tri_arg = tri;
        int *count = L.count;
        int *rk = rank;
        int *list = L.list;
        PeelCtrl *ctrl = L.ctrl;
        TileBands tbv = tb;
        void *args[] = {&D, &row_words, &nchunks, &nn, &tri_arg, &count, &rk, &perm, &list, &ctrl, &tbv, &front0_limit};
        EVOXB_CUDA(cudaLaunchCooperativeKernel(peel_kern, dim3(peel_blocks), dim3(kPeelThreads), args, 0, s));
        ++g_launches;
    }
    return 0;
}

// ---- multi-GPU: NCCL (loaded at run time from the process, torch ships libnccl.so.2) --------------
struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(void *) = nullptr;
    int (*CommInitRank)(void **, int, NcclId, int) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    bool ok = false;
};

static NcclApi &nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *env = getenv("EVOXB_NCCL_LIB");
        const char *names[] = {env, "libnccl.so.2", "libnccl.so"};
        for (const char *nm : names) {
            if (!nm) continue;
            api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) return;
        api.GetUniqueId = (int (*)(void *))dlsym(api.handle, "ncclGetUniqueId");
        api.CommInitRank = (int (*)(void **, int, NcclId, int))dlsym(api.handle, "ncclCommInitRank");
        api.CommDestroy = (int (*)(void *))dlsym(api.handle, "ncclCommDestroy");
        api.AllGather = (int (*)(const void *, void *, size_t, int, void *, cudaStream_t))dlsym(api.handle, "ncclAllGather");
        api.GetErrorString = (const char *(*)(int))dlsym(api.handle, "ncclGetErrorString");
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.GetErrorString;
    });
    return api;
}

#define EVOXB_NCCL(call)                                                                            \
    do {                                                                                            \
        int r__ = (call);                                                                           \
        if (r__ != 0)                                                                               \
            return ::evoxb::fail(EVOXB_E_NCCL, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, \
                                 nccl_api().GetErrorString(r__));                                   \
    } while (0)

// Which fused kernel peels across GPUs.  The data-flow kernel needs every rank to own at least one chunk (a
// rank without columns would publish nothing, and nobody would wait for it to have read a buffer before it is
// reused).  Measured on 2 and 8 B200s at N = 262144, M = 5 (10 000 members per front) the grid-wide kernel
// is faster (8 GPUs: peel 0.98 ms against 1.59 ms), so the data-flow kernel is kept for the populations where
// it wins on one GPU too: many small fronts.  Must be the same on every rank.
static bool sharded_use_flow(int64_t n, int m, int nchunks, int nranks) {
    if (nchunks < nranks || nchunks > kFlowMaxChunks) return false;
    static const int mode = [] {
        const char *ev = getenv("EVOXB_PEEL_MODE");
        return !ev ? 0 : !strcmp(ev, "flow") ? 1 : !strcmp(ev, "grid") ? 2 : 0;
    }();
    if (mode) return mode == 1;
    return peel_use_flow(n, m);
}

// Row-sharded^T ranking: every GPU presorts (same permutation everywhere), builds and keeps the
// matrix columns of its own chunks for ALL rows, and owns count[] of those columns.  Per front the
// GPUs exchange one bit per individual (ncclAllGather of N/8 bytes in total); no other traffic.
// The host checks "front empty?" once per kFrontBatch fronts (one 4-byte-per-front D2H + stream sync).
static int run_nds_sharded(const float *fitness, int64_t n, int m, int32_t *rank, const NdsLayout &L, evoxb_comm *comm,
                           cudaStream_t s, Profiler &prof) {
    if (comm->h_err && *reinterpret_cast<volatile unsigned int *>(comm->h_err))
        return fail(EVOXB_E_NCCL, "an earlier call on this communicator gave up waiting for a peer GPU (its ranks are -1); "
                                  "the communicator cannot be used any more: destroy it and create a new one on every rank");
    if (!L.tri) return fail(EVOXB_E_UNSUPPORTED, "the multi-GPU path supports m <= 16 (got %d)", m);
    if (L.nchunks > kMaxChunks)
        return fail(EVOXB_E_UNSUPPORTED, "n = %lld exceeds %d rows", (long long)n, kMaxChunks * kChunkCols);
    NcclApi &nc = nccl_api();
    prof.mark("prep", s);
    EVOXB_CUDA(cudaMemsetAsync(L.ctrl, 0, sizeof(PeelCtrl), s));
    EVOXB_CUDA(cudaMemsetAsync(L.count, 0, sizeof(int) * (size_t)L.npad, s));
    if (int rc = run_presort(fitness, n, m, L, s)) return rc;

    prof.mark("dominance_build", s);
    if (L.levels) {
        if (int rc = run_level_build(L, n, m, s, &L.cm)) return rc;
    } else if (int rc = launch_dominance_build_tri(L.xt, L.npad, n, m, L.uid_row, L.uid_col, L.tile_min1, L.tile_max1, L.D, make_rows(L), s, &L.cm)) {
        return rc;
    }

    const void *p2p_kern = L.nchunks <= 32    ? (const void *)peel_p2p_kernel<1>
                           : L.nchunks <= 128 ? (const void *)peel_p2p_kernel<4>
                                              : (const void *)peel_p2p_kernel<kMaxChunks / 32>;
    int blocks = 0;
    if (int rc = coop_blocks(p2p_kern, kPeelThreads, 3, &blocks)) return rc;
    const TileBands tb{L.tile_min1, L.tile_max1};
    prof.mark("dominator_count", s);
    if (L.cm.nlocal > 0) {
        dominator_count_kernel<<<blocks, kPeelThreads, 0, s>>>(L.D, make_rows(L), L.nchunks, n, 1, L.cm, L.count, tb,
                                                               L.levels ? &L.ctrl->level_fallback : nullptr);
        EVOXB_LAUNCH_CHECK("dominator_count_kernel");
    }

    prof.mark("peel", s);
    const size_t slice_words = (size_t)L.slice_chunks * kChunkWords;
    if (comm->p2p && sharded_use_flow(n, m, L.nchunks, L.nranks)) {
        // data-flow peel across the GPUs: the same kernel as on one GPU, its tagged words and flags also stored
        // into every peer's exchange region over NVLink (peel_flow.cuh)
        FlowArgs a{};
        bool all_ok = true;
        FlowLaunch mine{};
        for (int q = 0; q < L.nranks && all_ok; ++q) {
            const ChunkMap cq{q, L.nranks, q < L.nchunks ? (L.nchunks - q + L.nranks - 1) / L.nranks : 0};
            FlowLaunch fq{};
            if (int rc = flow_prepare(L.nchunks, cq, true, &fq)) return rc;
            all_ok = all_ok && fq.ok;
            a.grids[q] = fq.plan.grid;
            if (q == comm->rank) mine = fq;
        }
        if (all_ok) {
            a.D = L.D;
            a.rows = make_rows(L);
            a.nchunks = L.nchunks;
            a.n = (int)n;
            a.tri = 1;
            a.cm = L.cm;
            a.count = L.count;
            a.rank_sorted = L.list;
            a.tb = tb;
            for (int q = 0; q < L.nranks; ++q) a.peers.base[q] = comm->peer_ptr[q] + comm->flow_off;
            a.me = comm->rank;
            a.nranks = L.nranks;
            a.slot_words = (int64_t)kFlowMaxChunks * kChunkWords;
                    a.max_words = mine.plan.max_words;
            a.target = mine.plan.target;
            a.fronts_done = comm->fronts_dev;
            a.host_error = comm->h_err;
            a.timeout_ns = comm->timeout_ns;
            a.timing = nullptr;
            a.front0 = nullptr;
            a.front0_limit = -1;
            void *args[] = {&a};
            EVOXB_CUDA(cudaLaunchCooperativeKernel((const void *)peel_flow_kernel, dim3(mine.plan.grid), dim3(kFlowThreads), args,
                                                   mine.smem, s));
            ++g_launches;
            unpermute_rank_kernel<<<(unsigned)(L.npad / 256), 256, 0, s>>>(L.list, L.perm, (int)L.npad, comm->fronts_dev, nullptr, -1,
                                                                           rank);
            EVOXB_LAUNCH_CHECK("unpermute_rank_kernel");
            comm->last_path = 2;
            return 0;
        }
    }
    if (comm->p2p && (size_t)L.nranks * slice_words <= (size_t)kXchgSliceWordsCap) {
        comm->last_path = 1;
        // fused path: one persistent kernel per GPU exchanges the front bitmask over peer memory
        const uint32_t *D = L.D;
        TriRows row_words = make_rows(L);
        int nchunks = L.nchunks, nn = (int)n;
        ChunkMap cm = L.cm;
        int *count = L.count, *rk = rank, *list = L.list, *cnt = L.cnt, *fronts = comm->fronts_dev;
        const int *perm = L.perm;
        XchgCtx x{comm->peers_dev, comm->rank, comm->nranks, L.slice_chunks};
        TileBands tbv = tb;
        unsigned int *herr = comm->h_err;
        unsigned long long tmo = comm->timeout_ns;
        void *args[] = {&D, &row_words, &nchunks, &nn, &cm, &count, &rk, &perm, &list, &cnt, &x, &fronts, &tbv, &herr, &tmo};
        // across GPUs every rank still streams all n member rows and the extra CTAs also widen the
        // exchange, so the full grid wins there (measured at 4 and 8 GPUs: 0.93 / 0.83 ms against
        // 1.13 / 0.96 ms with one CTA per SM)
        const int p2p_blocks = L.nranks > 1 ? blocks : peel_grid_for(n, blocks);
        EVOXB_CUDA(cudaLaunchCooperativeKernel(p2p_kern, dim3(p2p_blocks), dim3(kPeelThreads), args, 0, s));
        ++g_launches;
        poison_ranks_kernel<<<(unsigned)div_up(n, 256), 256, 0, s>>>(rank, (int)n, comm->fronts_dev);  // -1 everywhere after a time-out
        EVOXB_LAUNCH_CHECK("poison_ranks_kernel");
        return 0;
    }
    comm->last_path = 0;
    const unsigned detect_blocks = (unsigned)div_up((int64_t)(L.cm.nlocal > 0 ? L.cm.nlocal : 1) * kChunkWords * 32, 256);
    int r = 0;
    for (;;) {
        EVOXB_CUDA(cudaMemsetAsync(L.totals, 0, sizeof(int) * kFrontBatch, s));
        for (int b = 0; b < kFrontBatch; ++b) {
            EVOXB_CUDA(cudaMemsetAsync(L.send_bits, 0, sizeof(uint32_t) * slice_words, s));
            if (L.cm.nlocal > 0) {
                front_detect_owned_kernel<<<detect_blocks, 256, 0, s>>>(L.count, (int)n, L.cm, L.send_bits);
                EVOXB_LAUNCH_CHECK("front_detect_owned_kernel");
            }
            EVOXB_NCCL(nc.AllGather(L.send_bits, L.recv_bits, slice_words, /*ncclUint32*/ 3, comm->nccl, s));
            front_expand_kernel<<<L.nchunks, kChunkWords, 0, s>>>(L.recv_bits, L.nranks, L.slice_chunks, L.nchunks, r + b,
                                                                  rank, L.perm, L.list, L.cnt, L.totals + b);
            EVOXB_LAUNCH_CHECK("front_expand_kernel");
            if (L.cm.nlocal > 0) {
                front_subtract_kernel<<<blocks, kPeelThreads, 0, s>>>(L.D, make_rows(L), L.nchunks, (int)n, 1, L.cm, L.list,
                                                                      L.cnt, L.count, tb);
                EVOXB_LAUNCH_CHECK("front_subtract_kernel");
            }
        }
        EVOXB_CUDA(cudaMemcpyAsync(comm->h_flags, L.totals, sizeof(int) * kFrontBatch, cudaMemcpyDeviceToHost, s));
        EVOXB_CUDA(cudaStreamSynchronize(s));
        bool done = false;
        for (int b = 0; b < kFrontBatch; ++b) done = done || (comm->h_flags[b] == 0);  // identical on every GPU
        r += kFrontBatch;
        if (done) break;
        if (r > n + kFrontBatch) return fail(EVOXB_E_CUDA, "front loop did not terminate");
    }
    return 0;
}

// Maps every rank's exchange buffer into this process (CUDA IPC over NVLink peer access).  Handles
// travel through one ncclAllGather.  Any failure leaves comm->p2p == false (NCCL path); the decision
// is made collectively so that all ranks take the same path.
static void setup_peer_exchange(evoxb_comm *c) {
    NcclApi &nc = nccl_api();
    const int P = c->nranks;
    if (P > kXchgMaxRanks || getenv("EVOXB_NO_P2P")) return;
    cudaStream_t s = nullptr;
    bool ok = true;
    unsigned char *hbuf_dev = nullptr;
    struct Slot { cudaIpcMemHandle_t h; int ok; int pad[15]; };
    static_assert(sizeof(Slot) == 128, "slot size");
    Slot mine{};
    std::vector<Slot> all((size_t)P);
    c->flow_off = (xchg_buffer_bytes() + 255) & ~(size_t)255;
    const size_t xbytes = c->flow_off + flow_ring_bytes((int64_t)kFlowMaxChunks * kChunkWords);
    ok = ok && cudaMalloc((void **)&c->xbuf, xbytes) == cudaSuccess;
    ok = ok && cudaMemset(c->xbuf, 0, xbytes) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&c->fronts_dev, sizeof(int)) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&c->peers_dev, sizeof(unsigned char *) * kXchgMaxRanks) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&hbuf_dev, sizeof(Slot) * (size_t)(P + 1)) == cudaSuccess;
    if (ok && P > 1) ok = cudaIpcGetMemHandle(&mine.h, c->xbuf) == cudaSuccess;
    mine.ok = ok ? 1 : 0;
    if (hbuf_dev) {
        // collective: everybody takes part even if its own setup failed, so nobody hangs
        cudaMemcpy(hbuf_dev + sizeof(Slot) * (size_t)P, &mine, sizeof(Slot), cudaMemcpyHostToDevice);
        int r = nc.AllGather(hbuf_dev + sizeof(Slot) * (size_t)P, hbuf_dev, sizeof(Slot), /*ncclUint8*/ 1, c->nccl, s);
        cudaError_t e = cudaStreamSynchronize(s);
        if (r != 0 || e != cudaSuccess) ok = false;
        else cudaMemcpy(all.data(), hbuf_dev, sizeof(Slot) * (size_t)P, cudaMemcpyDeviceToHost);
    }
    for (int q = 0; ok && q < P; ++q) ok = all[(size_t)q].ok == 1;
    for (int q = 0; ok && q < P; ++q) {
        if (q == c->rank) {
            c->peer_ptr[q] = c->xbuf;
        } else {
            void *ptr = nullptr;
            if (cudaIpcOpenMemHandle(&ptr, all[(size_t)q].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                ok = false;
                cudaGetLastError();
            }
            c->peer_ptr[q] = static_cast<unsigned char *>(ptr);
        }
    }
    // second collective round: the fused path is used only if EVERY rank mapped every peer
    if (hbuf_dev) {
        mine.ok = ok ? 1 : 0;
        cudaMemcpy(hbuf_dev + sizeof(Slot) * (size_t)P, &mine, sizeof(Slot), cudaMemcpyHostToDevice);
        int r = nc.AllGather(hbuf_dev + sizeof(Slot) * (size_t)P, hbuf_dev, sizeof(Slot), 1, c->nccl, s);
        cudaError_t e = cudaStreamSynchronize(s);
        if (r != 0 || e != cudaSuccess) ok = false;
        else {
            cudaMemcpy(all.data(), hbuf_dev, sizeof(Slot) * (size_t)P, cudaMemcpyDeviceToHost);
            for (int q = 0; q < P; ++q) ok = ok && all[(size_t)q].ok == 1;
        }
        cudaFree(hbuf_dev);
    } else {
        ok = false;
    }
    if (ok) ok = cudaMemcpy(c->peers_dev, c->peer_ptr, sizeof(unsigned char *) * kXchgMaxRanks, cudaMemcpyHostToDevice) == cudaSuccess;
    cudaGetLastError();
    c->p2p = ok;
}

struct SelectLayout {
    NdsLayout nds;
    CrowdLayout crowd;
    SortBufs sb;
    // fused tail (tail.cuh)
    int *tile_a, *tile_b, *info;
    uint32_t *list_a, *list_b, *tbuf0, *tbuf1, *thist;
    float *contrib;
    int64_t tcap;
    int *hist;
    int *worst;
    uint8_t *mask;
    int *rank;
    float *crowd_out;
    size_t bytes;
};

struct ArgsortLayout {
    SortBufs sb;
    size_t bytes;
};

static ArgsortLayout argsort_layout(void *ws, int64_t n) {
    ArgsortLayout A{};
    Carver c(ws);
    const int64_t n1 = n > 0 ? n : 1;
    A.sb.cap = round_up(n1, 4);
    A.sb.tiles = (int)div_up(A.sb.cap, kSortTile);
    for (int b = 0; b < 2; ++b) {
        A.sb.keys[b] = c.take<uint32_t>((size_t)A.sb.cap);
        A.sb.vals[b] = c.take<uint32_t>((size_t)A.sb.cap);
    }
    A.sb.hist = c.take<uint32_t>(sort_hist_words(A.sb.cap, 1));
    A.bytes = c.total();
    return A;
}

static SelectLayout select_layout(void *ws, int64_t n, int m, int nranks = 1, int rank = 0, bool sharded = false) {
    SelectLayout S{};
    Carver c(ws);
    const int64_t n1 = n > 0 ? n : 1;
    S.worst = c.take<int>(64);
    S.hist = c.take<int>((size_t)n1);
    S.mask = c.take<uint8_t>((size_t)n1);
    S.rank = c.take<int>((size_t)n1);
    S.crowd_out = c.take<float>((size_t)n1);
    S.sb.cap = round_up(n1, 4);
    S.sb.tiles = (int)div_up(S.sb.cap, kSortTile);
    for (int b = 0; b < 2; ++b) {
        S.sb.keys[b] = c.take<uint32_t>((size_t)S.sb.cap);
        S.sb.vals[b] = c.take<uint32_t>((size_t)S.sb.cap);
    }
    S.sb.hist = c.take<uint32_t>(sort_hist_words(S.sb.cap, 1));
    {
        const size_t tiles = (size_t)div_up(n1, kCoopTile);
        const size_t segs = (size_t)(m > 2 ? m : 2);
        S.tcap = round_up(n1, 4);
        S.tile_a = c.take<int>(tiles);
        S.tile_b = c.take<int>(tiles);
        S.info = c.take<int>(64);
        S.list_a = c.take<uint32_t>((size_t)n1);
        S.list_b = c.take<uint32_t>((size_t)n1);
        S.tbuf0 = c.take<uint32_t>(segs * (size_t)S.tcap);
        S.tbuf1 = c.take<uint32_t>(segs * (size_t)S.tcap);
        S.thist = c.take<uint32_t>(segs * tiles * kRadix);
        S.contrib = c.take<float>((size_t)m * (size_t)n1);
    }
    {
        char *sub = c.take<char>(0);
        S.crowd = crowd_layout(sub, n, m);
        c.take<char>(S.crowd.bytes);
    }
    {
        char *sub = c.take<char>(0);
        S.nds = nds_layout(sub, n, m, n, nranks, rank, sharded);
        c.take<char>(S.nds.bytes);
    }
    S.bytes = c.total();
    return S;
}

}  // namespace evoxb

using namespace evoxb;

extern "C" {

EVOXB_API int evoxb_version(void) { return EVOXB_VERSION; }

EVOXB_API const char *evoxb_last_error(void) { return err_buf(); }

EVOXB_API int evoxb_profile_enable(int on) {
    profiler().enabled = on != 0;
    return 0;
}

#if EVOXB_FLOW_TIMING
EVOXB_API int evoxb_flow_timing_read(long long *out) {  // out[192 * 8 + 8]
    cudaDeviceSynchronize();
    return cudaMemcpy(out, flow_timing_buffer(), sizeof(long long) * (192 * 8 + 8), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -3;
}
#endif

EVOXB_API unsigned long long evoxb_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

EVOXB_API int evoxb_profile_read(float *ms, const char **names, int cap) {
    Profiler &p = profiler();
    if (!p.enabled || p.n < 2) return 0;
    const int stages = p.n - 1;
    if (cudaEventSynchronize(p.ev[p.n - 1]) != cudaSuccess)
        return fail(EVOXB_E_CUDA, "cudaEventSynchronize failed: %s", cudaGetErrorString(cudaGetLastError()));
    for (int i = 0; i < stages && i < cap; ++i) {
        float t = 0.f;
        cudaEventElapsedTime(&t, p.ev[i], p.ev[i + 1]);
        if (ms) ms[i] = t;
        if (names) names[i] = p.name[i];
    }
    return stages;
}

EVOXB_API int evoxb_dominate_relation_f32(const float *x, int64_t nx, const float *y, int64_t ny, int m, uint8_t *out,
                                evoxb_stream_t stream) {
    if (nx < 0 || ny < 0) return fail(EVOXB_E_INVALID, "negative size");
    if (m < 1 || m > EVOXB_MAX_OBJECTIVES) return fail(EVOXB_E_INVALID, "m must be in [1, %d]", EVOXB_MAX_OBJECTIVES);
    if (nx == 0 || ny == 0) return 0;
    if (!x || !y || !out) return fail(EVOXB_E_INVALID, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    for (int64_t i0 = 0; i0 < nx; i0 += 32768) {  // gridDim.y limit
        const int64_t rows = (nx - i0) < 32768 ? (nx - i0) : 32768;
        dim3 grid((unsigned)div_up(ny, 256), (unsigned)rows);
        dominate_relation_dense_kernel<<<grid, 256, 0, s>>>(x + i0 * m, rows, y, ny, m, out + i0 * ny);
        EVOXB_LAUNCH_CHECK("dominate_relation_dense_kernel");
    }
    return 0;
}

EVOXB_API int evoxb_nds_workspace_bytes(int64_t n, int m, size_t *bytes) {
    if (!bytes) return fail(EVOXB_E_INVALID, "bytes is null");
    if (int rc = check_common((const void *)1, n, m)) return rc;
    *bytes = nds_layout(nullptr, n, m, n).bytes;
    return 0;
}

EVOXB_API int evoxb_nds_rank_f32(const float *fitness, int64_t n, int m, int32_t *rank, void *ws, size_t ws_bytes,
                       evoxb_stream_t stream) {
    if (int rc = check_common(fitness, n, m)) return rc;
    if (n == 0) return 0;
    if (!rank) return fail(EVOXB_E_INVALID, "rank is null");
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255)) return fail(EVOXB_E_WORKSPACE, "workspace is null or not 256-byte aligned");
    NdsLayout L = nds_layout(ws, n, m, n);
    if (ws_bytes < L.bytes)
        return fail(EVOXB_E_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, L.bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Profiler &prof = profiler();
    prof.begin();
    if (int rc = run_nds(fitness, n, m, rank, L, s, prof)) return rc;
    prof.mark("end", s);
    return 0;
}

EVOXB_API int evoxb_nds_build_path(const void *ws, int64_t n, int m, evoxb_stream_t stream, int *path) {
    if (!path) return fail(EVOXB_E_INVALID, "path is null");
    if (int rc = check_common((const void *)1, n, m)) return rc;
    *path = 0;
    if (n == 0 || !nds_use_tri(m) || !nds_use_levels(n, m)) return 0;
    if (!ws) {  // host-side answer only: the levels would be used for this shape
        *path = 1;
        return 0;
    }
    int flag = 0;  // PeelCtrl is the first block of every ranking workspace
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    EVOXB_CUDA(cudaMemcpyAsync(&flag, &static_cast<const PeelCtrl *>(ws)->level_fallback, sizeof(int), cudaMemcpyDeviceToHost, s));
    EVOXB_CUDA(cudaStreamSynchronize(s));
    *path = flag ? 2 : 1;
    return 0;
}

EVOXB_API int evoxb_crowding_workspace_bytes(int64_t n, int m, size_t *bytes) {
    if (!bytes) return fail(EVOXB_E_INVALID, "bytes is null");
    if (int rc = check_common((const void *)1, n, m)) return rc;
    *bytes = crowd_layout(nullptr, n, m).bytes;
    return 0;
}

EVOXB_API int evoxb_crowding_f32(const float *fitness, const uint8_t *mask, int64_t n, int m, float *dist, void *ws,
                       size_t ws_bytes, evoxb_stream_t stream) {
    if (int rc = check_common(fitness, n, m)) return rc;
    if (n == 0) return 0;
    if (!dist) return fail(EVOXB_E_INVALID, "dist is null");
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255)) return fail(EVOXB_E_WORKSPACE, "workspace is null or not 256-byte aligned");
    CrowdLayout L = crowd_layout(ws, n, m);
    if (ws_bytes < L.bytes)
        return fail(EVOXB_E_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, L.bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Profiler &prof = profiler();
    prof.begin();
    prof.mark("crowding", s);
    if (int rc = launch_crowding(fitness, mask, n, m, dist, L, s)) return rc;
    prof.mark("end", s);
    return 0;
}

EVOXB_API int evoxb_nsga2_select_workspace_bytes(int64_t n, int m, size_t *bytes) {
    if (!bytes) return fail(EVOXB_E_INVALID, "bytes is null");
    if (int rc = check_common((const void *)1, n, m)) return rc;
    *bytes = select_layout(nullptr, n, m).bytes;
    return 0;
}

// variant 0: NSGA2.tell (nsga2.py:90-96): k-th rank at index pop_size, secondary key -crowding distance.
// variant 1: operator non_dominate (non_dominate.py:181-195): k-th rank at index topk-1, and the
//            secondary key is -argsort(crowding distance) taken positionally (the reference's quirk, :192-194).
static int select_impl(int variant, const float *fitness, int64_t n, int m, int64_t pop_size, int32_t *order,
                       int32_t *rank_out, float *crowd_out, void *ws, size_t ws_bytes, evoxb_stream_t stream,
                       evoxb_comm *comm = nullptr) {
    if (int rc = check_common(fitness, n, m)) return rc;
    const int64_t kth = variant == 0 ? pop_size : pop_size - 1;
    // rank[order[kth]] must exist (nsga2.py:92 / non_dominate.py:190)
    if (kth < 0 || kth >= n || pop_size > n)
        return fail(EVOXB_E_INVALID, "%s out of range (got %lld, n = %lld)", variant == 0 ? "pop_size" : "topk",
                    (long long)pop_size, (long long)n);
    if (!order && pop_size > 0) return fail(EVOXB_E_INVALID, "order is null");
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255)) return fail(EVOXB_E_WORKSPACE, "workspace is null or not 256-byte aligned");
    SelectLayout S = select_layout(ws, n, m, comm ? comm->nranks : 1, comm ? comm->rank : 0, comm != nullptr);
    if (ws_bytes < S.bytes)
        return fail(EVOXB_E_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, S.bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Profiler &prof = profiler();
    prof.begin();
    int *rank = rank_out ? rank_out : S.rank;
    float *crowd = crowd_out ? crowd_out : S.crowd_out;
    if (comm) {
        if (int rc = run_nds_sharded(fitness, n, m, rank, S.nds, comm, s, prof)) return rc;
    } else {
        if (int rc = run_nds(fitness, n, m, rank, S.nds, s, prof)) return rc;
    }

    if (variant == 0) {
        // cut-off rank, mask, crowding and survivor order in one cooperative launch (tail.cuh)
        prof.mark("selection_tail", s);
        EVOXB_CUDA(cudaMemsetAsync(S.hist, 0, sizeof(int) * (size_t)n, s));
        TailArgs ta{};
        ta.fitness = fitness; ta.rank = rank; ta.n = (int)n; ta.m = m; ta.kth = (int)kth; ta.pop = (int)pop_size;
        ta.order = order; ta.crowd = crowd; ta.hist = S.hist; ta.tile_a = S.tile_a; ta.tile_b = S.tile_b;
        ta.list_a = S.list_a; ta.list_b = S.list_b; ta.buf0 = S.tbuf0; ta.buf1 = S.tbuf1; ta.hist_g = S.thist;
        ta.contrib = S.contrib; ta.cap = S.tcap; ta.info = S.info;
        int tb = 0;
        if (int rc = coop_blocks((const void *)select_tail_kernel, 256, 2, &tb)) return rc;
        const int want = (int)div_up(n, kCoopTile) * (m > 2 ? m : 2);
        if (tb > want) tb = want < 1 ? 1 : want;
        void *targs[] = {&ta};
        EVOXB_CUDA(cudaLaunchCooperativeKernel((const void *)select_tail_kernel, dim3(tb), dim3(256), targs, 0, s));
        ++g_launches;
        prof.mark("end", s);
        return 0;
    }

    prof.mark("worst_rank", s);
    const unsigned gb = (unsigned)div_up(n, 256);
    EVOXB_CUDA(cudaMemsetAsync(S.hist, 0, sizeof(int) * (size_t)n, s));
    rank_hist_kernel<<<gb, 256, 0, s>>>(rank, n, S.hist);
    EVOXB_LAUNCH_CHECK("rank_hist_kernel");
    kth_rank_kernel<<<1, 1024, 0, s>>>(S.hist, n, kth, S.worst);
    EVOXB_LAUNCH_CHECK("kth_rank_kernel");
    rank_mask_kernel<<<gb, 256, 0, s>>>(rank, n, S.worst, S.mask);
    EVOXB_LAUNCH_CHECK("rank_mask_kernel");

    prof.mark("crowding", s);
    if (int rc = launch_crowding(fitness, S.mask, n, m, crowd, S.crowd, s)) return rc;

    prof.mark("select_sort", s);
    // variant 1 only from here on (variant 0 returned above)
    int cur = 0;
    crowd_keys_kernel<<<gb, 256, 0, s>>>(crowd, n, S.sb.keys[0], S.sb.vals[0]);
    EVOXB_LAUNCH_CHECK("crowd_keys_kernel");
    if (int rc = radix_sort_pairs(S.sb, 0, 1, n, nullptr, 0, 32, s, &cur)) return rc;  // vals = argsort(crowd)
    neg_index_keys_kernel<<<gb, 256, 0, s>>>(S.sb.vals[cur], n, S.sb.keys[cur ^ 1], S.sb.vals[cur ^ 1]);
    EVOXB_LAUNCH_CHECK("neg_index_keys_kernel");
    cur ^= 1;
    if (int rc = radix_sort_pairs(S.sb, cur, 1, n, nullptr, 0, round_up(bits_for(n), 8), s, &cur)) return rc;
    gather_rank_keys_kernel<<<gb, 256, 0, s>>>(rank, S.sb.vals[cur], n, S.sb.keys[cur]);
    EVOXB_LAUNCH_CHECK("gather_rank_keys_kernel");
    if (int rc = radix_sort_pairs(S.sb, cur, 1, n, nullptr, 0, round_up(bits_for(n), 8), s, &cur)) return rc;
    if (pop_size > 0) {
        take_prefix_kernel<<<(unsigned)div_up(pop_size, 256), 256, 0, s>>>(S.sb.vals[cur], pop_size, order);
        EVOXB_LAUNCH_CHECK("take_prefix_kernel");
    }
    prof.mark("end", s);
    return 0;
}

EVOXB_API int evoxb_nsga2_select_f32(const float *fitness, int64_t n, int m, int64_t pop_size, int32_t *order, int32_t *rank_out,
                           float *crowd_out, void *ws, size_t ws_bytes, evoxb_stream_t stream) {
    return select_impl(0, fitness, n, m, pop_size, order, rank_out, crowd_out, ws, ws_bytes, stream);
}

EVOXB_API int evoxb_non_dominate_select_f32(const float *fitness, int64_t n, int m, int64_t topk, int32_t *order, void *ws,
                                  size_t ws_bytes, evoxb_stream_t stream) {
    return select_impl(1, fitness, n, m, topk, order, nullptr, nullptr, ws, ws_bytes, stream);
}

EVOXB_API int evoxb_comm_unique_id(void *id128) {
    if (!id128) return fail(EVOXB_E_INVALID, "id128 is null");
    NcclApi &nc = nccl_api();
    if (!nc.ok) return fail(EVOXB_E_NCCL, "libnccl.so.2 could not be loaded (set EVOXB_NCCL_LIB)");
    EVOXB_NCCL(nc.GetUniqueId(id128));
    return 0;
}

EVOXB_API int evoxb_comm_create(const void *id128, int rank, int nranks, evoxb_comm **out) {
    if (!id128 || !out) return fail(EVOXB_E_INVALID, "null pointer");
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(EVOXB_E_INVALID, "bad rank %d of %d", rank, nranks);
    NcclApi &nc = nccl_api();
    if (!nc.ok) return fail(EVOXB_E_NCCL, "libnccl.so.2 could not be loaded (set EVOXB_NCCL_LIB)");
    evoxb_comm *c = new evoxb_comm();
    c->rank = rank;
    c->nranks = nranks;
    NcclId id;
    memcpy(id.internal, id128, sizeof(id.internal));
    int r = nc.CommInitRank(&c->nccl, nranks, id, rank);
    if (r != 0) {
        delete c;
        return fail(EVOXB_E_NCCL, "ncclCommInitRank failed: %s", nc.GetErrorString(r));
    }
    cudaError_t e = cudaHostAlloc((void **)&c->h_flags, sizeof(int) * 64, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        nc.CommDestroy(c->nccl);
        delete c;
        return fail(EVOXB_E_CUDA, "cudaHostAlloc failed: %s", cudaGetErrorString(e));
    }
    c->h_err = reinterpret_cast<unsigned int *>(c->h_flags + 32);
    *c->h_err = 0u;
    setup_peer_exchange(c);  // best effort: c->p2p says whether the fused path is available
    *out = c;
    return 0;
}

EVOXB_API int evoxb_comm_status(evoxb_comm *comm, int *p2p_enabled, int *error, int *fronts) {
    if (!comm) return fail(EVOXB_E_INVALID, "comm is null");
    if (p2p_enabled) *p2p_enabled = comm->p2p ? 1 : 0;
    if (error) *error = 0;
    if (fronts) *fronts = 0;
    if (comm->p2p) {
        XchgHeader h;
        FlowHeader fh;
        EVOXB_CUDA(cudaMemcpy(&h, comm->xbuf, sizeof(h), cudaMemcpyDeviceToHost));
        EVOXB_CUDA(cudaMemcpy(&fh, comm->xbuf + comm->flow_off, sizeof(fh), cudaMemcpyDeviceToHost));
        if (error) *error = (h.error || fh.error || (comm->h_err && *comm->h_err)) ? 1 : 0;
        if (fronts) EVOXB_CUDA(cudaMemcpy(fronts, comm->fronts_dev, sizeof(int), cudaMemcpyDeviceToHost));
    }
    return 0;
}

EVOXB_API int evoxb_comm_poll_error(evoxb_comm *comm) {
    return (comm && comm->h_err && *reinterpret_cast<volatile unsigned int *>(comm->h_err)) ? 1 : 0;
}

EVOXB_API int evoxb_comm_set_timeout_ms(evoxb_comm *comm, int64_t ms) {
    if (!comm) return fail(EVOXB_E_INVALID, "comm is null");
    if (ms < 1) return fail(EVOXB_E_INVALID, "time-out must be at least 1 ms (got %lld)", (long long)ms);
    comm->timeout_ns = (unsigned long long)ms * 1000000ull;
    return 0;
}

EVOXB_API int evoxb_comm_last_path(evoxb_comm *comm) { return comm ? comm->last_path : -1; }

EVOXB_API int evoxb_comm_destroy(evoxb_comm *comm) {
    if (!comm) return 0;
    if (comm->h_flags) cudaFreeHost(comm->h_flags);
    for (int q = 0; q < comm->nranks && q < 16; ++q)
        if (q != comm->rank && comm->peer_ptr[q]) cudaIpcCloseMemHandle(comm->peer_ptr[q]);
    if (comm->peers_dev) cudaFree(comm->peers_dev);
    if (comm->fronts_dev) cudaFree(comm->fronts_dev);
    if (comm->xbuf) cudaFree(comm->xbuf);
    if (comm->nccl) nccl_api().CommDestroy(comm->nccl);
    delete comm;
    return 0;
}

// Column chunks (4096 columns each) owned by `rank`: first, stride, count -- pure host arithmetic.
EVOXB_API int evoxb_shard_plan(int64_t n, int rank, int nranks, int *first_chunk, int *chunk_stride, int *local_chunks,
                     int *total_chunks) {
    if (n < 0 || nranks < 1 || rank < 0 || rank >= nranks) return fail(EVOXB_E_INVALID, "bad arguments");
    NdsLayout L = nds_layout(nullptr, n, 1, 0, nranks, rank);
    if (first_chunk) *first_chunk = L.cm.first;
    if (chunk_stride) *chunk_stride = L.cm.stride;
    if (local_chunks) *local_chunks = L.cm.nlocal;
    if (total_chunks) *total_chunks = L.nchunks;
    return 0;
}

// Column ownership of the data-flow peel kernel (peel_flow.cuh) for one rank: pure host arithmetic, the same the
// kernel redoes on the device.  pieces (optional, 4 ints each: cta, chunk, first unit, units) lists every piece.
EVOXB_API int evoxb_flow_plan(int64_t n, int rank, int nranks, int tri, int max_ctas, int *grid, int *npieces, int *max_words,
                              int32_t *pieces, int pieces_cap) {
    if (n < 1 || nranks < 1 || rank < 0 || rank >= nranks || max_ctas < 1) return fail(EVOXB_E_INVALID, "bad arguments");
    NdsLayout L = nds_layout(nullptr, n, 1, 0, nranks, rank);
    if (L.nchunks > kFlowMaxChunks) return fail(EVOXB_E_UNSUPPORTED, "n = %lld exceeds the data-flow kernel's %d chunks", (long long)n, kFlowMaxChunks);
    const FlowPlan p = flow_make_plan(L.nchunks, L.cm, tri != 0, max_ctas);
    if (grid) *grid = p.grid;
    if (npieces) *npieces = p.npieces;
    if (max_words) *max_words = p.max_words;
    if (pieces && p.grid > 0) {
        int idx = 0;
        for (int lc = L.cm.nlocal - 1; lc >= 0; --lc) {
            const int nu = flow_piece_units(flow_unit_weight(L.cm.global(lc), L.nchunks, tri != 0), p.target);
            for (int i = 0; i < kUnitsPerChunk / nu; ++i, ++idx) {
                if (idx >= pieces_cap) return fail(EVOXB_E_INVALID, "pieces_cap too small (%d pieces)", p.npieces);
                pieces[4 * idx] = idx % p.grid;
                pieces[4 * idx + 1] = L.cm.global(lc);
                pieces[4 * idx + 2] = i * nu;
                pieces[4 * idx + 3] = nu;
            }
        }
    }
    return 0;
}

EVOXB_API int evoxb_nds_sharded_workspace_bytes(int64_t n, int m, int rank, int nranks, size_t *bytes) {
    if (!bytes) return fail(EVOXB_E_INVALID, "bytes is null");
    if (int rc = check_common((const void *)1, n, m)) return rc;
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(EVOXB_E_INVALID, "bad rank %d of %d", rank, nranks);
    *bytes = select_layout(nullptr, n, m, nranks, rank, true).bytes;
    return 0;
}

EVOXB_API int evoxb_nds_rank_sharded_f32(const float *fitness, int64_t n, int m, int32_t *rank, void *ws, size_t ws_bytes,
                               evoxb_comm *comm, evoxb_stream_t stream) {
    if (int rc = check_common(fitness, n, m)) return rc;
    if (!comm) return fail(EVOXB_E_INVALID, "comm is null");
    if (n == 0) return 0;
    if (!rank) return fail(EVOXB_E_INVALID, "rank is null");
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255)) return fail(EVOXB_E_WORKSPACE, "workspace is null or not 256-byte aligned");
    SelectLayout S = select_layout(ws, n, m, comm->nranks, comm->rank, true);
    if (ws_bytes < S.bytes) return fail(EVOXB_E_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, S.bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Profiler &prof = profiler();
    prof.begin();
    if (int rc = run_nds_sharded(fitness, n, m, rank, S.nds, comm, s, prof)) return rc;
    prof.mark("end", s);
    return 0;
}

EVOXB_API int evoxb_nsga2_select_sharded_f32(const float *fitness, int64_t n, int m, int64_t pop_size, int32_t *order,
                                   int32_t *rank_out, float *crowd_out, void *ws, size_t ws_bytes, evoxb_comm *comm,
                                   evoxb_stream_t stream) {
    if (!comm) return fail(EVOXB_E_INVALID, "comm is null");
    return select_impl(0, fitness, n, m, pop_size, order, rank_out, crowd_out, ws, ws_bytes, stream, comm);
}

EVOXB_API int evoxb_sbx_pm_clip_f32(const float *pop, int64_t n, int d, const float *lb, const float *ub, float pro_c,
                          float dis_c, float pro_m, float dis_m, uint32_t seed, uint32_t generation, float *out,
                          evoxb_stream_t stream) {
    if (n < 0 || d < 1) return fail(EVOXB_E_INVALID, "bad shape (%lld, %d)", (long long)n, d);
    if (n == 0) return 0;
    if (!pop || !lb || !ub || !out) return fail(EVOXB_E_INVALID, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int64_t work = (n / 2) * d + ((n & 1) ? d : 0);
    sbx_pm_clip_kernel<<<(unsigned)div_up(work > 0 ? work : 1, 256), 256, 0, s>>>(pop, n, d, lb, ub, pro_c, dis_c, pro_m,
                                                                                dis_m, seed, generation, out);
    EVOXB_LAUNCH_CHECK("sbx_pm_clip_kernel");
    return 0;
}

EVOXB_API int evoxb_dtlz2_f32(const float *x, int64_t n, int d, int m, float *f, evoxb_stream_t stream) {
    if (n < 0 || m < 1 || d < m) return fail(EVOXB_E_INVALID, "bad shape (n=%lld, d=%d, m=%d)", (long long)n, d, m);
    if (n == 0) return 0;
    if (!x || !f) return fail(EVOXB_E_INVALID, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    dtlz2_kernel<<<(unsigned)div_up(n, 128), 128, 0, s>>>(x, n, d, m, f);
    EVOXB_LAUNCH_CHECK("dtlz2_kernel");
    return 0;
}

EVOXB_API int evoxb_lsmop1_f32(const float *x, int64_t n, int d, int m, const int32_t *sublen, const int32_t *start, float *f,
                               evoxb_stream_t stream) {
    if (n < 0 || m < 2 || m > EVOXB_MAX_OBJECTIVES || d < m)
        return fail(EVOXB_E_INVALID, "bad shape (n=%lld, d=%d, m=%d; 2 <= m <= %d <= d)", (long long)n, d, m, EVOXB_MAX_OBJECTIVES);
    if (n == 0) return 0;
    if (!x || !f || !sublen || !start) return fail(EVOXB_E_INVALID, "null pointer");
    LsmopGroups gr{};
    for (int i = 0; i < m; ++i) {
        gr.sublen[i] = sublen[i];
        gr.start[i] = start[i];
        if (sublen[i] < 0 || start[i] < 0 || (int64_t)start[i] + m - 1 + (int64_t)kLsmopNk * sublen[i] > d)
            return fail(EVOXB_E_INVALID, "group %d (start %d, sublen %d) does not fit d = %d", i, start[i], sublen[i], d);
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    lsmop1_kernel<<<(unsigned)div_up(n * 32, 256), 256, 0, s>>>(x, n, d, m, gr, f);
    EVOXB_LAUNCH_CHECK("lsmop1_kernel");
    return 0;
}

EVOXB_API int evoxb_kth_rank_workspace_bytes(int64_t n, size_t *bytes) {
    if (!bytes) return fail(EVOXB_E_INVALID, "bytes is null");
    if (n < 0) return fail(EVOXB_E_INVALID, "n must be >= 0");
    Carver c(nullptr);
    c.take<int>((size_t)(n > 0 ? n : 1));
    *bytes = c.total();
    return 0;
}

EVOXB_API int evoxb_kth_rank_i32(const int32_t *rank, int64_t n, int64_t k, int32_t *out, void *ws, size_t ws_bytes,
                       evoxb_stream_t stream) {
    if (n <= 0 || k < 0 || k >= n) return fail(EVOXB_E_INVALID, "k must be in [0, n) (k = %lld, n = %lld)", (long long)k, (long long)n);
    if (!rank || !out) return fail(EVOXB_E_INVALID, "null pointer");
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255) || ws_bytes < sizeof(int) * (size_t)n)
        return fail(EVOXB_E_WORKSPACE, "workspace missing, misaligned or too small");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int *hist = static_cast<int *>(ws);
    EVOXB_CUDA(cudaMemsetAsync(hist, 0, sizeof(int) * (size_t)n, s));
    rank_hist_kernel<<<(unsigned)div_up(n, 256), 256, 0, s>>>(rank, n, hist);
    EVOXB_LAUNCH_CHECK("rank_hist_kernel");
    kth_rank_kernel<<<1, 1024, 0, s>>>(hist, n, k, out);
    EVOXB_LAUNCH_CHECK("kth_rank_kernel");
    return 0;
}

EVOXB_API int evoxb_argsort_workspace_bytes(int64_t n, size_t *bytes) {
    if (!bytes) return fail(EVOXB_E_INVALID, "bytes is null");
    if (n < 0) return fail(EVOXB_E_INVALID, "n must be >= 0");
    *bytes = argsort_layout(nullptr, n).bytes;
    return 0;
}

EVOXB_API int evoxb_argsort_f32(const float *x, int64_t n, int32_t *idx, void *ws, size_t ws_bytes, evoxb_stream_t stream) {
    if (n < 0) return fail(EVOXB_E_INVALID, "n must be >= 0");
    if (n == 0) return 0;
    if (!x || !idx) return fail(EVOXB_E_INVALID, "null pointer");
    if (n > (int64_t)1 << 24) return fail(EVOXB_E_UNSUPPORTED, "n exceeds 2^24");
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255)) return fail(EVOXB_E_WORKSPACE, "workspace is null or not 256-byte aligned");
    ArgsortLayout A = argsort_layout(ws, n);
    if (ws_bytes < A.bytes) return fail(EVOXB_E_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, A.bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const unsigned gb = (unsigned)div_up(n, 256);
    crowd_keys_kernel<<<gb, 256, 0, s>>>(x, n, A.sb.keys[0], A.sb.vals[0]);
    EVOXB_LAUNCH_CHECK("crowd_keys_kernel");
    int cur = 0;
    if (int rc = radix_sort_pairs(A.sb, 0, 1, n, nullptr, 0, 32, s, &cur)) return rc;
    take_prefix_kernel<<<gb, 256, 0, s>>>(A.sb.vals[cur], n, idx);
    EVOXB_LAUNCH_CHECK("take_prefix_kernel");
    return 0;
}

EVOXB_API int evoxb_gather_rows_f32(const float *src, int64_t n, int64_t d, const int32_t *idx, int64_t k, float *dst,
                          evoxb_stream_t stream) {
    if (n < 0 || d < 0 || k < 0) return fail(EVOXB_E_INVALID, "negative size");
    if (k == 0 || d == 0) return 0;
    if (!src || !idx || !dst) return fail(EVOXB_E_INVALID, "null pointer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    gather_rows_kernel<<<(unsigned)div_up(k * d, 256), 256, 0, s>>>(src, d, idx, k, dst);
    EVOXB_LAUNCH_CHECK("gather_rows_kernel");
    return 0;
}

}  // extern "C"

// ---- NSGA-III environmental selection (nsga3.py:93-211) ---------------------------------------------------
namespace evoxb {
struct Nsga3Full {
    Nsga3Layout n3;
    NdsLayout nds;
    int *rank;
    size_t bytes;
};
static Nsga3Full nsga3_full_layout(void *ws, int64_t n, int m, int64_t r) {
    Nsga3Full F{};
    Carver c(ws);
    F.rank = c.take<int>((size_t)(n > 0 ? n : 1));
    {
        char *sub = c.take<char>(0);
        F.n3 = nsga3_layout(sub, n, m, r);
        c.take<char>(F.n3.bytes);
    }
    {
        char *sub = c.take<char>(0);
        F.nds = nds_layout(sub, n, m, n);
        c.take<char>(F.nds.bytes);
    }
    F.bytes = c.total();
    return F;
}
}  // namespace evoxb

extern "C" {

EVOXB_API int evoxb_nsga3_select_workspace_bytes(int64_t n, int m, int64_t r, size_t *bytes) {
    if (!bytes) return fail(EVOXB_E_INVALID, "bytes is null");
    if (int rc = check_common((const void *)1, n, m)) return rc;
    if (r < 1) return fail(EVOXB_E_INVALID, "at least one reference point is needed (got %lld)", (long long)r);
    *bytes = nsga3_full_layout(nullptr, n, m, r).bytes;
    return 0;
}

EVOXB_API int evoxb_nsga3_select_f32(const float *fitness, int64_t n, int m, int64_t pop_size, const float *ref, int64_t r,
                                     const uint32_t key[2], int32_t *selected_idx, uint32_t next_key[2], int32_t *rank_out,
                                     int32_t *group_id_out, float *group_dist_out, float *norm_out, void *ws, size_t ws_bytes,
                                     evoxb_stream_t stream) {
    if (int rc = check_common(fitness, n, m)) return rc;
    if (!nds_use_tri(m)) return fail(EVOXB_E_UNSUPPORTED, "NSGA-III selection supports m <= 16 (got %d)", m);
    if (r < 1 || !ref) return fail(EVOXB_E_INVALID, "reference points missing (r = %lld)", (long long)r);
    if (pop_size < 1 || pop_size >= n)
        return fail(EVOXB_E_INVALID, "pop_size must be in [1, n) (got %lld, n = %lld)", (long long)pop_size, (long long)n);
    if (!key || !selected_idx) return fail(EVOXB_E_INVALID, "null pointer");
    if (!ws || (reinterpret_cast<uintptr_t>(ws) & 255)) return fail(EVOXB_E_WORKSPACE, "workspace is null or not 256-byte aligned");
    Nsga3Full F = nsga3_full_layout(ws, n, m, r);
    if (ws_bytes < F.bytes) return fail(EVOXB_E_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, F.bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    Profiler &prof = profiler();
    prof.begin();
    int *rank = rank_out ? rank_out : F.rank;
    if (int rc = run_nds(fitness, n, m, rank, F.nds, s, prof)) return rc;

    prof.mark("nsga3_normalise", s);
    const Nsga3Layout &L = F.n3;
    const int nn = (int)n, rr = (int)r;
    const unsigned gb = (unsigned)div_up(n, 256);
    EVOXB_CUDA(cudaMemsetAsync(L.hist, 0, sizeof(int) * (size_t)n, s));
    EVOXB_CUDA(cudaMemsetAsync(L.info, 0, sizeof(int) * 64, s));
    EVOXB_CUDA(cudaMemsetAsync(L.ideal_key, 0xff, sizeof(uint32_t) * (size_t)m, s));
    EVOXB_CUDA(cudaMemsetAsync(L.worst_key, 0, sizeof(uint32_t) * (size_t)m, s));
    EVOXB_CUDA(cudaMemsetAsync(L.ext_key, 0xff, sizeof(unsigned long long) * (size_t)m, s));
    EVOXB_CUDA(cudaMemsetAsync(L.rho, 0, sizeof(int) * (size_t)r, s));
    EVOXB_CUDA(cudaMemsetAsync(L.rho_last, 0, sizeof(int) * (size_t)r, s));
    EVOXB_CUDA(cudaMemsetAsync(L.cursor, 0, sizeof(int) * (size_t)r, s));
    rank_hist_kernel<<<gb, 256, 0, s>>>(rank, n, L.hist);
    EVOXB_LAUNCH_CHECK("rank_hist_kernel");
    kth_rank_kernel<<<1, 1024, 0, s>>>(L.hist, n, pop_size, L.info);  // last_rank = rank[argsort(rank)[pop_size]], nsga3.py:97-98
    EVOXB_LAUNCH_CHECK("kth_rank_kernel");
    n3_ideal_kernel<<<gb, 256, 0, s>>>(fitness, rank, nn, m, L.info, L.ideal_key);
    EVOXB_LAUNCH_CHECK("n3_ideal_kernel");
    n3_extreme_kernel<<<gb, 256, 0, s>>>(fitness, rank, nn, m, L.info, L.ideal_key, L.ext_key, L.worst_key);
    EVOXB_LAUNCH_CHECK("n3_extreme_kernel");
    n3_nadir_kernel<<<1, 32, 0, s>>>(fitness, m, L.ideal_key, L.ext_key, L.worst_key, L.norm, L.info);
    EVOXB_LAUNCH_CHECK("n3_nadir_kernel");
    int *gid = group_id_out ? group_id_out : L.gid;
    float *gdist = group_dist_out ? group_dist_out : L.gdist;
    n3_unit_rows_kernel<<<gb, 256, 0, s>>>(fitness, rank, nn, m, L.info, L.norm, L.xu, L.xnorm, L.kept_list, gid, gdist, rr);
    EVOXB_LAUNCH_CHECK("n3_unit_rows_kernel");
    n3_unit_refs_kernel<<<(unsigned)div_up(r, 256), 256, 0, s>>>(ref, rr, m, L.yu);
    EVOXB_LAUNCH_CHECK("n3_unit_refs_kernel");

    prof.mark("nsga3_associate", s);
    const unsigned ga = (unsigned)div_up(n, kN3RowsPerCta);
#define EVOXB_N3_CASE(MM)                                                                          \
    case MM:                                                                                       \
        n3_associate_kernel<MM><<<ga, 256, 0, s>>>(L.xu, L.xnorm, L.kept_list, L.info, L.yu, rr, gid, gdist); \
        break;
    switch (m) {
        EVOXB_N3_CASE(1) EVOXB_N3_CASE(2) EVOXB_N3_CASE(3) EVOXB_N3_CASE(4) EVOXB_N3_CASE(5) EVOXB_N3_CASE(6)
        EVOXB_N3_CASE(7) EVOXB_N3_CASE(8) EVOXB_N3_CASE(9) EVOXB_N3_CASE(10) EVOXB_N3_CASE(11) EVOXB_N3_CASE(12)
        EVOXB_N3_CASE(13) EVOXB_N3_CASE(14) EVOXB_N3_CASE(15) EVOXB_N3_CASE(16)
    }
#undef EVOXB_N3_CASE
    EVOXB_LAUNCH_CHECK("n3_associate_kernel");

    prof.mark("nsga3_niche", s);
    n3_count_kernel<<<gb, 256, 0, s>>>(rank, gid, nn, rr, L.info, L.rho, L.rho_last, L.keep);
    EVOXB_LAUNCH_CHECK("n3_count_kernel");
    n3_offsets_kernel<<<1, 1024, 0, s>>>(L.rho_last, rr, L.off);
    EVOXB_LAUNCH_CHECK("n3_offsets_kernel");
    n3_fill_kernel<<<gb, 256, 0, s>>>(rank, gid, nn, rr, L.info, L.off, L.cursor, L.members);
    EVOXB_LAUNCH_CHECK("n3_fill_kernel");
    NicheArgs na{};
    na.n = nn; na.r = rr; na.pop = (int)pop_size;
    na.key0 = key[0]; na.key1 = key[1];
    na.rho = L.rho; na.rho_last = L.rho_last; na.off = L.off; na.members = L.members;
    na.taken = L.taken; na.keep = L.keep; na.gdist = gdist; na.info = L.info; na.selected_idx = selected_idx;
    n3_niche_kernel<<<1, 1024, 0, s>>>(na);
    EVOXB_LAUNCH_CHECK("n3_niche_kernel");
    if (norm_out) EVOXB_CUDA(cudaMemcpyAsync(norm_out, L.norm, sizeof(float) * 3 * (size_t)m, cudaMemcpyDeviceToDevice, s));
    if (next_key) {  // keys[0] of jax.random.split(key, pop_size + 1), nsga3.py:154, :209
        next_key[0] = jax_split_word(key[0], key[1], (uint32_t)pop_size + 1u, 0u);
        next_key[1] = jax_split_word(key[0], key[1], (uint32_t)pop_size + 1u, 1u);
    }
    prof.mark("end", s);
    return 0;
}

}  // extern "C"

// ---- crowding distance for callers that cannot allocate device memory -------------------------
// Scratch for evoxb_crowding_noalloc_f32: part of the library image, so every GPU that loads
// libevoxb.so holds these 8 MiB until the process exits.  In host code the name below is only a
// host-side handle of the symbol, not a device address; the address is looked up once per
// device with cudaGetSymbolAddress (noalloc_scratch_for_current_device) and that pointer is what
// the kernels receive.
#define EVOXB_NOALLOC_SCRATCH_BYTES EVOXB_CROWDING_NOALLOC_MAX_BYTES
static __device__ __align__(256) unsigned char g_noalloc_scratch[EVOXB_NOALLOC_SCRATCH_BYTES];

namespace evoxb {

constexpr int kMaxDevices = 64;
static std::mutex g_noalloc_mu;
static void *g_noalloc_ptr[kMaxDevices] = {};

// Device address of g_noalloc_scratch on the CURRENT device.  The first call on a device asks the
// runtime for the symbol's address (which may load the module on that device); later calls only
// read the cached pointer.
static int noalloc_scratch_for_current_device(void **out) {
    int dev = -1;
    EVOXB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(EVOXB_E_UNSUPPORTED, "device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return fail(EVOXB_E_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) failed on device %d: %s", dev,
                        cudaGetErrorString(e));
        g_noalloc_ptr[dev] = ptr;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

}  // namespace evoxb

extern "C" {

// Same result as evoxb_crowding_f32, with the sort buffers taken from the library's own static
// device array instead of a caller workspace.  See include/evoxb.h for the contract.
EVOXB_API int evoxb_crowding_noalloc_f32(const float *fitness, const uint8_t *mask, int64_t n, int m, float *dist,
                                         evoxb_stream_t stream) {
    if (int rc = check_common(fitness, n, m)) return rc;
    if (n == 0) return 0;
    if (!dist) return fail(EVOXB_E_INVALID, "dist is null");
    const size_t need = crowd_layout(nullptr, n, m).bytes;
    if (need > EVOXB_NOALLOC_SCRATCH_BYTES)
        return fail(EVOXB_E_WORKSPACE,
                    "evoxb_crowding_noalloc_f32: n = %lld, m = %d needs %zu bytes of scratch but the built-in array "
                    "holds %u; use evoxb_crowding_f32 with a caller workspace for this size",
                    (long long)n, m, need, (unsigned)EVOXB_NOALLOC_SCRATCH_BYTES);
    void *scratch = nullptr;
    if (int rc = noalloc_scratch_for_current_device(&scratch)) return rc;
    // Nothing in the array has to start at zero: every word a call reads (segment length, tile
    // counts, keys, indices, histograms, per-objective contributions) is written earlier in the
    // same call, so the previous call's leftovers are never observed.
    CrowdLayout L = crowd_layout(scratch, n, m);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    return launch_crowding(fitness, mask, n, m, dist, L, s);
}

}  // extern "C"
