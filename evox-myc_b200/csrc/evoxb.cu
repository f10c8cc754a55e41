// C-ABI entry points of libevoxb.so (see include/evoxb.h).  Hand-written sm_100a CUDA only:
// no CPU fallback, no library kernels on the compute path.
#include "common.cuh"
#include "dominance.cuh"
#include "peel.cuh"

namespace evoxb {

bool g_profile_enabled = false;

struct NdsLayout {
    int64_t n, npad, row_words;
    int m, nchunks;
    float *xt;
    uint32_t *D;
    int *count;
    int *list;
    PeelCtrl *ctrl;
    size_t bytes;
};

static NdsLayout nds_layout(void *ws, int64_t n, int m, int64_t n_rows) {
    NdsLayout L{};
    L.n = n;
    L.m = m;
    L.npad = round_up(n > 0 ? n : 1, kColPad);
    L.row_words = L.npad / 32;
    L.nchunks = (int)(L.npad / kChunkCols);
    Carver c(ws);
    L.ctrl = c.take<PeelCtrl>(1);
    L.xt = c.take<float>((size_t)L.npad * (size_t)m);
    L.count = c.take<int>((size_t)L.npad);
    L.list = c.take<int>((size_t)L.npad);
    L.D = c.take<uint32_t>((size_t)n_rows * (size_t)L.row_words);
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

static int peel_grid(int *blocks) {
    static thread_local int cached_dev = -1, cached_blocks = 0;
    int dev = 0;
    EVOXB_CUDA(cudaGetDevice(&dev));
    if (dev != cached_dev) {
        int sms = 0, per_sm = 0;
        EVOXB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        EVOXB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, peel_persistent_kernel, kPeelThreads, 0));
        if (per_sm < 1) return fail(EVOXB_E_CUDA, "peel kernel does not fit on an SM");
        if (per_sm > 3) per_sm = 3;
        cached_blocks = sms * per_sm;
        cached_dev = dev;
    }
    *blocks = cached_blocks;
    return 0;
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

    prof.mark("prep", s);
    EVOXB_CUDA(cudaMemsetAsync(L.ctrl, 0, sizeof(PeelCtrl), s));
    EVOXB_CUDA(cudaMemsetAsync(L.count, 0, sizeof(int) * (size_t)L.npad, s));
    transpose_pad_kernel<<<(unsigned)div_up(L.npad, 256), 256, 0, s>>>(fitness, n, m, L.npad, L.xt);
    EVOXB_LAUNCH_CHECK("transpose_pad_kernel");

    prof.mark("dominance_build", s);
    if (int rc = launch_dominance_build(L.xt, L.npad, 0, n, m, L.D, L.row_words, s)) return rc;

    int blocks = 0;
    if (int rc = peel_grid(&blocks)) return rc;

    prof.mark("dominator_count", s);
    dominator_count_kernel<<<blocks, kPeelThreads, 0, s>>>(L.D, L.row_words, L.nchunks, n, L.count);
    EVOXB_LAUNCH_CHECK("dominator_count_kernel");

    prof.mark("peel", s);
    {
        const uint32_t *D = L.D;
        int64_t row_words = L.row_words;
        int nchunks = L.nchunks;
        int nn = (int)n;
        int *count = L.count;
        int *rk = rank;
        const int *perm = nullptr;
        int *list = L.list;
        PeelCtrl *ctrl = L.ctrl;
        void *args[] = {&D, &row_words, &nchunks, &nn, &count, &rk, &perm, &list, &ctrl};
        EVOXB_CUDA(cudaLaunchCooperativeKernel((const void *)peel_persistent_kernel, dim3(blocks), dim3(kPeelThreads),
                                               args, 0, s));
    }
    prof.mark("end", s);
    return 0;
}

}  // extern "C"
