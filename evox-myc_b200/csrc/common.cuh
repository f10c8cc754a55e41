// Shared helpers for the evoxb kernels (sm_100a only).
#pragma once

#include <atomic>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/evoxb.h"

namespace evoxb {

// ---- thread-local error message -----------------------------------------------------------
inline char *err_buf() {
    static thread_local char buf[512] = {0};
    return buf;
}
inline int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}

#define EVOXB_CUDA(call)                                                                         \
    do {                                                                                         \
        cudaError_t e__ = (call);                                                                \
        if (e__ != cudaSuccess)                                                                  \
            return ::evoxb::fail(EVOXB_E_CUDA, "%s failed at %s:%d: %s", #call, __FILE__,        \
                                 __LINE__, cudaGetErrorString(e__));                             \
    } while (0)

#define EVOXB_LAUNCH_CHECK(name)                                                                 \
    do {                                                                                         \
        ++::evoxb::g_launches;                                                                   \
        cudaError_t e__ = cudaGetLastError();                                                    \
        if (e__ != cudaSuccess)                                                                  \
            return ::evoxb::fail(EVOXB_E_CUDA, "launch of %s failed at %s:%d: %s", name,         \
                                 __FILE__, __LINE__, cudaGetErrorString(e__));                   \
    } while (0)

// ---- sizes ---------------------------------------------------------------------------------
constexpr int kColPad = 4096;  // dominance rows are padded to a multiple of 4096 columns (512 B)

__host__ __device__ inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
__host__ __device__ inline int64_t div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Which column chunks this GPU owns: global chunk of local chunk lc is first + lc * stride.
// Single GPU: {0, 1, nchunks}.  Row-sharded^T multi-GPU: rank p of P owns chunks p, p+P, ... and
// stores, for every row, only the words of its own chunks (row stride = nlocal * 128 words).
struct ChunkMap {
    int first, stride, nlocal;
    __host__ __device__ int global(int lc) const { return first + lc * stride; }
};

// Where the packed row of individual i starts.  Two layouts of the bit matrix:
//   rectangular (unsorted full matrix, m > 16): row i = row_vec uint4 starting at i * row_vec;
//   triangular  (presorted population): row i only has bits in the column chunks at or after its own chunk
//     a = i / 4096, and only those LOCAL chunks are stored (this GPU owns chunks first, first + stride, ...):
//     nl(a) * 32 uint4 per row of chunk a.  vstart() is the row's "virtual start" -- where it would begin if the
//     local chunks in front of its own were stored too -- so that the uint4 of local chunk lc sits at
//     vstart(i) + lc * 32 in both layouts.  Halves the matrix (N^2/16 bytes on one GPU).
// Pure arithmetic (no table to load before a tile can be addressed): with before(a) = number of local chunks in
// front of chunk a, the rows of chunks < a hold 4096 * 32 * (a * nlocal - sum_{a' < a} before(a')) uint4.
struct TriRows {
    int64_t row_vec;  // rectangular layout: uint4 per row
    int64_t row_words;  // the same in 32-bit words
    int tri;          // 1 = triangular storage
    int first, stride, nlocal;
    __host__ __device__ __forceinline__ int before(int a) const { return a <= first ? 0 : (a - first - 1) / stride + 1; }
    __host__ __device__ __forceinline__ int row_uint4(int a) const { return tri ? (nlocal - before(a)) * 32 : (int)row_vec; }
    // uint4 stored in front of the rows of chunk a
    __host__ __device__ __forceinline__ int64_t chunk_base(int a) const {
        const int64_t T = a - 1 - first > 0 ? a - 1 - first : 0;  // chunks first + 1 .. a - 1 have a local chunk in front
        const int64_t Q = T / stride, R = T - Q * stride;
        const int64_t g = T + stride * (Q * (Q - 1) / 2) + Q * R;  // sum of before(a') over a' < a
        return 4096ll * 32ll * ((int64_t)a * nlocal - g);
    }
    __host__ __device__ __forceinline__ int64_t vstart(int64_t i) const {
        if (!tri) return i * row_vec;
        const int a = (int)(i >> 12);
        const int bf = before(a);
        return chunk_base(a) + (i & 4095) * (int64_t)((nlocal - bf) * 32) - (int64_t)bf * 32;
    }
};

// Bump allocator over the caller's workspace; every block is 256-byte aligned.
struct Carver {
    char *base;
    size_t off = 0;
    explicit Carver(void *p) : base(static_cast<char *>(p)) {}
    template <typename T>
    T *take(size_t count) {
        off = (off + 255) & ~size_t(255);
        T *p = reinterpret_cast<T *>(base ? base + off : nullptr);
        off += count * sizeof(T);
        return p;
    }
    size_t total() const { return (off + 255) & ~size_t(255); }
};

// ---- stage profiler (events on the caller's stream) -------------------------------------------
struct Profiler {
    static constexpr int kMax = 24;
    bool enabled = false;
    int n = 0;  // number of marks (stages = n - 1)
    cudaEvent_t ev[kMax] = {};
    const char *name[kMax] = {};
    bool created = false;
    void begin() { n = 0; }
    void mark(const char *stage_name, cudaStream_t s) {
        if (!enabled || n >= kMax) return;
        if (!created) {
            for (int i = 0; i < kMax; ++i) cudaEventCreate(&ev[i]);
            created = true;
        }
        cudaEventRecord(ev[n], s);
        name[n] = stage_name;  // name of the stage that STARTS at this mark
        ++n;
    }
};
inline Profiler &profiler() {
    static thread_local Profiler p;
    return p;
}
extern std::atomic<unsigned long long> g_launches;  // kernels launched by this library since load

// ---- small device helpers ---------------------------------------------------------------------
__device__ __forceinline__ uint32_t lop3_xor3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t lop3_maj(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
// streaming 128-bit load that does not allocate in L1 (data is read exactly once)
__device__ __forceinline__ uint4 ldg_stream_v4(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// the same load, predicated in the instruction itself (no branch: a batch of such loads stays in
// flight together); returns zeros when `fetch` is false
__device__ __forceinline__ uint4 ldg_stream_v4_if(const uint4 *p, bool fetch) {
    uint4 r;
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "setp.ne.u32 q, %5, 0;\n\t"
        "mov.u32 %0, 0; mov.u32 %1, 0; mov.u32 %2, 0; mov.u32 %3, 0;\n\t"
        "@q ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n\t"
        "}"
        : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
        : "l"(p), "r"((unsigned)fetch));
    return r;
}

}  // namespace evoxb
