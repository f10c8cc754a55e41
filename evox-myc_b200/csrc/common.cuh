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
