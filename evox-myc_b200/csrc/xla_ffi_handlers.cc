// XLA FFI handlers: the thin custom-call layer between jax.ffi.ffi_call and the C ABI of libevoxb.so.
//
// NOT BUILT ON THIS IMAGE: jaxlib (which ships xla/ffi/api/ffi.h) is not installed and cannot be
// fetched, so this file is compiled only where jax is importable (evox-myc_b200/_build.py:build_ffi)
// and has never been exercised -- see DESIGN.md section 1 and INTEGRATION.md section 2.  It forwards
// XLA's operand/result buffers, XLA's scratch allocator and XLA's stream unchanged to the entry
// points declared in include/evoxb.h; there is no logic of its own.
//
// Replaces the bodies of (reference tree):
//   non_dominated_sort  src/evox/operators/selection/non_dominate.py:30-113
//   crowding_distance   src/evox/operators/selection/non_dominate.py:116-156
#if __has_include("xla/ffi/api/ffi.h")
#include <cuda_runtime.h>

#include "../../include/evoxb.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

static ffi::Error NdsImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::F32> x,
                          ffi::ResultBuffer<ffi::S32> rank) {
    if (x.dimensions().size() != 2) return ffi::Error::InvalidArgument("evoxb_nds: x must be (n, m)");
    const int64_t n = x.dimensions()[0];
    const int m = static_cast<int>(x.dimensions()[1]);
    size_t bytes = 0;
    if (evoxb_nds_workspace_bytes(n, m, &bytes)) return ffi::Error::InvalidArgument(evoxb_last_error());
    if (n == 0) return ffi::Error::Success();
    auto ws = scratch.Allocate(bytes, 256);
    if (!ws) return ffi::Error::ResourceExhausted("evoxb_nds: XLA could not provide the dominance-matrix scratch");
    if (evoxb_nds_rank_f32(x.typed_data(), n, m, rank->typed_data(), *ws, bytes, stream))
        return ffi::Error::Internal(evoxb_last_error());
    return ffi::Error::Success();
}

static ffi::Error CrowdingImpl(cudaStream_t stream, ffi::ScratchAllocator scratch, ffi::Buffer<ffi::F32> costs,
                               ffi::Buffer<ffi::PRED> mask, ffi::ResultBuffer<ffi::F32> dist) {
    if (costs.dimensions().size() != 2) return ffi::Error::InvalidArgument("evoxb_crowding: costs must be (n, m)");
    const int64_t n = costs.dimensions()[0];
    const int m = static_cast<int>(costs.dimensions()[1]);
    size_t bytes = 0;
    if (evoxb_crowding_workspace_bytes(n, m, &bytes)) return ffi::Error::InvalidArgument(evoxb_last_error());
    if (n == 0) return ffi::Error::Success();
    auto ws = scratch.Allocate(bytes, 256);
    if (!ws) return ffi::Error::ResourceExhausted("evoxb_crowding: XLA could not provide the sort scratch");
    // JAX's bool is one byte per element, which is what the C ABI's mask expects
    if (evoxb_crowding_f32(costs.typed_data(), reinterpret_cast<const uint8_t *>(mask.untyped_data()), n, m,
                           dist->typed_data(), *ws, bytes, stream))
        return ffi::Error::Internal(evoxb_last_error());
    return ffi::Error::Success();
}

XLA_FFI_DEFINE_HANDLER_SYMBOL(EvoxbNds, NdsImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Ctx<ffi::ScratchAllocator>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::S32>>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(EvoxbCrowding, CrowdingImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Ctx<ffi::ScratchAllocator>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::PRED>>()
                                  .Ret<ffi::Buffer<ffi::F32>>());
#endif  // __has_include("xla/ffi/api/ffi.h")
