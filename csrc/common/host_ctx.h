// Host-side helper for entry points that call the CUDA *driver* API (cuTensorMapEncodeTiled): the driver needs a
// context current on the calling thread. PyTorch's autograd worker thread for device 0 may not have one yet (torch skips
// cudaSetDevice when the target equals the thread's default device 0), so the first GEMM of a backward pass failed with
// CUDA_ERROR_INVALID_CONTEXT there. Any context-requiring runtime call binds the primary context; do it once per thread.
#pragma once
#include <cuda_runtime.h>

namespace dsb {
inline void ensure_context_on_this_thread() {
    static thread_local bool bound = false;
    if (!bound) {
        cudaFree(nullptr);
        bound = true;
    }
}
}  // namespace dsb
