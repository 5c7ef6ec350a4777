// flatten / unflatten of dense tensors on the host side of the C ABI (reference csrc/utils/flatten_unflatten.cpp:11-25
// exposes torch::utils::flatten_dense_tensors). The engine itself never needs them — parameters and gradients live in
// flat arenas from the start — but client code using `deepspeed.ops.op_builder.UtilsBuilder().load()` gets the same
// two functions; these raw-pointer versions back the CPU path.
#include <omp.h>
#include <stdint.h>
#include <string.h>

extern "C" {

// copies n tensors (ptrs[i], bytes[i]) back to back into dst
int dsb_flatten_bytes(const void* const* ptrs, const long* bytes, int n, void* dst) {
    long off = 0;
    for (int i = 0; i < n; ++i) {
        memcpy(static_cast<char*>(dst) + off, ptrs[i], bytes[i]);
        off += bytes[i];
    }
    return 0;
}

int dsb_unflatten_bytes(const void* src, void* const* ptrs, const long* bytes, int n) {
    long off = 0;
    for (int i = 0; i < n; ++i) {
        memcpy(ptrs[i], static_cast<const char*>(src) + off, bytes[i]);
        off += bytes[i];
    }
    return 0;
}

}  // extern "C"
