// Attention softmax forward/backward (LayerNorm lives in layernorm.cu).
// One CTA per row with the row cached in registers (16-byte accesses), warp-shuffle reductions; LayerNorm
// backward keeps per-thread dgamma/dbeta accumulators across a grid-stride loop over rows and emits per-CTA
// partials that a second tiny kernel folds, so gamma/beta gradients cost no atomics and no second pass over dy.
// Reference behaviour: csrc/transformer/normalize_kernels.cu:16-2103, csrc/transformer/softmax_kernels.cu:8-577.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"

namespace dsb {

// ------------------------------------------------------------------------------------------- softmax
constexpr int SM_THREADS = 256;
constexpr int SM_MAX_ITERS = 16;  // k_len up to 256 * 8 * 16 = 32768

template <typename T, int ITERS>
__global__ void __launch_bounds__(SM_THREADS) softmax_fwd_kernel(const T* __restrict__ x, const T* __restrict__ mask,
                                                                 T* __restrict__ y, int heads, int q_len, int k_len,
                                                                 float scale, int causal) {
    __shared__ float sh[33];
    const long row = blockIdx.x;
    const int q = row % q_len;
    const int b = row / ((long)q_len * heads);
    const int nvec = k_len >> 3;
    const int limit = causal ? q + (k_len - q_len) : k_len - 1;  // last visible key index
    const uint4* xr = reinterpret_cast<const uint4*>(x + row * k_len);
    const uint4* mr = mask ? reinterpret_cast<const uint4*>(mask + (long)b * k_len) : nullptr;
    float vals[ITERS][8];
    float mx = -INFINITY;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * SM_THREADS;
        if (i < nvec) {
            unpack8<T>(xr[i], vals[it]);
            float m8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (mr) unpack8<T>(__ldg(mr + i), m8);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float v = vals[it][k] * scale + m8[k];
                if (i * 8 + k > limit) v = -INFINITY;
                vals[it][k] = v;
                mx = fmaxf(mx, v);
            }
        }
    }
    mx = block_max(mx, sh);
    if (mx == -INFINITY) mx = 0.f;
    float sum = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * SM_THREADS;
        if (i < nvec) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                vals[it][k] = __expf(vals[it][k] - mx);
                sum += vals[it][k];
            }
        }
    }
    sum = block_sum(sum, sh);
    const float inv = 1.f / (sum + 1e-20f);
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * SM_THREADS;
        if (i < nvec) {
#pragma unroll
            for (int k = 0; k < 8; ++k) vals[it][k] *= inv;
            reinterpret_cast<uint4*>(y + row * k_len)[i] = pack8<T>(vals[it]);
        }
    }
}

template <typename T, int ITERS>
__global__ void __launch_bounds__(SM_THREADS) softmax_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                 T* __restrict__ dx, int k_len, float scale) {
    __shared__ float sh[33];
    const long row = blockIdx.x;
    const int nvec = k_len >> 3;
    float yv[ITERS][8], gv[ITERS][8];
    float dot = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * SM_THREADS;
        if (i < nvec) {
            unpack8<T>(reinterpret_cast<const uint4*>(y + row * k_len)[i], yv[it]);
            unpack8<T>(reinterpret_cast<const uint4*>(dy + row * k_len)[i], gv[it]);
#pragma unroll
            for (int k = 0; k < 8; ++k) dot += yv[it][k] * gv[it][k];
        }
    }
    dot = block_sum(dot, sh);
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * SM_THREADS;
        if (i < nvec) {
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = yv[it][k] * (gv[it][k] - dot) * scale;
            reinterpret_cast<uint4*>(dx + row * k_len)[i] = pack8<T>(o);
        }
    }
}

template <typename T>
static int softmax_fwd_launch(const void* x, const void* mask, void* y, long rows, int heads, int q_len, int k_len,
                              float scale, int causal, cudaStream_t s) {
    const int iters = ((k_len >> 3) + SM_THREADS - 1) / SM_THREADS;
    const T* xp = reinterpret_cast<const T*>(x);
    const T* mp = reinterpret_cast<const T*>(mask);
    T* yp = reinterpret_cast<T*>(y);
#define SMF(I) softmax_fwd_kernel<T, I><<<(unsigned)rows, SM_THREADS, 0, s>>>(xp, mp, yp, heads, q_len, k_len, scale, causal)
    if (iters <= 1) SMF(1); else if (iters <= 2) SMF(2); else if (iters <= 4) SMF(4); else if (iters <= 8) SMF(8); else SMF(16);
#undef SMF
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}
template <typename T>
static int softmax_bwd_launch(const void* dy, const void* y, void* dx, long rows, int k_len, float scale,
                              cudaStream_t s) {
    const int iters = ((k_len >> 3) + SM_THREADS - 1) / SM_THREADS;
    const T* gp = reinterpret_cast<const T*>(dy);
    const T* yp = reinterpret_cast<const T*>(y);
    T* dp = reinterpret_cast<T*>(dx);
#define SMB(I) softmax_bwd_kernel<T, I><<<(unsigned)rows, SM_THREADS, 0, s>>>(gp, yp, dp, k_len, scale)
    if (iters <= 1) SMB(1); else if (iters <= 2) SMB(2); else if (iters <= 4) SMB(4); else if (iters <= 8) SMB(8); else SMB(16);
#undef SMB
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace dsb

using namespace dsb;

extern "C" int dsb_softmax_fwd(const void* x, const void* mask, void* y, int batch, int heads, int q_len, int k_len,
                               float scale, int causal, int dtype, void* stream) {
    const long rows = (long)batch * heads * q_len;
    if (rows <= 0) return 0;
    if (k_len % 8 || k_len > SM_THREADS * 8 * SM_MAX_ITERS) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == DSB_DTYPE_F16) return softmax_fwd_launch<__half>(x, mask, y, rows, heads, q_len, k_len, scale, causal, s);
    if (dtype == DSB_DTYPE_BF16) return softmax_fwd_launch<__nv_bfloat16>(x, mask, y, rows, heads, q_len, k_len, scale, causal, s);
    return DSB_ERR_DTYPE;
}

extern "C" int dsb_softmax_bwd(const void* dy, const void* y, void* dx, long rows, int k_len, float scale, int dtype,
                               void* stream) {
    if (rows <= 0) return 0;
    if (k_len % 8 || k_len > SM_THREADS * 8 * SM_MAX_ITERS) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == DSB_DTYPE_F16) return softmax_bwd_launch<__half>(dy, y, dx, rows, k_len, scale, s);
    if (dtype == DSB_DTYPE_BF16) return softmax_bwd_launch<__nv_bfloat16>(dy, y, dx, rows, k_len, scale, s);
    return DSB_ERR_DTYPE;
}
