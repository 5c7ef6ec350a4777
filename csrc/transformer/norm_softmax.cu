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

// ---- short rows (k_len <= 2048): one row per group of G lanes (G = 8..32, a power of two), several rows per warp and 8
// warps per CTA; reductions are shuffles inside the group, no shared memory, no block barrier. The one-CTA-per-row
// kernels above leave 240 of 256 threads idle at k_len = 128 (BERT seq 128: 182 us per call instead of ~12).
template <int G>
DSB_DEVICE float group_max(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <int G>
DSB_DEVICE float group_sum(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <typename T, int ITERS, int G>
__global__ void __launch_bounds__(SM_THREADS) softmax_fwd_group_kernel(const T* __restrict__ x, const T* __restrict__ mask,
                                                                       T* __restrict__ y, long rows, int heads, int q_len,
                                                                       int k_len, float scale, int causal) {
    const int gl = threadIdx.x % G;                                   // lane inside the group
    const long row = (long)blockIdx.x * (SM_THREADS / G) + threadIdx.x / G;
    const bool live = row < rows;                                     // whole groups are live or dead together
    const long r = live ? row : 0;
    const int q = r % q_len;
    const int b = r / ((long)q_len * heads);
    const int nvec = k_len >> 3;
    const int limit = causal ? q + (k_len - q_len) : k_len - 1;
    const uint4* xr = reinterpret_cast<const uint4*>(x + r * k_len);
    const uint4* mr = mask ? reinterpret_cast<const uint4*>(mask + (long)b * k_len) : nullptr;
    float vals[ITERS][8];
    float mx = -INFINITY;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = gl + it * G;
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
    mx = group_max<G>(mx);
    if (mx == -INFINITY) mx = 0.f;
    float sum = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        if (gl + it * G < nvec) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                vals[it][k] = __expf(vals[it][k] - mx);
                sum += vals[it][k];
            }
        }
    }
    sum = group_sum<G>(sum);
    const float inv = 1.f / (sum + 1e-20f);
    if (!live) return;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = gl + it * G;
        if (i < nvec) {
#pragma unroll
            for (int k = 0; k < 8; ++k) vals[it][k] *= inv;
            reinterpret_cast<uint4*>(y + row * k_len)[i] = pack8<T>(vals[it]);
        }
    }
}

template <typename T, int ITERS, int G>
__global__ void __launch_bounds__(SM_THREADS) softmax_bwd_group_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                       T* __restrict__ dx, long rows, int k_len,
                                                                       float scale) {
    const int gl = threadIdx.x % G;
    const long row = (long)blockIdx.x * (SM_THREADS / G) + threadIdx.x / G;
    const bool live = row < rows;
    const long r = live ? row : 0;
    const int nvec = k_len >> 3;
    float yv[ITERS][8], gv[ITERS][8];
    float dot = 0.f;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = gl + it * G;
        if (i < nvec) {
            unpack8<T>(reinterpret_cast<const uint4*>(y + r * k_len)[i], yv[it]);
            unpack8<T>(reinterpret_cast<const uint4*>(dy + r * k_len)[i], gv[it]);
#pragma unroll
            for (int k = 0; k < 8; ++k) dot += yv[it][k] * gv[it][k];
        }
    }
    dot = group_sum<G>(dot);
    if (!live) return;
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = gl + it * G;
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
    const int nvec = k_len >> 3;
    if (nvec <= 64) {   // k_len <= 512: row per lane group (longer rows keep too many values per lane: 145 registers,
                        // 11 % occupancy at k_len 2048, profiles/ncu_ops_table.txt) -> one CTA per row above that
#define SMG(I, GG) softmax_fwd_group_kernel<T, I, GG><<<(unsigned)((rows + SM_THREADS / GG - 1) / (SM_THREADS / GG)), SM_THREADS, 0, s>>>(xp, mp, yp, rows, heads, q_len, k_len, scale, causal)
        if (nvec <= 8) SMG(1, 8); else if (nvec <= 16) SMG(1, 16); else if (nvec <= 32) SMG(1, 32);
        else if (nvec <= 64) SMG(2, 32); else if (nvec <= 128) SMG(4, 32); else SMG(8, 32);
#undef SMG
        cudaError_t e0 = cudaGetLastError();
        return e0 == cudaSuccess ? 0 : -(int)e0;
    }
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
    const int nvec = k_len >> 3;
    if (nvec <= 64) {
#define SBG(I, GG) softmax_bwd_group_kernel<T, I, GG><<<(unsigned)((rows + SM_THREADS / GG - 1) / (SM_THREADS / GG)), SM_THREADS, 0, s>>>(gp, yp, dp, rows, k_len, scale)
        if (nvec <= 8) SBG(1, 8); else if (nvec <= 16) SBG(1, 16); else if (nvec <= 32) SBG(1, 32);
        else if (nvec <= 64) SBG(2, 32); else if (nvec <= 128) SBG(4, 32); else SBG(8, 32);
#undef SBG
        cudaError_t e0 = cudaGetLastError();
        return e0 == cudaSuccess ? 0 : -(int)e0;
    }
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
