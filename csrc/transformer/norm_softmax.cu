// LayerNorm forward/backward (+fused residual add / fused extra gradient) and attention softmax forward/backward.
// One CTA per row with the row cached in registers (16-byte accesses), warp-shuffle reductions; LayerNorm
// backward keeps per-thread dgamma/dbeta accumulators across a grid-stride loop over rows and emits per-CTA
// partials that a second tiny kernel folds, so gamma/beta gradients cost no atomics and no second pass over dy.
// Reference behaviour: csrc/transformer/normalize_kernels.cu:16-2103, csrc/transformer/softmax_kernels.cu:8-577.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"

namespace dsb {

constexpr int LN_THREADS = 256;
constexpr int LN_MAX_ITERS = 4;  // rows up to 256 * 8 * 4 = 8192 elements stay in registers

// ------------------------------------------------------------------------------------------- LayerNorm forward
template <typename T, int ITERS>
__global__ void __launch_bounds__(LN_THREADS) ln_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                            const T* __restrict__ gamma, const T* __restrict__ beta,
                                                            T* __restrict__ y, T* __restrict__ sum_out,
                                                            float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                            long rows, int hidden, float eps) {
    __shared__ float sh[33];
    const int nvec = hidden >> 3;
    for (long row = blockIdx.x; row < rows; row += gridDim.x) {
        const uint4* xr = reinterpret_cast<const uint4*>(x + row * hidden);
        const uint4* rr = res ? reinterpret_cast<const uint4*>(res + row * hidden) : nullptr;
        float vals[ITERS][8];
        float s = 0.f;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int i = threadIdx.x + it * LN_THREADS;
            if (i < nvec) {
                unpack8<T>(xr[i], vals[it]);
                if (rr) {
                    float r8[8];
                    unpack8<T>(rr[i], r8);
#pragma unroll
                    for (int k = 0; k < 8; ++k) vals[it][k] += r8[k];
                    if (sum_out) {
                        // round through the storage dtype so backward sees exactly what was stored
                        uint4 pk = pack8<T>(vals[it]);
                        reinterpret_cast<uint4*>(sum_out + row * hidden)[i] = pk;
                        unpack8<T>(pk, vals[it]);
                    }
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) s += vals[it][k];
            }
        }
        const float mean = block_sum(s, sh) / hidden;
        float vs = 0.f;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int i = threadIdx.x + it * LN_THREADS;
            if (i < nvec) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    float d = vals[it][k] - mean;
                    vs += d * d;
                }
            }
        }
        const float var = block_sum(vs, sh) / hidden;
        const float rstd = rsqrtf(var + eps);
        if (threadIdx.x == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int i = threadIdx.x + it * LN_THREADS;
            if (i < nvec) {
                float g8[8], b8[8], o[8];
                unpack8<T>(__ldg(reinterpret_cast<const uint4*>(gamma) + i), g8);
                unpack8<T>(__ldg(reinterpret_cast<const uint4*>(beta) + i), b8);
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = (vals[it][k] - mean) * rstd * g8[k] + b8[k];
                reinterpret_cast<uint4*>(y + row * hidden)[i] = pack8<T>(o);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------- LayerNorm backward
template <typename T, int ITERS>
__global__ void __launch_bounds__(LN_THREADS) ln_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                            const float* __restrict__ mean,
                                                            const float* __restrict__ rstd,
                                                            const T* __restrict__ gamma, const T* __restrict__ dres,
                                                            T* __restrict__ dx, float* __restrict__ part_g,
                                                            float* __restrict__ part_b, long rows, int hidden) {
    __shared__ float sh[33];
    const int nvec = hidden >> 3;
    float accg[ITERS][8], accb[ITERS][8], g8[ITERS][8];
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * LN_THREADS;
#pragma unroll
        for (int k = 0; k < 8; ++k) { accg[it][k] = 0.f; accb[it][k] = 0.f; g8[it][k] = 0.f; }
        if (i < nvec) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(gamma) + i), g8[it]);
    }
    for (long row = blockIdx.x; row < rows; row += gridDim.x) {
        const float mu = mean[row], rs = rstd[row];
        float xh[ITERS][8], dyv[ITERS][8];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int i = threadIdx.x + it * LN_THREADS;
            if (i < nvec) {
                unpack8<T>(reinterpret_cast<const uint4*>(x + row * hidden)[i], xh[it]);
                unpack8<T>(reinterpret_cast<const uint4*>(dy + row * hidden)[i], dyv[it]);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    xh[it][k] = (xh[it][k] - mu) * rs;
                    accg[it][k] += dyv[it][k] * xh[it][k];
                    accb[it][k] += dyv[it][k];
                    const float dg = dyv[it][k] * g8[it][k];
                    s1 += dg;
                    s2 += dg * xh[it][k];
                }
            }
        }
        s1 = block_sum(s1, sh) / hidden;
        s2 = block_sum(s2, sh) / hidden;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int i = threadIdx.x + it * LN_THREADS;
            if (i < nvec) {
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = (dyv[it][k] * g8[it][k] - s1 - xh[it][k] * s2) * rs;
                if (dres) {
                    float r8[8];
                    unpack8<T>(reinterpret_cast<const uint4*>(dres + row * hidden)[i], r8);
#pragma unroll
                    for (int k = 0; k < 8; ++k) o[k] += r8[k];
                }
                reinterpret_cast<uint4*>(dx + row * hidden)[i] = pack8<T>(o);
            }
        }
    }
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * LN_THREADS;
        if (i < nvec) {
            float* pg = part_g + (long)blockIdx.x * hidden + i * 8;
            float* pb = part_b + (long)blockIdx.x * hidden + i * 8;
            *reinterpret_cast<float4*>(pg) = make_float4(accg[it][0], accg[it][1], accg[it][2], accg[it][3]);
            *reinterpret_cast<float4*>(pg + 4) = make_float4(accg[it][4], accg[it][5], accg[it][6], accg[it][7]);
            *reinterpret_cast<float4*>(pb) = make_float4(accb[it][0], accb[it][1], accb[it][2], accb[it][3]);
            *reinterpret_cast<float4*>(pb + 4) = make_float4(accb[it][4], accb[it][5], accb[it][6], accb[it][7]);
        }
    }
}
template <typename T>
__global__ void ln_bwd_finalize(const float* __restrict__ part_g, const float* __restrict__ part_b, int nparts,
                                int hidden, T* __restrict__ dgamma, T* __restrict__ dbeta) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= hidden) return;
    float g = 0.f, b = 0.f;
    for (int r = 0; r < nparts; ++r) {
        g += part_g[(long)r * hidden + c];
        b += part_b[(long)r * hidden + c];
    }
    dgamma[c] = Cvt<T>::from_f(g);
    dbeta[c] = Cvt<T>::from_f(b);
}

constexpr int LN_BWD_BLOCKS = 296;  // 148 * 2

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

extern "C" int dsb_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                                 void* sum_out, float* mean, float* rstd, long rows, int hidden, float eps, int dtype,
                                 void* stream) {
    if (rows <= 0) return 0;
    if (hidden % 8 || hidden > LN_THREADS * 8 * LN_MAX_ITERS) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const unsigned grid = (unsigned)(rows < 148L * 16 ? rows : 148L * 16);
    const int iters = ((hidden >> 3) + LN_THREADS - 1) / LN_THREADS;
#define LNF(T, I) ln_fwd_kernel<T, I><<<grid, LN_THREADS, 0, s>>>((const T*)x, (const T*)residual, (const T*)gamma, (const T*)beta, (T*)y, (T*)sum_out, mean, rstd, rows, hidden, eps)
#define LNF_D(T) do { if (iters <= 1) LNF(T, 1); else if (iters <= 2) LNF(T, 2); else LNF(T, 4); } while (0)
    if (dtype == DSB_DTYPE_F16) LNF_D(__half);
    else if (dtype == DSB_DTYPE_BF16) LNF_D(__nv_bfloat16);
    else return DSB_ERR_DTYPE;
#undef LNF_D
#undef LNF
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" long dsb_layernorm_bwd_scratch_bytes(int hidden) { return 2L * LN_BWD_BLOCKS * hidden * sizeof(float); }

extern "C" int dsb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd,
                                 const void* gamma, const void* dres, void* dx, void* dgamma, void* dbeta, long rows,
                                 int hidden, int dtype, void* scratch, void* stream) {
    if (rows <= 0) return 0;
    if (hidden % 8 || hidden > LN_THREADS * 8 * LN_MAX_ITERS) return DSB_ERR_SIZE;
    if (!scratch) return DSB_ERR_ARG;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int grid = (int)(rows < LN_BWD_BLOCKS ? rows : LN_BWD_BLOCKS);
    float* pg = reinterpret_cast<float*>(scratch);
    float* pb = pg + (long)LN_BWD_BLOCKS * hidden;
    const int iters = ((hidden >> 3) + LN_THREADS - 1) / LN_THREADS;
#define LNB(T, I) ln_bwd_kernel<T, I><<<grid, LN_THREADS, 0, s>>>((const T*)dy, (const T*)x, mean, rstd, (const T*)gamma, (const T*)dres, (T*)dx, pg, pb, rows, hidden)
#define LNB_D(T) do { if (iters <= 1) LNB(T, 1); else if (iters <= 2) LNB(T, 2); else LNB(T, 4); \
        ln_bwd_finalize<T><<<(hidden + 255) / 256, 256, 0, s>>>(pg, pb, grid, hidden, (T*)dgamma, (T*)dbeta); } while (0)
    if (dtype == DSB_DTYPE_F16) LNB_D(__half);
    else if (dtype == DSB_DTYPE_BF16) LNB_D(__nv_bfloat16);
    else return DSB_ERR_DTYPE;
#undef LNB_D
#undef LNB
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

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
