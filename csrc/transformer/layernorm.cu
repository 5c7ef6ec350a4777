// LayerNorm forward/backward (+fused residual add / fused extra gradient), bandwidth-bound design for B200.
//
// A CTA of 8 warps walks the rows in groups of R rows per iteration; thread t owns the same 8-element column slices
// in every row. Three things keep HBM busy (the first version of this file, one row per CTA iteration with two
// block barriers per row, reached only 30-35 % of the copy bandwidth, profiles/bench_ops_v1.json):
//   * the rows of a group are loaded back to back and the NEXT group's loads are issued before the current group is
//     reduced, so every thread always has R x ITERS x (2..3) independent 16-byte loads in flight;
//   * operands stay packed (raw 16-byte registers) between the passes and are re-unpacked instead of being held as
//     fp32, which keeps the kernel at 2-3 resident CTAs per SM;
//   * all row statistics of a group go through ONE block reduction (a single __syncthreads over a double-buffered
//     shared array).
// The backward keeps its dgamma/dbeta accumulators in registers across the whole grid-stride loop and emits one
// partial row per CTA; a second small kernel folds the partials with a [32 columns x 8 row-slices] CTA shape.
// Reference behaviour: csrc/transformer/normalize_kernels.cu:16-2103 (one CTA per row, two passes for gamma/beta).
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"

namespace dsb {

constexpr int LN_THREADS = 256;
constexpr int LN_WARPS = LN_THREADS / 32;
constexpr int LN_MAX_ITERS = 4;      // rows up to 256 * 8 * 4 = 8192 elements stay in registers
constexpr int LN_BWD_BLOCKS = 296;   // 148 SMs x 2 resident CTAs; also the number of dgamma/dbeta partial rows

// Sum NV independent values over the CTA with one barrier. `sh` holds 2 x LN_WARPS x NV floats (double buffered:
// a buffer is rewritten only after every thread passed the barrier of the following reduction).
template <int NV>
DSB_DEVICE void block_sum_n(float (&v)[NV], float* sh, int& flip) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
    float* buf = sh + flip * (LN_WARPS * NV);
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) buf[warp * NV + i] = v[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < LN_WARPS; ++w) s += buf[w * NV + i];
        v[i] = s;
    }
    flip ^= 1;
}

// Load the raw 16-byte slices of a group of R rows (zeros past the end).
template <typename T, int ITERS, int R>
DSB_DEVICE void load_group(uint4 (&dst)[R][ITERS], const T* __restrict__ src, long row0, long rows, int hidden,
                           int nvec) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const long row = row0 + r;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int i = threadIdx.x + it * LN_THREADS;
            dst[r][it] = make_uint4(0, 0, 0, 0);
            if (src != nullptr && row < rows && i < nvec)
                dst[r][it] = ld_nc_v4(reinterpret_cast<const uint4*>(src + row * hidden) + i);
        }
    }
}

// ------------------------------------------------------------------------------------------- forward
template <typename T, int ITERS, int R>
__global__ void __launch_bounds__(LN_THREADS, ITERS == 1 ? 3 : 2)
ln_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res, const T* __restrict__ gamma,
              const T* __restrict__ beta, T* __restrict__ y, T* __restrict__ sum_out, float* __restrict__ mean_out,
              float* __restrict__ rstd_out, long rows, int hidden, float eps) {
    __shared__ float sh[2 * LN_WARPS * R];
    int flip = 0;
    const int nvec = hidden >> 3;
    const float inv_h = 1.f / hidden;
    const long stride = (long)gridDim.x * R;
    uint4 cx[R][ITERS], cr[R][ITERS];
    load_group<T, ITERS, R>(cx, x, (long)blockIdx.x * R, rows, hidden, nvec);
    load_group<T, ITERS, R>(cr, res, (long)blockIdx.x * R, rows, hidden, nvec);
    for (long row0 = (long)blockIdx.x * R; row0 < rows; row0 += stride) {
        uint4 nx[R][ITERS], nr[R][ITERS];
        load_group<T, ITERS, R>(nx, x, row0 + stride, rows, hidden, nvec);
        load_group<T, ITERS, R>(nr, res, row0 + stride, rows, hidden, nvec);
        float s[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long row = row0 + r;
            s[r] = 0.f;
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int i = threadIdx.x + it * LN_THREADS;
                float v[8];
                unpack8<T>(cx[r][it], v);
                if (res) {
                    float r8[8];
                    unpack8<T>(cr[r][it], r8);
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] += r8[k];
                    // round through the storage dtype so backward sees exactly what was stored
                    cx[r][it] = pack8<T>(v);
                    if (sum_out && row < rows && i < nvec)
                        reinterpret_cast<uint4*>(sum_out + row * hidden)[i] = cx[r][it];
                    unpack8<T>(cx[r][it], v);
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) s[r] += v[k];
            }
        }
        block_sum_n<R>(s, sh, flip);
        float vs[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            s[r] *= inv_h;   // mean
            vs[r] = 0.f;
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                if (threadIdx.x + it * LN_THREADS < nvec) {
                    float v[8];
                    unpack8<T>(cx[r][it], v);
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float d = v[k] - s[r];
                        vs[r] += d * d;
                    }
                }
            }
        }
        block_sum_n<R>(vs, sh, flip);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long row = row0 + r;
            if (row < rows) {
                const float rstd = rsqrtf(vs[r] * inv_h + eps);
                if (threadIdx.x == 0) {
                    if (mean_out) mean_out[row] = s[r];
                    if (rstd_out) rstd_out[row] = rstd;
                }
#pragma unroll
                for (int it = 0; it < ITERS; ++it) {
                    const int i = threadIdx.x + it * LN_THREADS;
                    if (i < nvec) {
                        float v[8], g8[8], b8[8], o[8];
                        unpack8<T>(cx[r][it], v);
                        unpack8<T>(__ldg(reinterpret_cast<const uint4*>(gamma) + i), g8);
                        unpack8<T>(__ldg(reinterpret_cast<const uint4*>(beta) + i), b8);
#pragma unroll
                        for (int k = 0; k < 8; ++k) o[k] = (v[k] - s[r]) * rstd * g8[k] + b8[k];
                        reinterpret_cast<uint4*>(y + row * hidden)[i] = pack8<T>(o);
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
#pragma unroll
            for (int it = 0; it < ITERS; ++it) { cx[r][it] = nx[r][it]; cr[r][it] = nr[r][it]; }
        }
    }
}

// ------------------------------------------------------------------------------------------- forward, one warp per row
// For rows of 256 * VEC elements (VEC <= 16: up to 4096) a warp owns a whole row: lane l holds the 16-byte vectors
// l, l + 32, ... (coalesced 512-byte requests), the two row statistics are warp shuffles, and no block barrier exists,
// so the eight warps of a CTA stream independently. gamma / beta are staged
// once per CTA in shared memory. The block-cooperative kernel above (two barriers per row group) reached 50 % (plain) /
// 80 % (with residual) of the copy bandwidth on 8192 x 2048 (profiles/bench_ops_v4.json); this one removes the
// barriers from the streaming loop.
template <typename T, int VEC>
__global__ void __launch_bounds__(LN_THREADS, VEC <= 4 ? 4 : (VEC <= 8 ? 2 : 1))
ln_fwd_warp_kernel(const T* __restrict__ x, const T* __restrict__ res, const T* __restrict__ gamma,
                   const T* __restrict__ beta, T* __restrict__ y, T* __restrict__ sum_out, float* __restrict__ mean_out,
                   float* __restrict__ rstd_out, long rows, float eps) {
    constexpr int HIDDEN = 256 * VEC;
    __shared__ uint4 sg[32 * VEC], sb[32 * VEC];
    for (int i = threadIdx.x; i < 32 * VEC; i += LN_THREADS) {
        sg[i] = __ldg(reinterpret_cast<const uint4*>(gamma) + i);
        sb[i] = __ldg(reinterpret_cast<const uint4*>(beta) + i);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long warp_id = (long)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    const long nwarps = (long)gridDim.x * LN_WARPS;
    const float inv_h = 1.f / HIDDEN;
    uint4 cx[VEC], cr[VEC];
    auto load_row = [&](uint4 (&dx)[VEC], uint4 (&dr)[VEC], long row) {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            dx[i] = make_uint4(0, 0, 0, 0);
            dr[i] = make_uint4(0, 0, 0, 0);
            if (row < rows) {
                dx[i] = ld_nc_v4(reinterpret_cast<const uint4*>(x + row * HIDDEN) + lane + 32 * i);
                if (res != nullptr) dr[i] = ld_nc_v4(reinterpret_cast<const uint4*>(res + row * HIDDEN) + lane + 32 * i);
            }
        }
    };
    // no register prefetch of the next row: 24 - 32 resident warps per SM, each with its whole row in flight, already
    // keep several hundred KB outstanding per SM
    for (long row = warp_id; row < rows; row += nwarps) {
        load_row(cx, cr, row);
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            float v[8];
            unpack8<T>(cx[i], v);
            if (res != nullptr) {
                float r8[8];
                unpack8<T>(cr[i], r8);
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] += r8[k];
                cx[i] = pack8<T>(v);      // round through the storage dtype: backward sees exactly what is stored
                if (sum_out != nullptr) reinterpret_cast<uint4*>(sum_out + row * HIDDEN)[lane + 32 * i] = cx[i];
                unpack8<T>(cx[i], v);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) sum += v[k];
        }
        const float mu = warp_sum(sum) * inv_h;
        float var = 0.f;
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            float v[8];
            unpack8<T>(cx[i], v);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float d = v[k] - mu;
                var += d * d;
            }
        }
        const float rstd = rsqrtf(warp_sum(var) * inv_h + eps);
        if (lane == 0) {
            if (mean_out != nullptr) mean_out[row] = mu;
            if (rstd_out != nullptr) rstd_out[row] = rstd;
        }
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            float v[8], g8[8], b8[8], o[8];
            unpack8<T>(cx[i], v);
            unpack8<T>(sg[lane + 32 * i], g8);
            unpack8<T>(sb[lane + 32 * i], b8);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = (v[k] - mu) * rstd * g8[k] + b8[k];
            reinterpret_cast<uint4*>(y + row * HIDDEN)[lane + 32 * i] = pack8<T>(o);
        }
    }
}

// ------------------------------------------------------------------------------------------- backward
// dx = rstd * (dy*gamma - mean_h(dy*gamma) - xhat * mean_h(dy*gamma*xhat)) (+ dres); dgamma += dy*xhat; dbeta += dy.
template <typename T, int ITERS, int R, bool INV>
__global__ void __launch_bounds__(LN_THREADS, ITERS == 1 ? 2 : 1)
ln_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const float* __restrict__ mean,
              const float* __restrict__ rstd, const T* __restrict__ gamma, const T* __restrict__ dres,
              T* __restrict__ dx, float* __restrict__ part_g, float* __restrict__ part_b, long rows, int hidden,
              const T* __restrict__ beta_inv) {
    // beta_inv != nullptr: "invertible" mode. `x` then holds the layer's OUTPUT y and the normalised input is recovered
    // as (y - beta) / gamma, so the forward pass does not have to keep its input (mean is unused, rstd still needed).
    __shared__ float sh[2 * LN_WARPS * 2 * R];
    int flip = 0;
    const int nvec = hidden >> 3;
    const float inv_h = 1.f / hidden;
    const long stride = (long)gridDim.x * R;
    float accg[ITERS][8], accb[ITERS][8];
    uint4 graw[ITERS], braw[ITERS];
    constexpr bool inv = INV;      // compile-time: the common (non-invertible) path carries no beta registers / divisions
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
        const int i = threadIdx.x + it * LN_THREADS;
#pragma unroll
        for (int k = 0; k < 8; ++k) { accg[it][k] = 0.f; accb[it][k] = 0.f; }
        graw[it] = i < nvec ? __ldg(reinterpret_cast<const uint4*>(gamma) + i) : make_uint4(0, 0, 0, 0);
        braw[it] = make_uint4(0, 0, 0, 0);
        if (inv && i < nvec) braw[it] = __ldg(reinterpret_cast<const uint4*>(beta_inv) + i);
    }
    uint4 cx[R][ITERS], cdy[R][ITERS], cdr[R][ITERS];
    float cmu[R], crs[R];
    {
        const long row0 = (long)blockIdx.x * R;
        load_group<T, ITERS, R>(cx, x, row0, rows, hidden, nvec);
        load_group<T, ITERS, R>(cdy, dy, row0, rows, hidden, nvec);
        load_group<T, ITERS, R>(cdr, dres, row0, rows, hidden, nvec);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const bool ok = row0 + r < rows;
            cmu[r] = ok ? __ldg(mean + row0 + r) : 0.f;
            crs[r] = ok ? __ldg(rstd + row0 + r) : 0.f;
        }
    }
    for (long row0 = (long)blockIdx.x * R; row0 < rows; row0 += stride) {
        // next group's loads go out before this group's reduction
        uint4 nx[R][ITERS], ndy[R][ITERS], ndr[R][ITERS];
        float nmu[R], nrs[R];
        load_group<T, ITERS, R>(nx, x, row0 + stride, rows, hidden, nvec);
        load_group<T, ITERS, R>(ndy, dy, row0 + stride, rows, hidden, nvec);
        load_group<T, ITERS, R>(ndr, dres, row0 + stride, rows, hidden, nvec);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const bool ok = row0 + stride + r < rows;
            nmu[r] = ok ? __ldg(mean + row0 + stride + r) : 0.f;
            nrs[r] = ok ? __ldg(rstd + row0 + stride + r) : 0.f;
        }
        float red[2 * R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float s1 = 0.f, s2 = 0.f;
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                float xv[8], dyv[8], g8[8], b8[8];
                unpack8<T>(cx[r][it], xv);
                unpack8<T>(cdy[r][it], dyv);
                unpack8<T>(graw[it], g8);
                unpack8<T>(braw[it], b8);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    // rows past the end have rs = 0 and dy = 0 -> contribute nothing
                    const float xh = inv ? (g8[k] != 0.f ? (xv[k] - b8[k]) / g8[k] : 0.f) : (xv[k] - cmu[r]) * crs[r];
                    accg[it][k] += dyv[k] * xh;
                    accb[it][k] += dyv[k];
                    const float dg = dyv[k] * g8[k];
                    s1 += dg;
                    s2 += dg * xh;
                }
            }
            red[2 * r] = s1;
            red[2 * r + 1] = s2;
        }
        block_sum_n<2 * R>(red, sh, flip);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long row = row0 + r;
            if (row < rows) {
                const float s1 = red[2 * r] * inv_h, s2 = red[2 * r + 1] * inv_h;
#pragma unroll
                for (int it = 0; it < ITERS; ++it) {
                    const int i = threadIdx.x + it * LN_THREADS;
                    if (i < nvec) {
                        float xv[8], dyv[8], g8[8], b8[8], o[8];
                        unpack8<T>(cx[r][it], xv);
                        unpack8<T>(cdy[r][it], dyv);
                        unpack8<T>(graw[it], g8);
                        unpack8<T>(braw[it], b8);
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const float xh = inv ? (g8[k] != 0.f ? (xv[k] - b8[k]) / g8[k] : 0.f)
                                                 : (xv[k] - cmu[r]) * crs[r];
                            o[k] = (dyv[k] * g8[k] - s1 - xh * s2) * crs[r];
                        }
                        if (dres) {
                            float r8[8];
                            unpack8<T>(cdr[r][it], r8);
#pragma unroll
                            for (int k = 0; k < 8; ++k) o[k] += r8[k];
                        }
                        reinterpret_cast<uint4*>(dx + row * hidden)[i] = pack8<T>(o);
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            cmu[r] = nmu[r];
            crs[r] = nrs[r];
#pragma unroll
            for (int it = 0; it < ITERS; ++it) { cx[r][it] = nx[r][it]; cdy[r][it] = ndy[r][it]; cdr[r][it] = ndr[r][it]; }
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

// Fold the per-CTA partial rows: CTA = 32 columns x 32 row slices (1024 threads), one output column per lane, shared
// memory cross-slice sum. (With 8 slices every thread walked 37 rows back to back and the kernel took 10.5 us, a third
// of the backward itself: profiles/ncu_ops_table.txt.)
constexpr int FIN_SLICES = 32;
template <typename T>
__global__ void __launch_bounds__(32 * FIN_SLICES) ln_bwd_finalize(const float* __restrict__ part_g,
                                                                   const float* __restrict__ part_b, int nparts,
                                                                   int hidden, T* __restrict__ dgamma,
                                                                   T* __restrict__ dbeta, int accumulate) {
    __shared__ float sg[FIN_SLICES][33], sb[FIN_SLICES][33];
    const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    float g = 0.f, b = 0.f;
    if (c < hidden) {
#pragma unroll 4
        for (int r = slice; r < nparts; r += FIN_SLICES) {
            g += part_g[(long)r * hidden + c];
            b += part_b[(long)r * hidden + c];
        }
    }
    sg[slice][lane] = g;
    sb[slice][lane] = b;
    __syncthreads();
    if (slice == 0 && c < hidden) {
#pragma unroll
        for (int s = 1; s < FIN_SLICES; ++s) { g += sg[s][lane]; b += sb[s][lane]; }
        if (accumulate) {   // add into the parameters' existing gradient (views of the flat gradient arena)
            g += Cvt<T>::to_f(dgamma[c]);
            b += Cvt<T>::to_f(dbeta[c]);
        }
        dgamma[c] = Cvt<T>::from_f(g);
        dbeta[c] = Cvt<T>::from_f(b);
    }
}

}  // namespace dsb

using namespace dsb;

extern "C" int dsb_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                                 void* sum_out, float* mean, float* rstd, long rows, int hidden, float eps, int dtype,
                                 void* stream) {
    if (rows <= 0) return 0;
    if (hidden % 8 || hidden > LN_THREADS * 8 * LN_MAX_ITERS) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (hidden % 256 == 0 && hidden <= 4096 && (hidden / 256 == 1 || hidden / 256 == 2 || hidden / 256 == 3 ||
                                                  hidden / 256 == 4 || hidden / 256 == 6 || hidden / 256 == 8 ||
                                                  hidden / 256 == 16)) {
        // one warp per row (no block barriers in the streaming loop)
        const long want = (rows + LN_WARPS - 1) / LN_WARPS;
        const unsigned g = (unsigned)(want < 148L * 4 ? want : 148L * 4);
#define LNW(T, V) ln_fwd_warp_kernel<T, V><<<g, LN_THREADS, 0, s>>>((const T*)x, (const T*)residual, (const T*)gamma, (const T*)beta, (T*)y, (T*)sum_out, mean, rstd, rows, eps)
#define LNW_D(T) do { switch (hidden / 256) { case 1: LNW(T, 1); break; case 2: LNW(T, 2); break; case 3: LNW(T, 3); break; \
        case 4: LNW(T, 4); break; case 6: LNW(T, 6); break; case 8: LNW(T, 8); break; default: LNW(T, 16); break; } } while (0)
        if (dtype == DSB_DTYPE_F16) LNW_D(__half);
        else if (dtype == DSB_DTYPE_BF16) LNW_D(__nv_bfloat16);
        else return DSB_ERR_DTYPE;
#undef LNW_D
#undef LNW
        cudaError_t we = cudaGetLastError();
        return we == cudaSuccess ? 0 : -(int)we;
    }
    const int iters = ((hidden >> 3) + LN_THREADS - 1) / LN_THREADS;
    const int R = iters <= 1 ? 4 : (iters <= 2 ? 2 : 1);
    const long groups = (rows + R - 1) / R;
    const long cap = 148L * (iters <= 1 ? 3 : 2);
    const unsigned grid = (unsigned)(groups < cap ? groups : cap);
#define LNF(T, I, RR) ln_fwd_kernel<T, I, RR><<<grid, LN_THREADS, 0, s>>>((const T*)x, (const T*)residual, (const T*)gamma, (const T*)beta, (T*)y, (T*)sum_out, mean, rstd, rows, hidden, eps)
#define LNF_D(T) do { if (iters <= 1) LNF(T, 1, 4); else if (iters <= 2) LNF(T, 2, 2); else LNF(T, 4, 1); } while (0)
    if (dtype == DSB_DTYPE_F16) LNF_D(__half);
    else if (dtype == DSB_DTYPE_BF16) LNF_D(__nv_bfloat16);
    else return DSB_ERR_DTYPE;
#undef LNF_D
#undef LNF
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" long dsb_layernorm_bwd_scratch_bytes(int hidden) { return 2L * LN_BWD_BLOCKS * hidden * sizeof(float); }

// accumulate != 0: dgamma/dbeta += (they point at the parameters' gradient buffers) instead of being overwritten.
static int ln_bwd_impl(const void* dy, const void* x, const float* mean, const float* rstd, const void* gamma,
                       const void* beta_inv, const void* dres, void* dx, void* dgamma, void* dbeta, long rows, int hidden,
                       int dtype, int accumulate, void* scratch, void* stream) {
    if (rows <= 0) return 0;
    if (hidden % 8 || hidden > LN_THREADS * 8 * LN_MAX_ITERS) return DSB_ERR_SIZE;
    if (!scratch) return DSB_ERR_ARG;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int iters = ((hidden >> 3) + LN_THREADS - 1) / LN_THREADS;
    const int R = iters <= 1 ? 2 : 1;
    const long groups = (rows + R - 1) / R;
    const long cap = iters <= 1 ? LN_BWD_BLOCKS : 148;
    const int grid = (int)(groups < cap ? groups : cap);
    float* pg = reinterpret_cast<float*>(scratch);
    float* pb = pg + (long)LN_BWD_BLOCKS * hidden;
#define LNB1(T, I, RR, V) ln_bwd_kernel<T, I, RR, V><<<grid, LN_THREADS, 0, s>>>((const T*)dy, (const T*)x, mean, rstd, (const T*)gamma, (const T*)dres, (T*)dx, pg, pb, rows, hidden, (const T*)beta_inv)
#define LNB(T, I, RR) do { if (beta_inv != nullptr) LNB1(T, I, RR, true); else LNB1(T, I, RR, false); } while (0)
#define LNB_D(T) do { if (iters <= 1) LNB(T, 1, 2); else if (iters <= 2) LNB(T, 2, 1); else LNB(T, 4, 1); \
        ln_bwd_finalize<T><<<(hidden + 31) / 32, 32 * FIN_SLICES, 0, s>>>(pg, pb, grid, hidden, (T*)dgamma, (T*)dbeta, accumulate); } while (0)
    if (dtype == DSB_DTYPE_F16) LNB_D(__half);
    else if (dtype == DSB_DTYPE_BF16) LNB_D(__nv_bfloat16);
    else return DSB_ERR_DTYPE;
#undef LNB_D
#undef LNB
#undef LNB1
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_layernorm_bwd_acc(const void* dy, const void* x, const float* mean, const float* rstd,
                                     const void* gamma, const void* dres, void* dx, void* dgamma, void* dbeta,
                                     long rows, int hidden, int dtype, int accumulate, void* scratch, void* stream) {
    return ln_bwd_impl(dy, x, mean, rstd, gamma, nullptr, dres, dx, dgamma, dbeta, rows, hidden, dtype, accumulate,
                       scratch, stream);
}

// Invertible variant: takes the layer's OUTPUT y (plus gamma, beta, rstd) instead of its input; the forward pass can
// drop the input. Reference: the *_invertible kernels of csrc/transformer/normalize_kernels.cu:602-672,1056-1348.
extern "C" int dsb_layernorm_bwd_inv(const void* dy, const void* y, const float* rstd, const void* gamma,
                                     const void* beta, const void* dres, void* dx, void* dgamma, void* dbeta, long rows,
                                     int hidden, int dtype, int accumulate, void* scratch, void* stream) {
    if (!beta) return DSB_ERR_ARG;
    return ln_bwd_impl(dy, y, rstd /* unused mean */, rstd, gamma, beta, dres, dx, dgamma, dbeta, rows, hidden, dtype,
                       accumulate, scratch, stream);
}

extern "C" int dsb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd,
                                 const void* gamma, const void* dres, void* dx, void* dgamma, void* dbeta, long rows,
                                 int hidden, int dtype, void* scratch, void* stream) {
    return dsb_layernorm_bwd_acc(dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, rows, hidden, dtype, 0, scratch,
                                 stream);
}
