// Block-sparse attention kernels for sm_100a: SDD (sparse = dense x dense^T), DSD (dense = sparse x dense, with
// optional transposed sparse operand) and block-sparse softmax forward / backward.
// They replace the reference's Triton-C 0.4 kernels (deepspeed/ops/sparse_attention/trsrc/matmul.tr:4-201,
// softmax_fwd.tr:4-136, softmax_bwd.tr:4-54), which cannot be built any more.
//
// Storage: dense operands [B, H, S, D] (D contiguous); sparse operand [B, nnz, block, block] in layout order
// (head-major, then block-row, then block-column). LUTs (built on the host by csrc/sparse_attn/lut.cpp):
//   blocks  : int32 [nnz, 4]  = (head, block_row, block_col, 0)                     -> SDD
//   row_ptr : int32 [H*nb + 1], row_idx int32 [nnz, 2] = (block_col, block_id)      -> DSD, softmax
//   col_ptr : int32 [H*nb + 1], col_idx int32 [nnz, 2] = (block_row, block_id)      -> DSD with transposed sparse
// Tiles are 16 x 16 x 16 tensor-core fragments (fp16 / bf16 in, fp32 accumulate); one warp owns one output fragment.
// Round-1 note: fragments go through the warp-level MMA path; packing several 16/32-wide blocks into tcgen05 tiles
//  is the follow-up once the LUT format is settled.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"
#include <mma.h>

using namespace nvcuda;

namespace dsb {

template <typename T> struct WmmaT;
template <> struct WmmaT<__half> { using type = __half; };
template <> struct WmmaT<__nv_bfloat16> { using type = __nv_bfloat16; };

// ------------------------------------------------------------------------------------------------ SDD
// C[b, blk] = alpha * A[b, h, i*bs : (i+1)*bs, :] * B[b, h, j*bs : (j+1)*bs, :]^T        grid (nnz, B)
template <typename T>
__global__ void __launch_bounds__(512) bs_sdd_kernel(const T* __restrict__ A, const T* __restrict__ Bm,
                                                     T* __restrict__ C, const int* __restrict__ blocks, int H, int S,
                                                     int D, int bs, long nnz, float alpha) {
    const int blk = blockIdx.x, b = blockIdx.y;
    const int h = blocks[blk * 4 + 0], bi = blocks[blk * 4 + 1], bj = blocks[blk * 4 + 2];
    const int warp = threadIdx.x >> 5;
    const int tiles = bs >> 4;
    if (warp >= tiles * tiles) return;
    const int tr = warp / tiles, tc = warp % tiles;
    const T* a = A + (((long)b * H + h) * S + (long)bi * bs + tr * 16) * D;
    const T* bb = Bm + (((long)b * H + h) * S + (long)bj * bs + tc * 16) * D;
    wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc;
    wmma::fill_fragment(acc, 0.f);
    for (int k = 0; k < D; k += 16) {
        wmma::fragment<wmma::matrix_a, 16, 16, 16, T, wmma::row_major> fa;
        wmma::fragment<wmma::matrix_b, 16, 16, 16, T, wmma::col_major> fb;
        wmma::load_matrix_sync(fa, a + k, D);
        wmma::load_matrix_sync(fb, bb + k, D);   // (k, n) -> B[row n, k]: col-major with ld = D
        wmma::mma_sync(acc, fa, fb, acc);
    }
    __shared__ float stage[32][16 * 16];
    float* st = stage[warp];
    wmma::store_matrix_sync(st, acc, 16, wmma::mem_row_major);
    __syncwarp();
    T* c = C + (((long)b * nnz + blk) * bs + tr * 16) * bs + tc * 16;
    const int lane = threadIdx.x & 31;
    for (int e = lane; e < 256; e += 32) c[(e >> 4) * bs + (e & 15)] = Cvt<T>::from_f(st[e] * alpha);
}

// ------------------------------------------------------------------------------------------------ DSD
// out[b, h, r*bs : (r+1)*bs, :] = sum over LUT entries (o, blk) of op(P[b, blk]) * X[b, h, o*bs : (o+1)*bs, :]
// transposed = 0: LUT = row lists, op = identity;  transposed = 1: LUT = column lists, op = transpose.
// grid (H*nb, D/64 (ceil), B); warps tile the (bs x 64) output slab into 16x16 fragments.
template <typename T>
__global__ void __launch_bounds__(512) bs_dsd_kernel(const T* __restrict__ P, const T* __restrict__ X,
                                                     T* __restrict__ O, const int* __restrict__ ptr,
                                                     const int* __restrict__ idx, int H, int S, int D, int bs, long nnz,
                                                     int nb, int transposed) {
    const int hr = blockIdx.x, b = blockIdx.z;
    const int h = hr / nb, r = hr % nb;
    const int d0 = blockIdx.y * 64;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rt = bs >> 4;                               // fragment rows
    const int ct = min(4, (D - d0 + 15) >> 4);            // fragment columns in this 64-wide slab
    if (warp >= rt * ct) return;
    const int tr = warp / ct, tc = warp % ct;
    wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc;
    wmma::fill_fragment(acc, 0.f);
    const int beg = ptr[hr], end = ptr[hr + 1];
    for (int e = beg; e < end; ++e) {
        const int o = idx[e * 2 + 0], blk = idx[e * 2 + 1];
        const T* p = P + ((long)b * nnz + blk) * bs * bs;
        const T* x = X + (((long)b * H + h) * S + (long)o * bs) * D + d0 + tc * 16;
        for (int k = 0; k < bs; k += 16) {
            wmma::fragment<wmma::matrix_b, 16, 16, 16, T, wmma::row_major> fb;
            wmma::load_matrix_sync(fb, x + (long)k * D, D);
            if (!transposed) {
                wmma::fragment<wmma::matrix_a, 16, 16, 16, T, wmma::row_major> fa;
                wmma::load_matrix_sync(fa, p + (tr * 16) * bs + k, bs);
                wmma::mma_sync(acc, fa, fb, acc);
            } else {
                // op(P)[m, k] = P[k, m]: read the stored block column-major
                wmma::fragment<wmma::matrix_a, 16, 16, 16, T, wmma::col_major> fa;
                wmma::load_matrix_sync(fa, p + k * bs + tr * 16, bs);
                wmma::mma_sync(acc, fa, fb, acc);
            }
        }
    }
    __shared__ float stage[32][16 * 16];
    float* st = stage[warp];
    wmma::store_matrix_sync(st, acc, 16, wmma::mem_row_major);
    __syncwarp();
    T* o = O + (((long)b * H + h) * S + (long)r * bs + tr * 16) * D + d0 + tc * 16;
    for (int e = lane; e < 256; e += 32) {
        const int col = d0 + tc * 16 + (e & 15);
        if (col < D) o[(long)(e >> 4) * D + (e & 15)] = Cvt<T>::from_f(st[e]);
    }
}

// ------------------------------------------------------------------------------------------------ softmax
// One warp per (batch, head, block-row, row-in-block): the row's elements live in the row's non-zero blocks.
// x = x * scale + rpe[h, row, col] + kp_mask[b, col] (add) or -inf where kp_mask == 0 (mul) + attn_mask[row, col].
template <typename T>
__global__ void __launch_bounds__(256) bs_softmax_fwd_kernel(T* __restrict__ X, const int* __restrict__ ptr,
                                                             const int* __restrict__ idx, int H, int nb, int bs,
                                                             long nnz, int S, float scale, const T* __restrict__ rpe,
                                                             const T* __restrict__ kpm, int kpm_mul,
                                                             const T* __restrict__ am, int am_mul) {
    const int warp_global = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.y;
    const int rows_total = H * nb * bs;
    if (warp_global >= rows_total) return;
    const int hr = warp_global / bs, rr = warp_global % bs;
    const int h = hr / nb, r = hr % nb;
    const int beg = ptr[hr], end = ptr[hr + 1];
    const int n = (end - beg) * bs;
    const long row = (long)r * bs + rr;
    auto load = [&](int e) -> float {
        const int k = beg + e / bs, c = e % bs;
        const int col_blk = idx[k * 2 + 0], blk = idx[k * 2 + 1];
        const long col = (long)col_blk * bs + c;
        float v = Cvt<T>::to_f(X[(((long)b * nnz + blk) * bs + rr) * bs + c]) * scale;
        if (rpe) v += Cvt<T>::to_f(rpe[((long)h * S + row) * S + col]);
        if (kpm) {
            const float m = Cvt<T>::to_f(kpm[(long)b * S + col]);
            if (kpm_mul) { if (m == 0.f) v = -INFINITY; } else v += m;
        }
        if (am) {
            const float m = Cvt<T>::to_f(am[row * S + col]);
            if (am_mul) { if (m == 0.f) v = -INFINITY; } else v += m;
        }
        return v;
    };
    float mx = -INFINITY;
    for (int e = lane; e < n; e += 32) mx = fmaxf(mx, load(e));
    mx = warp_max(mx);
    if (mx == -INFINITY) mx = 0.f;
    float sum = 0.f;
    for (int e = lane; e < n; e += 32) sum += __expf(load(e) - mx);
    sum = warp_sum(sum);
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
    for (int e = lane; e < n; e += 32) {
        const int k = beg + e / bs, c = e % bs;
        const int blk = idx[k * 2 + 1];
        const float v = __expf(load(e) - mx) * inv;
        X[(((long)b * nnz + blk) * bs + rr) * bs + c] = Cvt<T>::from_f(v);
    }
}

// dX = Y * (dY - sum(Y * dY)) * scale, in place on dY
template <typename T>
__global__ void __launch_bounds__(256) bs_softmax_bwd_kernel(T* __restrict__ dY, const T* __restrict__ Y,
                                                             const int* __restrict__ ptr, const int* __restrict__ idx,
                                                             int H, int nb, int bs, long nnz, float scale) {
    const int warp_global = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.y;
    if (warp_global >= H * nb * bs) return;
    const int hr = warp_global / bs, rr = warp_global % bs;
    const int beg = ptr[hr], end = ptr[hr + 1];
    const int n = (end - beg) * bs;
    float dot = 0.f;
    for (int e = lane; e < n; e += 32) {
        const int blk = idx[(beg + e / bs) * 2 + 1];
        const long off = (((long)b * nnz + blk) * bs + rr) * bs + e % bs;
        dot += Cvt<T>::to_f(Y[off]) * Cvt<T>::to_f(dY[off]);
    }
    dot = warp_sum(dot);
    for (int e = lane; e < n; e += 32) {
        const int blk = idx[(beg + e / bs) * 2 + 1];
        const long off = (((long)b * nnz + blk) * bs + rr) * bs + e % bs;
        dY[off] = Cvt<T>::from_f(Cvt<T>::to_f(Y[off]) * (Cvt<T>::to_f(dY[off]) - dot) * scale);
    }
}

}  // namespace dsb

using namespace dsb;

#define BS_RET()                                   \
    do {                                           \
        cudaError_t e__ = cudaGetLastError();      \
        return e__ == cudaSuccess ? 0 : -(int)e__; \
    } while (0)

extern "C" int dsb_bs_sdd(const void* A, const void* B, void* C, const int* blocks, int batch, int H, int S, int D,
                          int block, long nnz, float alpha, int dtype, void* stream) {
    if (block % 16 || block > 64 || D % 16 || nnz <= 0) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int warps = (block / 16) * (block / 16);
    dim3 grid((unsigned)nnz, batch);
    if (dtype == DSB_DTYPE_F16)
        bs_sdd_kernel<__half><<<grid, warps * 32, 0, s>>>((const __half*)A, (const __half*)B, (__half*)C, blocks, H, S, D, block, nnz, alpha);
    else if (dtype == DSB_DTYPE_BF16)
        bs_sdd_kernel<__nv_bfloat16><<<grid, warps * 32, 0, s>>>((const __nv_bfloat16*)A, (const __nv_bfloat16*)B, (__nv_bfloat16*)C, blocks, H, S, D, block, nnz, alpha);
    else return DSB_ERR_DTYPE;
    BS_RET();
}

extern "C" int dsb_bs_dsd(const void* P, const void* X, void* O, const int* ptr, const int* idx, int batch, int H,
                          int S, int D, int block, long nnz, int transposed, int dtype, void* stream) {
    if (block % 16 || block > 64 || D % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int nb = S / block;
    const int warps = (block / 16) * 4;
    dim3 grid(H * nb, (D + 63) / 64, batch);
    if (dtype == DSB_DTYPE_F16)
        bs_dsd_kernel<__half><<<grid, warps * 32, 0, s>>>((const __half*)P, (const __half*)X, (__half*)O, ptr, idx, H, S, D, block, nnz, nb, transposed);
    else if (dtype == DSB_DTYPE_BF16)
        bs_dsd_kernel<__nv_bfloat16><<<grid, warps * 32, 0, s>>>((const __nv_bfloat16*)P, (const __nv_bfloat16*)X, (__nv_bfloat16*)O, ptr, idx, H, S, D, block, nnz, nb, transposed);
    else return DSB_ERR_DTYPE;
    BS_RET();
}

extern "C" int dsb_bs_softmax_fwd(void* X, const int* ptr, const int* idx, int batch, int H, int S, int block,
                                  long nnz, float scale, const void* rpe, const void* kp_mask, int kp_mul,
                                  const void* attn_mask, int attn_mul, int dtype, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int nb = S / block;
    const int rows = H * nb * block;
    dim3 grid((rows + 7) / 8, batch);
    if (dtype == DSB_DTYPE_F16)
        bs_softmax_fwd_kernel<__half><<<grid, 256, 0, s>>>((__half*)X, ptr, idx, H, nb, block, nnz, S, scale, (const __half*)rpe, (const __half*)kp_mask, kp_mul, (const __half*)attn_mask, attn_mul);
    else if (dtype == DSB_DTYPE_BF16)
        bs_softmax_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((__nv_bfloat16*)X, ptr, idx, H, nb, block, nnz, S, scale, (const __nv_bfloat16*)rpe, (const __nv_bfloat16*)kp_mask, kp_mul, (const __nv_bfloat16*)attn_mask, attn_mul);
    else return DSB_ERR_DTYPE;
    BS_RET();
}

extern "C" int dsb_bs_softmax_bwd(void* dY, const void* Y, const int* ptr, const int* idx, int batch, int H, int S,
                                  int block, long nnz, float scale, int dtype, void* stream) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int nb = S / block;
    const int rows = H * nb * block;
    dim3 grid((rows + 7) / 8, batch);
    if (dtype == DSB_DTYPE_F16)
        bs_softmax_bwd_kernel<__half><<<grid, 256, 0, s>>>((__half*)dY, (const __half*)Y, ptr, idx, H, nb, block, nnz, scale);
    else if (dtype == DSB_DTYPE_BF16)
        bs_softmax_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((__nv_bfloat16*)dY, (const __nv_bfloat16*)Y, ptr, idx, H, nb, block, nnz, scale);
    else return DSB_ERR_DTYPE;
    BS_RET();
}
