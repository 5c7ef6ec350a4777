// Bandwidth-bound transformer helpers: bias+GeLU fwd/bwd, Philox dropout (+bias, +residual) fwd/bwd, QKV
// bias/split/transpose, [0213] transposes, column-sum (bias grad), fused adds, rotary embedding, and a fused
// token-level softmax cross entropy. All use 16-byte accesses with fp32 math in registers.
// Reference behaviour: csrc/transformer/gelu_kernels.cu:38-330, dropout_kernels.cu:5-721,
// transform_kernels.cu:56-575, general_kernels.cu:6-323.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"

namespace dsb {

static int ew_sms() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}
static unsigned ew_grid(long vecs, int threads) {
    long b = (vecs + threads - 1) / threads;
    long cap = (long)ew_sms() * 16;
    if (b > cap) b = cap;
    return (unsigned)(b < 1 ? 1 : b);
}

// ------------------------------------------------------------------------------------------- bias + GeLU
template <typename T>
__global__ void __launch_bounds__(256) bias_gelu_fwd_kernel(const T* __restrict__ x, const T* __restrict__ bias,
                                                            T* __restrict__ y, long nvec, int cvec) {
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float v[8];
        unpack8<T>(reinterpret_cast<const uint4*>(x)[i], v);
        if (bias) {
            float b[8];
            unpack8<T>(__ldg(reinterpret_cast<const uint4*>(bias) + (i % cvec)), b);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += b[k];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = gelu_tanh(v[k]);
        reinterpret_cast<uint4*>(y)[i] = pack8<T>(v);
    }
}
template <typename T>
__global__ void __launch_bounds__(256) bias_gelu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                            const T* __restrict__ bias, T* __restrict__ dx, long nvec,
                                                            int cvec) {
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float v[8], g[8];
        unpack8<T>(reinterpret_cast<const uint4*>(x)[i], v);
        unpack8<T>(reinterpret_cast<const uint4*>(dy)[i], g);
        if (bias) {
            float b[8];
            unpack8<T>(__ldg(reinterpret_cast<const uint4*>(bias) + (i % cvec)), b);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += b[k];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] *= gelu_tanh_grad(v[k]);
        reinterpret_cast<uint4*>(dx)[i] = pack8<T>(g);
    }
}

// ------------------------------------------------------------------------------------------- Philox dropout
struct Philox {
    uint2 key;
    uint4 ctr;
    DSB_DEVICE Philox(unsigned long long seed, unsigned long long subseq, unsigned long long offset) {
        key = make_uint2((unsigned)seed, (unsigned)(seed >> 32));
        ctr = make_uint4((unsigned)offset, (unsigned)(offset >> 32), (unsigned)subseq, (unsigned)(subseq >> 32));
    }
    DSB_DEVICE static uint2 mulhilo(unsigned a, unsigned b) {
        unsigned long long p = (unsigned long long)a * b;
        return make_uint2((unsigned)p, (unsigned)(p >> 32));
    }
    DSB_DEVICE uint4 operator()() const {
        uint4 c = ctr;
        uint2 k = key;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint2 r0 = mulhilo(0xD2511F53u, c.x), r1 = mulhilo(0xCD9E8D57u, c.z);
            c = make_uint4(r1.y ^ c.y ^ k.x, r1.x, r0.y ^ c.w ^ k.y, r0.x);
            k.x += 0x9E3779B9u;
            k.y += 0xBB67AE85u;
        }
        return c;
    }
};

// 8 elements per thread-iteration; two Philox calls give 8 x 32 random bits.
template <typename T>
__global__ void __launch_bounds__(256) dropout_fwd_kernel(const T* __restrict__ x, const T* __restrict__ bias,
                                                          const T* __restrict__ res, T* __restrict__ out,
                                                          unsigned char* __restrict__ mask, long nvec, int cvec,
                                                          float ratio, unsigned long long seed,
                                                          unsigned long long offset) {
    const float scale = 1.f / (1.f - ratio);
    const unsigned thresh = (unsigned)(ratio * 4294967295.0);
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        Philox ph(seed, (unsigned long long)i, offset);
        uint4 ra = ph();
        ph.ctr.x ^= 0x5bd1e995u;
        uint4 rb = ph();
        const unsigned rnd[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
        float v[8];
        unpack8<T>(reinterpret_cast<const uint4*>(x)[i], v);
        if (bias) {
            float b[8];
            unpack8<T>(__ldg(reinterpret_cast<const uint4*>(bias) + (i % cvec)), b);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += b[k];
        }
        unsigned long long mbits = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const bool keep = rnd[k] >= thresh;
            v[k] = keep ? v[k] * scale : 0.f;
            mbits |= (unsigned long long)(keep ? 1 : 0) << (8 * k);
        }
        if (res) {
            float r[8];
            unpack8<T>(reinterpret_cast<const uint4*>(res)[i], r);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += r[k];
        }
        reinterpret_cast<uint4*>(out)[i] = pack8<T>(v);
        reinterpret_cast<unsigned long long*>(mask)[i] = mbits;
    }
}
template <typename T>
__global__ void __launch_bounds__(256) dropout_bwd_kernel(const T* __restrict__ dy,
                                                          const unsigned char* __restrict__ mask, T* __restrict__ dx,
                                                          long nvec, float ratio) {
    const float scale = 1.f / (1.f - ratio);
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float g[8];
        unpack8<T>(reinterpret_cast<const uint4*>(dy)[i], g);
        const unsigned long long mb = reinterpret_cast<const unsigned long long*>(mask)[i];
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = ((mb >> (8 * k)) & 1ull) ? g[k] * scale : 0.f;
        reinterpret_cast<uint4*>(dx)[i] = pack8<T>(g);
    }
}

// ------------------------------------------------------------------------------------------- transforms
// qkv: [B, S, 3, H, D] -> q/k/v: [B, H, S, D], bias: [3, H, D]
template <typename T>
__global__ void __launch_bounds__(256) qkv_split_kernel(const T* __restrict__ qkv, const T* __restrict__ bias,
                                                        T* __restrict__ q, T* __restrict__ k, T* __restrict__ v, int B,
                                                        int S, int H, int D) {
    const int dv = D >> 3;
    const long total = (long)B * S * 3 * H * dv;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        long t = i;
        const int d = t % dv; t /= dv;
        const int h = t % H; t /= H;
        const int w = t % 3; t /= 3;
        const int s = t % S;
        const int b = t / S;
        float val[8];
        unpack8<T>(reinterpret_cast<const uint4*>(qkv)[i], val);
        if (bias) {
            float bb[8];
            unpack8<T>(__ldg(reinterpret_cast<const uint4*>(bias) + ((long)w * H + h) * dv + d), bb);
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) val[kk] += bb[kk];
        }
        T* dst = w == 0 ? q : (w == 1 ? k : v);
        reinterpret_cast<uint4*>(dst)[(((long)b * H + h) * S + s) * dv + d] = pack8<T>(val);
    }
}
// x: [B, d1, d2, D] -> y: [B, d2, d1, D]
template <typename T>
__global__ void __launch_bounds__(256) transpose_0213_kernel(const T* __restrict__ x, T* __restrict__ y, int B,
                                                             int d1, int d2, int D) {
    const int dv = D >> 3;
    const long total = (long)B * d1 * d2 * dv;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        long t = i;
        const int d = t % dv; t /= dv;
        const int j = t % d2; t /= d2;
        const int a = t % d1;
        const int b = t / d1;
        reinterpret_cast<uint4*>(y)[(((long)b * d2 + j) * d1 + a) * dv + d] = reinterpret_cast<const uint4*>(x)[i];
    }
}

// column sum: each CTA owns 64 columns (8 vec-lanes x 32 row-lanes), grid.y splits rows, fp32 atomics at the end
template <typename T>
__global__ void __launch_bounds__(256) column_sum_kernel(const T* __restrict__ x, float* __restrict__ acc, long rows,
                                                         int cols) {
    __shared__ float sh[32][65];
    const int cvec = cols >> 3;
    const int lane_c = threadIdx.x & 7;    // which 8-col vector inside the 64-col stripe
    const int lane_r = threadIdx.x >> 3;   // 0..31
    const int vcol = blockIdx.x * 8 + lane_c;
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (vcol < cvec) {
        for (long r = (long)blockIdx.y * 32 + lane_r; r < rows; r += (long)gridDim.y * 32) {
            float v[8];
            unpack8<T>(reinterpret_cast<const uint4*>(x)[r * cvec + vcol], v);
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] += v[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) sh[lane_r][lane_c * 8 + k] = s[k];
    __syncthreads();
    if (threadIdx.x < 64) {
        float t = 0.f;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) t += sh[r][threadIdx.x];
        const int c = blockIdx.x * 64 + threadIdx.x;
        if (c < cols) atomicAdd(acc + c, t);
    }
}
template <typename T>
__global__ void colsum_finalize(const float* __restrict__ acc, T* __restrict__ out, int cols, int accumulate) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < cols) out[c] = Cvt<T>::from_f(acc[c] + (accumulate ? Cvt<T>::to_f(out[c]) : 0.f));
}

template <typename T>
__global__ void __launch_bounds__(256) fused_add_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                        const T* __restrict__ c, const T* __restrict__ d,
                                                        T* __restrict__ out, long nvec) {
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float va[8], vb[8];
        unpack8<T>(reinterpret_cast<const uint4*>(a)[i], va);
        unpack8<T>(reinterpret_cast<const uint4*>(b)[i], vb);
#pragma unroll
        for (int k = 0; k < 8; ++k) va[k] += vb[k];
        if (c) {
            unpack8<T>(reinterpret_cast<const uint4*>(c)[i], vb);
#pragma unroll
            for (int k = 0; k < 8; ++k) va[k] += vb[k];
        }
        if (d) {
            unpack8<T>(reinterpret_cast<const uint4*>(d)[i], vb);
#pragma unroll
            for (int k = 0; k < 8; ++k) va[k] += vb[k];
        }
        reinterpret_cast<uint4*>(out)[i] = pack8<T>(va);
    }
}

// ------------------------------------------------------------------------------------------- rotary (NeoX half split)
template <typename T>
__global__ void __launch_bounds__(256) rotary_kernel(T* __restrict__ q, T* __restrict__ k, int B, int S, int H, int D,
                                                     int rot, long sb, long ss, long sh_, float log2_base, int inverse) {
    const int half = rot >> 1;
    const long total = (long)B * S * H * half;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        long t = i;
        const int j = t % half; t /= half;
        const int h = t % H; t /= H;
        const int s = t % S;
        const int b = t / S;
        const float inv_freq = exp2f(-log2_base * (2.f * j) / rot);
        float sn, cs;
        sincosf(s * inv_freq, &sn, &cs);
        if (inverse) sn = -sn;
        const long off = b * sb + s * ss + h * sh_;
        {
            float x0 = Cvt<T>::to_f(q[off + j]), x1 = Cvt<T>::to_f(q[off + j + half]);
            q[off + j] = Cvt<T>::from_f(x0 * cs - x1 * sn);
            q[off + j + half] = Cvt<T>::from_f(x1 * cs + x0 * sn);
        }
        {
            float x0 = Cvt<T>::to_f(k[off + j]), x1 = Cvt<T>::to_f(k[off + j + half]);
            k[off + j] = Cvt<T>::from_f(x0 * cs - x1 * sn);
            k[off + j + half] = Cvt<T>::from_f(x1 * cs + x0 * sn);
        }
    }
}

// ------------------------------------------------------------------------------------------- cross entropy
constexpr int CE_THREADS = 512;
template <typename T>
__global__ void __launch_bounds__(CE_THREADS) ce_fwd_kernel(const T* __restrict__ logits,
                                                            const long* __restrict__ targets, float* __restrict__ loss,
                                                            float* __restrict__ lse_out, int vocab, long ld,
                                                            long ignore_index) {
    __shared__ float sh[33];
    const long row = blockIdx.x;
    const T* xr = logits + row * ld;
    const int nvec = vocab >> 3;
    // online max / sum-exp in one pass
    float mx = -INFINITY, sum = 0.f;
    for (int i = threadIdx.x; i < nvec; i += CE_THREADS) {
        float v[8];
        unpack8<T>(ld_nc_v4(reinterpret_cast<const uint4*>(xr) + i), v);
        float lm = v[0];
#pragma unroll
        for (int k = 1; k < 8; ++k) lm = fmaxf(lm, v[k]);
        const float nm = fmaxf(mx, lm);
        float ls = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) ls += __expf(v[k] - nm);
        sum = sum * __expf(mx - nm) + ls;
        mx = nm;
    }
    for (int i = (nvec << 3) + threadIdx.x; i < vocab; i += CE_THREADS) {
        const float v = Cvt<T>::to_f(xr[i]);
        const float nm = fmaxf(mx, v);
        sum = sum * __expf(mx - nm) + __expf(v - nm);
        mx = nm;
    }
    const float gmx = block_max(mx, sh);
    sum = (mx == -INFINITY) ? 0.f : sum * __expf(mx - gmx);
    sum = block_sum(sum, sh);
    if (threadIdx.x == 0) {
        const float lse = gmx + __logf(sum);
        lse_out[row] = lse;
        const long t = targets[row];
        loss[row] = (t == ignore_index || t < 0 || t >= vocab) ? 0.f : lse - Cvt<T>::to_f(xr[t]);
    }
}
template <typename T>
__global__ void __launch_bounds__(CE_THREADS) ce_bwd_kernel(T* __restrict__ logits, const long* __restrict__ targets,
                                                            const float* __restrict__ lse,
                                                            const float* __restrict__ gs_dev, float gs, int vocab,
                                                            long ld, long ignore_index) {
    const long row = blockIdx.x;
    T* xr = logits + row * ld;
    const long t = targets[row];
    if (gs_dev) gs *= *gs_dev;
    if (t == ignore_index || t < 0 || t >= vocab) gs = 0.f;
    const float l = lse[row];
    const int nvec = vocab >> 3;
    for (int i = threadIdx.x; i < nvec; i += CE_THREADS) {
        float v[8];
        unpack8<T>(reinterpret_cast<const uint4*>(xr)[i], v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float p = __expf(v[k] - l);
            if (i * 8 + k == t) p -= 1.f;
            v[k] = p * gs;
        }
        reinterpret_cast<uint4*>(xr)[i] = pack8<T>(v);
    }
    for (int i = (nvec << 3) + threadIdx.x; i < vocab; i += CE_THREADS) {
        float p = __expf(Cvt<T>::to_f(xr[i]) - l);
        if (i == t) p -= 1.f;
        xr[i] = Cvt<T>::from_f(p * gs);
    }
}

}  // namespace dsb

using namespace dsb;

#define DSB_RET()                                  \
    do {                                           \
        cudaError_t e__ = cudaGetLastError();      \
        return e__ == cudaSuccess ? 0 : -(int)e__; \
    } while (0)
#define DSB_DISPATCH_HALF(dtype, CALL)                           \
    if (dtype == DSB_DTYPE_F16) { using T = __half; CALL; }      \
    else if (dtype == DSB_DTYPE_BF16) { using T = __nv_bfloat16; CALL; } \
    else return DSB_ERR_DTYPE;

extern "C" int dsb_bias_gelu_fwd(const void* x, const void* bias, void* y, long rows, int cols, int dtype,
                                 void* stream) {
    if (rows <= 0) return 0;
    if (cols % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long nvec = rows * (cols >> 3);
    DSB_DISPATCH_HALF(dtype, (bias_gelu_fwd_kernel<T><<<ew_grid(nvec, 256), 256, 0, s>>>((const T*)x, (const T*)bias, (T*)y, nvec, cols >> 3)));
    DSB_RET();
}
extern "C" int dsb_bias_gelu_bwd(const void* dy, const void* x, const void* bias, void* dx, long rows, int cols,
                                 int dtype, void* stream) {
    if (rows <= 0) return 0;
    if (cols % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long nvec = rows * (cols >> 3);
    DSB_DISPATCH_HALF(dtype, (bias_gelu_bwd_kernel<T><<<ew_grid(nvec, 256), 256, 0, s>>>((const T*)dy, (const T*)x, (const T*)bias, (T*)dx, nvec, cols >> 3)));
    DSB_RET();
}
extern "C" int dsb_dropout_fwd(const void* x, const void* bias, const void* residual, void* out, unsigned char* mask,
                               long rows, int cols, float ratio, unsigned long long seed, unsigned long long offset,
                               int dtype, void* stream) {
    if (rows <= 0) return 0;
    if (cols % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long nvec = rows * (cols >> 3);
    DSB_DISPATCH_HALF(dtype, (dropout_fwd_kernel<T><<<ew_grid(nvec, 256), 256, 0, s>>>((const T*)x, (const T*)bias, (const T*)residual, (T*)out, mask, nvec, cols >> 3, ratio, seed, offset)));
    DSB_RET();
}
extern "C" int dsb_dropout_bwd(const void* dy, const unsigned char* mask, void* dx, long n, float ratio, int dtype,
                               void* stream) {
    if (n <= 0) return 0;
    if (n % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long nvec = n >> 3;
    DSB_DISPATCH_HALF(dtype, (dropout_bwd_kernel<T><<<ew_grid(nvec, 256), 256, 0, s>>>((const T*)dy, mask, (T*)dx, nvec, ratio)));
    DSB_RET();
}
extern "C" int dsb_qkv_bias_split_transpose(const void* qkv, const void* bias, void* q, void* k, void* v, int B, int S,
                                            int H, int D, int dtype, void* stream) {
    if (D % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long nvec = (long)B * S * 3 * H * (D >> 3);
    if (nvec <= 0) return 0;
    DSB_DISPATCH_HALF(dtype, (qkv_split_kernel<T><<<ew_grid(nvec, 256), 256, 0, s>>>((const T*)qkv, (const T*)bias, (T*)q, (T*)k, (T*)v, B, S, H, D)));
    DSB_RET();
}
extern "C" int dsb_transpose_0213(const void* x, void* y, int B, int d1, int d2, int D, int dtype, void* stream) {
    if (D % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long nvec = (long)B * d1 * d2 * (D >> 3);
    if (nvec <= 0) return 0;
    DSB_DISPATCH_HALF(dtype, (transpose_0213_kernel<T><<<ew_grid(nvec, 256), 256, 0, s>>>((const T*)x, (T*)y, B, d1, d2, D)));
    DSB_RET();
}

extern "C" int dsb_column_sum(const void* x, void* out, long rows, int cols, int dtype, int accumulate,
                              float* scratch, void* stream) {
    if (cols % 8) return DSB_ERR_SIZE;
    if (rows <= 0 || cols <= 0) return 0;
    if (!scratch) return DSB_ERR_ARG;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    float* acc = scratch;
    cudaMemsetAsync(acc, 0, sizeof(float) * cols, s);
    const int gx = (cols + 63) / 64;
    long gy = (long)ew_sms() * 4 / gx;
    if (gy < 1) gy = 1;
    if (gy > (rows + 31) / 32) gy = (rows + 31) / 32;
    dim3 grid(gx, (unsigned)gy);
    DSB_DISPATCH_HALF(dtype, (column_sum_kernel<T><<<grid, 256, 0, s>>>((const T*)x, acc, rows, cols),
                              colsum_finalize<T><<<(cols + 255) / 256, 256, 0, s>>>(acc, (T*)out, cols, accumulate)));
    DSB_RET();
}
extern "C" int dsb_fused_add(const void* a, const void* b, const void* c, const void* d, void* out, long n, int dtype,
                             void* stream) {
    if (n <= 0) return 0;
    if (n % 8) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long nvec = n >> 3;
    DSB_DISPATCH_HALF(dtype, (fused_add_kernel<T><<<ew_grid(nvec, 256), 256, 0, s>>>((const T*)a, (const T*)b, (const T*)c, (const T*)d, (T*)out, nvec)));
    DSB_RET();
}
extern "C" int dsb_rotary(void* q, void* k, int B, int S, int H, int D, int rot_dim, long stride_b, long stride_s,
                          long stride_h, float base, int inverse, int dtype, void* stream) {
    if (rot_dim % 2 || rot_dim > D) return DSB_ERR_SIZE;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const long total = (long)B * S * H * (rot_dim / 2);
    if (total <= 0) return 0;
    const float l2b = log2f(base);
    DSB_DISPATCH_HALF(dtype, (rotary_kernel<T><<<ew_grid(total, 256), 256, 0, s>>>((T*)q, (T*)k, B, S, H, D, rot_dim, stride_b, stride_s, stride_h, l2b, inverse)));
    DSB_RET();
}
extern "C" int dsb_cross_entropy_fwd(const void* logits, const long* targets, float* loss, float* lse, long tokens,
                                     int vocab, long ld, long ignore_index, int dtype, void* stream) {
    if (tokens <= 0) return 0;
    if (ld % 8) return DSB_ERR_ALIGN;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    DSB_DISPATCH_HALF(dtype, (ce_fwd_kernel<T><<<(unsigned)tokens, CE_THREADS, 0, s>>>((const T*)logits, targets, loss, lse, vocab, ld, ignore_index)));
    DSB_RET();
}
extern "C" int dsb_cross_entropy_bwd(void* logits, const long* targets, const float* lse, const float* grad_scale_dev,
                                     float grad_scale, long tokens, int vocab, long ld, long ignore_index, int dtype,
                                     void* stream) {
    if (tokens <= 0) return 0;
    if (ld % 8) return DSB_ERR_ALIGN;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    DSB_DISPATCH_HALF(dtype, (ce_bwd_kernel<T><<<(unsigned)tokens, CE_THREADS, 0, s>>>((T*)logits, targets, lse, grad_scale_dev, grad_scale, vocab, ld, ignore_index)));
    DSB_RET();
}

// ------------------------------------------------------------------------------------------- embedding
// out[t, :] = table[ids[t], :] and its gradient dtable[ids[t], :] += dout[t, :]. The backward adds with packed 16-bit
// atomics straight into the caller's gradient buffer (the flat gradient arena), so there is no dense [vocab, dim]
// temporary, no sort and no separate accumulate pass (torch's embedding_dense_backward: sort + segmented sum into a
// fresh tensor, then AccumulateGrad adds it: 1.3 ms per 8192 x 2048 micro-batch against 50432 rows). Tokens that
// repeat add in arbitrary order (16-bit rounding may differ in the last bit between runs).
namespace dsb {
template <typename T>
__global__ void __launch_bounds__(256) embedding_fwd_kernel(const T* __restrict__ table, const long* __restrict__ ids,
                                                            T* __restrict__ out, long ntok, int dim, long vocab) {
    const int vec = dim >> 3;
    for (long t = blockIdx.x; t < ntok; t += gridDim.x) {
        const long row = ids[t];
        if (row < 0 || row >= vocab) continue;      // out-of-range ids read as zeros would hide bugs: leave untouched
        const uint4* src = reinterpret_cast<const uint4*>(table + row * dim);
        uint4* dst = reinterpret_cast<uint4*>(out + t * dim);
        for (int i = threadIdx.x; i < vec; i += blockDim.x) dst[i] = __ldg(src + i);
    }
}

template <typename T2>
__global__ void __launch_bounds__(256) embedding_bwd_kernel(const T2* __restrict__ dout, const long* __restrict__ ids,
                                                            T2* __restrict__ dtable, long ntok, int dim2, long vocab) {
    for (long t = blockIdx.x; t < ntok; t += gridDim.x) {
        const long row = ids[t];
        if (row < 0 || row >= vocab) continue;
        const T2* src = dout + t * dim2;
        T2* dst = dtable + row * dim2;
        for (int i = threadIdx.x; i < dim2; i += blockDim.x) atomicAdd(dst + i, src[i]);
    }
}
}  // namespace dsb

extern "C" int dsb_embedding_fwd(const void* table, const long* ids, void* out, long ntok, int dim, long vocab,
                                 int dtype, void* stream) {
    if (dim % 8) return DSB_ERR_SIZE;
    if (ntok <= 0) return 0;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const unsigned grid = (unsigned)(ntok < (long)ew_sms() * 16 ? ntok : (long)ew_sms() * 16);
    if (dtype == DSB_DTYPE_F16)
        dsb::embedding_fwd_kernel<__half><<<grid, 256, 0, s>>>((const __half*)table, ids, (__half*)out, ntok, dim, vocab);
    else if (dtype == DSB_DTYPE_BF16)
        dsb::embedding_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)table, ids,
                                                                     (__nv_bfloat16*)out, ntok, dim, vocab);
    else
        return DSB_ERR_DTYPE;
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_embedding_bwd_acc(const void* dout, const long* ids, void* dtable, long ntok, int dim, long vocab,
                                     int dtype, void* stream) {
    if (dim % 2) return DSB_ERR_SIZE;
    if (ntok <= 0) return 0;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const unsigned grid = (unsigned)(ntok < (long)ew_sms() * 16 ? ntok : (long)ew_sms() * 16);
    if (dtype == DSB_DTYPE_F16)
        dsb::embedding_bwd_kernel<__half2><<<grid, 256, 0, s>>>((const __half2*)dout, ids, (__half2*)dtable, ntok,
                                                                dim / 2, vocab);
    else if (dtype == DSB_DTYPE_BF16)
        dsb::embedding_bwd_kernel<__nv_bfloat162><<<grid, 256, 0, s>>>((const __nv_bfloat162*)dout, ids,
                                                                       (__nv_bfloat162*)dtable, ntok, dim / 2, vocab);
    else
        return DSB_ERR_DTYPE;
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

// ------------------------------------------------------------------------------------------- allocation-free column sum
// Same reduction as dsb_column_sum, for callers that cannot hand over (or obtain) scratch memory: a plug-in or host
// callback that is not allowed to touch the CUDA allocation API, or a process whose HBM has been given wholesale to
// another allocator so that cudaMalloc would fail. The fp32 accumulators live in a fixed file-scope device array
// instead. Every GPU that loads this library carries that array (DSB_COLSUM_NOALLOC_MAX_COLS floats = 256 KiB)
// until the process exits.
//
// The array's name is only a host-side handle; the device address is looked up with cudaGetSymbolAddress the first
// time the entry point runs on a given device (with that device current) and cached per device. Make that first
// call outside of any stream capture. All calls on one GPU share the array and it still holds the previous call's
// sums, so each call zeroes the columns it uses first (stream-ordered); two calls that may run at the same time on
// one GPU (different streams) would corrupt each other and must be serialised by the caller.
namespace dsb {
static __device__ __align__(256) float g_colsum_noalloc_scratch[DSB_COLSUM_NOALLOC_MAX_COLS];
static float* g_colsum_noalloc_ptr[64] = {nullptr};
}  // namespace dsb

extern "C" int dsb_column_sum_noalloc(const void* x, void* out, long rows, int cols, int dtype, int accumulate,
                                      void* stream) {
    if (cols % 8) return DSB_ERR_SIZE;
    if (cols > DSB_COLSUM_NOALLOC_MAX_COLS) return DSB_ERR_SIZE;  // beyond the built-in scratch: use dsb_column_sum
    if (rows <= 0 || cols <= 0) return 0;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return -(int)e;
    if (dev < 0 || dev >= 64) return DSB_ERR_ARG;
    if (dsb::g_colsum_noalloc_ptr[dev] == nullptr) {
        float* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), dsb::g_colsum_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) return e != cudaSuccess ? -(int)e : DSB_ERR_ARG;
        dsb::g_colsum_noalloc_ptr[dev] = p;
    }
    float* acc = dsb::g_colsum_noalloc_ptr[dev];
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    e = cudaMemsetAsync(acc, 0, sizeof(float) * cols, s);
    if (e != cudaSuccess) return -(int)e;
    const int gx = (cols + 63) / 64;
    long gy = (long)ew_sms() * 4 / gx;
    if (gy < 1) gy = 1;
    if (gy > (rows + 31) / 32) gy = (rows + 31) / 32;
    dim3 grid(gx, (unsigned)gy);
    DSB_DISPATCH_HALF(dtype, (column_sum_kernel<T><<<grid, 256, 0, s>>>((const T*)x, acc, rows, cols),
                              colsum_finalize<T><<<(cols + 255) / 256, 256, 0, s>>>(acc, (T*)out, cols, accumulate)));
    DSB_RET();
}
