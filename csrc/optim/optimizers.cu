// Fused optimizer kernels for sm_100a: flat-arena Adam/AdamW, multi-tensor Adam, LAMB, casts, sum-of-squares.
// All are HBM-bandwidth bound, so the design goal is: one pass, 16-byte accesses, fp32 math in registers,
// grid sized to a multiple of the 148 SMs, no host synchronisation (scale / skip flags are read on the device).
// Reference behaviour: csrc/adam/multi_tensor_adam.cu:28-121, csrc/lamb/fused_lamb_cuda_kernel.cu:186-310,
// csrc/adam/custom_cuda_kernel.cu:5-20.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"

namespace dsb {

static int g_sms = 0;
static int num_sms() {
    if (!g_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_sms <= 0) g_sms = 148;
    }
    return g_sms;
}
static int grid_for(long work_items, int per_block, int waves = 8) {
    long b = (work_items + per_block - 1) / per_block;
    long cap = (long)num_sms() * waves;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

template <typename T> DSB_DEVICE void load4(const T* p, long i, float (&o)[4]);
template <> DSB_DEVICE void load4<float>(const float* p, long i, float (&o)[4]) {
    float4 t = *reinterpret_cast<const float4*>(p + i);
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
}
template <> DSB_DEVICE void load4<__half>(const __half* p, long i, float (&o)[4]) {
    uint2 t = *reinterpret_cast<const uint2*>(p + i);
    float2 a = Cvt<__half>::unpack2(t.x), b = Cvt<__half>::unpack2(t.y);
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
template <> DSB_DEVICE void load4<__nv_bfloat16>(const __nv_bfloat16* p, long i, float (&o)[4]) {
    uint2 t = *reinterpret_cast<const uint2*>(p + i);
    float2 a = Cvt<__nv_bfloat16>::unpack2(t.x), b = Cvt<__nv_bfloat16>::unpack2(t.y);
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
template <typename T> DSB_DEVICE void store4(T* p, long i, const float (&o)[4]);
template <> DSB_DEVICE void store4<float>(float* p, long i, const float (&o)[4]) {
    *reinterpret_cast<float4*>(p + i) = make_float4(o[0], o[1], o[2], o[3]);
}
template <> DSB_DEVICE void store4<__half>(__half* p, long i, const float (&o)[4]) {
    uint2 t;
    t.x = Cvt<__half>::pack2(o[0], o[1]);
    t.y = Cvt<__half>::pack2(o[2], o[3]);
    *reinterpret_cast<uint2*>(p + i) = t;
}
template <> DSB_DEVICE void store4<__nv_bfloat16>(__nv_bfloat16* p, long i, const float (&o)[4]) {
    uint2 t;
    t.x = Cvt<__nv_bfloat16>::pack2(o[0], o[1]);
    t.y = Cvt<__nv_bfloat16>::pack2(o[2], o[3]);
    *reinterpret_cast<uint2*>(p + i) = t;
}

struct AdamHyper {
    float lr, beta1, beta2, eps, wd, bc1, bc2, inv_scale;
    int adamw;
};

DSB_DEVICE void adam_math(float& p, float& m, float& v, float g, const AdamHyper& h) {
    if (!h.adamw) g += h.wd * p;
    m = h.beta1 * m + (1.f - h.beta1) * g;
    v = h.beta2 * v + (1.f - h.beta2) * g * g;
    float upd = (m / h.bc1) / (sqrtf(v / h.bc2) + h.eps);
    if (h.adamw) upd += h.wd * p;
    p -= h.lr * upd;
}

// --------------------------------------------------------------------------------------------- flat Adam
template <typename G, typename LP>
__global__ void __launch_bounds__(512) adam_flat_kernel(float* __restrict__ p, float* __restrict__ m,
                                                        float* __restrict__ v, const G* __restrict__ g,
                                                        LP* __restrict__ lp, long n, AdamHyper h,
                                                        const float* __restrict__ inv_scale_dev,
                                                        const int* __restrict__ skip_dev) {
    if (skip_dev != nullptr && *skip_dev != 0) return;
    if (inv_scale_dev != nullptr) h.inv_scale *= *inv_scale_dev;
    const long n4 = n >> 2;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float pv[4], mv[4], vv[4], gv[4];
        load4(p, i * 4, pv);
        load4(m, i * 4, mv);
        load4(v, i * 4, vv);
        load4(g, i * 4, gv);
#pragma unroll
        for (int k = 0; k < 4; ++k) adam_math(pv[k], mv[k], vv[k], gv[k] * h.inv_scale, h);
        store4(p, i * 4, pv);
        store4(m, i * 4, mv);
        store4(v, i * 4, vv);
        if (lp != nullptr) store4(lp, i * 4, pv);
    }
    // tail (< 4 elements)
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        long i = (n4 << 2) + threadIdx.x;
        float pv = p[i], mv = m[i], vv = v[i];
        adam_math(pv, mv, vv, Cvt<G>::to_f(g[i]) * h.inv_scale, h);
        p[i] = pv; m[i] = mv; v[i] = vv;
        if (lp != nullptr) lp[i] = Cvt<LP>::from_f(pv);
    }
}

template <typename G>
static int adam_flat_dispatch(float* p, float* m, float* v, const void* g, void* lp, int lp_dtype, long n,
                              const AdamHyper& h, const float* isd, const int* sd, cudaStream_t s) {
    const int grid = grid_for(n / 4 + 1, 512);
    const G* gg = reinterpret_cast<const G*>(g);
    if (lp == nullptr || lp_dtype == DSB_DTYPE_F16)
        adam_flat_kernel<G, __half><<<grid, 512, 0, s>>>(p, m, v, gg, reinterpret_cast<__half*>(lp), n, h, isd, sd);
    else if (lp_dtype == DSB_DTYPE_BF16)
        adam_flat_kernel<G, __nv_bfloat16><<<grid, 512, 0, s>>>(p, m, v, gg, reinterpret_cast<__nv_bfloat16*>(lp), n, h, isd, sd);
    else
        adam_flat_kernel<G, float><<<grid, 512, 0, s>>>(p, m, v, gg, reinterpret_cast<float*>(lp), n, h, isd, sd);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

// --------------------------------------------------------------------------------------------- multi-tensor Adam
constexpr int MT_MAX = 24;
struct MTList {
    void* p[MT_MAX];
    void* g[MT_MAX];
    void* m[MT_MAX];
    void* v[MT_MAX];
    long n[MT_MAX];
};

template <typename P, typename G, typename S>
__global__ void __launch_bounds__(512) adam_multi_kernel(MTList l, AdamHyper h) {
    const int t = blockIdx.y;
    P* p = reinterpret_cast<P*>(l.p[t]);
    const G* g = reinterpret_cast<const G*>(l.g[t]);
    S* m = reinterpret_cast<S*>(l.m[t]);
    S* v = reinterpret_cast<S*>(l.v[t]);
    const long n = l.n[t];
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float pv = Cvt<P>::to_f(p[i]), mv = Cvt<S>::to_f(m[i]), vv = Cvt<S>::to_f(v[i]);
        adam_math(pv, mv, vv, Cvt<G>::to_f(g[i]) * h.inv_scale, h);
        p[i] = Cvt<P>::from_f(pv);
        m[i] = Cvt<S>::from_f(mv);
        v[i] = Cvt<S>::from_f(vv);
    }
}

// --------------------------------------------------------------------------------------------- LAMB
// phase 1: moments + per-CTA partial sums of w^2 and u^2 ; phase 2: trust ratio + apply (update recomputed).
constexpr int LAMB_BLOCKS = 592;  // 148 * 4
struct LambHyper {
    float lr, beta1, beta2, eps, wd, max_coeff, min_coeff, step_size, grad_scale;
    int eps_inside;
};
DSB_DEVICE float lamb_update(float p, float m, float v, const LambHyper& h) {
    float denom = h.eps_inside ? sqrtf(v + h.eps) : sqrtf(v) + h.eps;
    return m / denom + h.wd * p;
}
template <typename P, typename G>
__global__ void __launch_bounds__(512) lamb_phase1(const P* __restrict__ p, const G* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, long n, LambHyper h,
                                                   float* __restrict__ partials) {
    __shared__ float sh[33];
    float w2 = 0.f, u2 = 0.f;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float pv = Cvt<P>::to_f(p[i]);
        float gv = Cvt<G>::to_f(g[i]) / h.grad_scale;
        float mv = h.beta1 * m[i] + (1.f - h.beta1) * gv;
        float vv = h.beta2 * v[i] + (1.f - h.beta2) * gv * gv;
        m[i] = mv;
        v[i] = vv;
        float u = lamb_update(pv, mv, vv, h);
        w2 += pv * pv;
        u2 += u * u;
    }
    w2 = block_sum(w2, sh);
    u2 = block_sum(u2, sh);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = w2;
        partials[gridDim.x + blockIdx.x] = u2;
    }
}
template <typename P, typename C>
__global__ void __launch_bounds__(512) lamb_phase2(P* __restrict__ p, C* __restrict__ pcopy,
                                                   const float* __restrict__ m, const float* __restrict__ v, long n,
                                                   LambHyper h, const float* __restrict__ partials, int nparts,
                                                   float* __restrict__ coeff_out) {
    __shared__ float sh[33];
    float w2 = 0.f, u2 = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) {
        w2 += partials[i];
        u2 += partials[nparts + i];
    }
    w2 = block_sum(w2, sh);
    u2 = block_sum(u2, sh);
    float coeff = 1.f;
    if (w2 > 0.f && u2 > 0.f) coeff = fminf(fmaxf(sqrtf(w2) / sqrtf(u2), h.min_coeff), h.max_coeff);
    if (blockIdx.x == 0 && threadIdx.x == 0 && coeff_out != nullptr) *coeff_out = coeff;
    const float s = h.step_size * coeff;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float pv = Cvt<P>::to_f(p[i]);
        pv -= s * lamb_update(pv, m[i], v[i], h);
        p[i] = Cvt<P>::from_f(pv);
        if (pcopy != nullptr) pcopy[i] = Cvt<C>::from_f(pv);
    }
}

// --------------------------------------------------------------------------------------------- cast / scale / sumsq
template <typename S, typename D>
__global__ void __launch_bounds__(512) cast_kernel(const S* __restrict__ s, D* __restrict__ d, long n) {
    const long n4 = n >> 2;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float t[4];
        load4(s, i * 4, t);
        store4(d, i * 4, t);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        long i = (n4 << 2) + threadIdx.x;
        d[i] = Cvt<D>::from_f(Cvt<S>::to_f(s[i]));
    }
}
template <typename T>
__global__ void __launch_bounds__(512) scale_kernel(T* __restrict__ x, long n, float s, const float* __restrict__ sd) {
    if (sd != nullptr) s *= *sd;
    const long n4 = n >> 2;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float t[4];
        load4(x, i * 4, t);
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] *= s;
        store4(x, i * 4, t);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        long i = (n4 << 2) + threadIdx.x;
        x[i] = Cvt<T>::from_f(Cvt<T>::to_f(x[i]) * s);
    }
}

constexpr int SUMSQ_BLOCKS = 1184;  // 148 * 8
template <typename T>
__global__ void __launch_bounds__(512) sumsq_phase1(const T* __restrict__ x, long n, float* __restrict__ partials,
                                                    int* __restrict__ flags) {
    __shared__ float sh[33];
    float acc = 0.f;
    int bad = 0;
    const long n4 = n >> 2;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float t[4];
        load4(x, i * 4, t);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            acc += t[k] * t[k];
            bad |= !isfinite(t[k]);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        float t = Cvt<T>::to_f(x[(n4 << 2) + threadIdx.x]);
        acc += t * t;
        bad |= !isfinite(t);
    }
    acc = block_sum(acc, sh);
    bad = __syncthreads_or(bad);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = acc;
        flags[blockIdx.x] = bad;
    }
}
__global__ void __launch_bounds__(512) sumsq_phase2(const float* __restrict__ partials, const int* __restrict__ flags,
                                                    int nparts, float* __restrict__ out, int* __restrict__ flag_out,
                                                    int accumulate) {
    __shared__ float sh[33];
    float acc = 0.f;
    int bad = 0;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) {
        acc += partials[i];
        bad |= flags[i];
    }
    acc = block_sum(acc, sh);
    bad = __syncthreads_or(bad);
    if (threadIdx.x == 0) {
        if (!isfinite(acc)) bad = 1;
        *out = accumulate ? *out + acc : acc;
        if (flag_out != nullptr && bad) *flag_out = 1;
    }
}

}  // namespace dsb

using namespace dsb;

#define DSB_CHECK_LAUNCH()                         \
    do {                                           \
        cudaError_t e__ = cudaGetLastError();      \
        return e__ == cudaSuccess ? 0 : -(int)e__; \
    } while (0)

extern "C" int dsb_adam_flat(float* p, float* m, float* v, const void* g, int g_dtype, void* p_lp, int lp_dtype,
                             long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                             int adamw_mode, int bias_correction, int step, float inv_scale,
                             const float* inv_scale_dev, const int* skip_dev, void* stream) {
    if (n <= 0) return 0;
    if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
         reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(p_lp)) & 15)
        return DSB_ERR_ALIGN;
    AdamHyper h;
    h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.wd = weight_decay; h.adamw = adamw_mode;
    h.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
    h.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
    h.inv_scale = inv_scale;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    switch (g_dtype) {
        case DSB_DTYPE_F16: return adam_flat_dispatch<__half>(p, m, v, g, p_lp, lp_dtype, n, h, inv_scale_dev, skip_dev, s);
        case DSB_DTYPE_BF16: return adam_flat_dispatch<__nv_bfloat16>(p, m, v, g, p_lp, lp_dtype, n, h, inv_scale_dev, skip_dev, s);
        case DSB_DTYPE_F32: return adam_flat_dispatch<float>(p, m, v, g, p_lp, lp_dtype, n, h, inv_scale_dev, skip_dev, s);
    }
    return DSB_ERR_DTYPE;
}

template <typename P, typename G>
static int adam_multi_launch(const MTList& l, int cnt, long maxn, int state_dtype, const AdamHyper& h, cudaStream_t s) {
    dim3 grid(grid_for(maxn, 512, 4), cnt);
    if (state_dtype == DSB_DTYPE_F32) adam_multi_kernel<P, G, float><<<grid, 512, 0, s>>>(l, h);
    else if (state_dtype == DSB_DTYPE_F16) adam_multi_kernel<P, G, __half><<<grid, 512, 0, s>>>(l, h);
    else adam_multi_kernel<P, G, __nv_bfloat16><<<grid, 512, 0, s>>>(l, h);
    DSB_CHECK_LAUNCH();
}

extern "C" int dsb_adam_multi(int ntensors, void* const* p, void* const* g, void* const* m, void* const* v,
                              const long* numel, int p_dtype, int g_dtype, int state_dtype, float lr, float beta1,
                              float beta2, float eps, float weight_decay, int adamw_mode, int bias_correction,
                              int step, float inv_scale, void* stream) {
    AdamHyper h;
    h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.wd = weight_decay; h.adamw = adamw_mode;
    h.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
    h.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
    h.inv_scale = inv_scale;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    for (int base = 0; base < ntensors; base += MT_MAX) {
        MTList l;
        int cnt = ntensors - base < MT_MAX ? ntensors - base : MT_MAX;
        long maxn = 0;
        for (int i = 0; i < cnt; ++i) {
            l.p[i] = p[base + i]; l.g[i] = g[base + i]; l.m[i] = m[base + i]; l.v[i] = v[base + i];
            l.n[i] = numel[base + i];
            if (l.n[i] > maxn) maxn = l.n[i];
        }
        int rc = DSB_ERR_DTYPE;
#define MT_CASE(PD, PT, GD, GT) \
    if (p_dtype == PD && g_dtype == GD) rc = adam_multi_launch<PT, GT>(l, cnt, maxn, state_dtype, h, s);
        MT_CASE(DSB_DTYPE_F32, float, DSB_DTYPE_F32, float)
        MT_CASE(DSB_DTYPE_F16, __half, DSB_DTYPE_F16, __half)
        MT_CASE(DSB_DTYPE_BF16, __nv_bfloat16, DSB_DTYPE_BF16, __nv_bfloat16)
        MT_CASE(DSB_DTYPE_F32, float, DSB_DTYPE_F16, __half)
        MT_CASE(DSB_DTYPE_F32, float, DSB_DTYPE_BF16, __nv_bfloat16)
#undef MT_CASE
        if (rc) return rc;
    }
    return 0;
}

extern "C" long dsb_lamb_scratch_bytes(long n) { (void)n; return 2L * LAMB_BLOCKS * sizeof(float); }

template <typename P, typename G>
static int lamb_launch(void* p, void* p_copy, int copy_dtype, const void* g, float* m, float* v, long n,
                       const LambHyper& h, float* coeff_out, float* partials, cudaStream_t s) {
    int grid = grid_for(n, 512, 4);
    if (grid > LAMB_BLOCKS) grid = LAMB_BLOCKS;
    lamb_phase1<P, G><<<grid, 512, 0, s>>>(reinterpret_cast<const P*>(p), reinterpret_cast<const G*>(g), m, v, n, h, partials);
    if (p_copy == nullptr || copy_dtype == DSB_DTYPE_F16)
        lamb_phase2<P, __half><<<grid, 512, 0, s>>>(reinterpret_cast<P*>(p), reinterpret_cast<__half*>(p_copy), m, v, n, h, partials, grid, coeff_out);
    else if (copy_dtype == DSB_DTYPE_BF16)
        lamb_phase2<P, __nv_bfloat16><<<grid, 512, 0, s>>>(reinterpret_cast<P*>(p), reinterpret_cast<__nv_bfloat16*>(p_copy), m, v, n, h, partials, grid, coeff_out);
    else
        lamb_phase2<P, float><<<grid, 512, 0, s>>>(reinterpret_cast<P*>(p), reinterpret_cast<float*>(p_copy), m, v, n, h, partials, grid, coeff_out);
    DSB_CHECK_LAUNCH();
}

extern "C" int dsb_lamb_step(void* p, int p_dtype, void* p_copy, int copy_dtype, const void* g, int g_dtype, float* m,
                             float* v, long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                             float max_coeff, float min_coeff, int eps_inside_sqrt, int bias_correction, int step,
                             float grad_scale, float* coeff_out, void* scratch, void* stream) {
    if (n <= 0) return 0;
    if (scratch == nullptr) return DSB_ERR_ARG;
    LambHyper h;
    h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.wd = weight_decay;
    h.max_coeff = max_coeff; h.min_coeff = min_coeff; h.eps_inside = eps_inside_sqrt;
    h.grad_scale = grad_scale == 0.f ? 1.f : grad_scale;
    if (bias_correction) {
        float bc1 = 1.f - powf(beta1, (float)step), bc2 = 1.f - powf(beta2, (float)step);
        h.step_size = lr * sqrtf(bc2) / bc1;
    } else {
        h.step_size = lr;
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    float* partials = reinterpret_cast<float*>(scratch);
#define LAMB_CASE(PD, PT, GD, GT) \
    if (p_dtype == PD && g_dtype == GD) return lamb_launch<PT, GT>(p, p_copy, copy_dtype, g, m, v, n, h, coeff_out, partials, s);
    LAMB_CASE(DSB_DTYPE_F32, float, DSB_DTYPE_F32, float)
    LAMB_CASE(DSB_DTYPE_F32, float, DSB_DTYPE_F16, __half)
    LAMB_CASE(DSB_DTYPE_F32, float, DSB_DTYPE_BF16, __nv_bfloat16)
    LAMB_CASE(DSB_DTYPE_F16, __half, DSB_DTYPE_F16, __half)
    LAMB_CASE(DSB_DTYPE_BF16, __nv_bfloat16, DSB_DTYPE_BF16, __nv_bfloat16)
#undef LAMB_CASE
    return DSB_ERR_DTYPE;
}

template <typename S>
static int cast_from(const void* src, void* dst, int dst_dtype, long n, cudaStream_t s) {
    const int grid = grid_for(n / 4 + 1, 512);
    const S* sp = reinterpret_cast<const S*>(src);
    if (dst_dtype == DSB_DTYPE_F16) cast_kernel<S, __half><<<grid, 512, 0, s>>>(sp, reinterpret_cast<__half*>(dst), n);
    else if (dst_dtype == DSB_DTYPE_BF16) cast_kernel<S, __nv_bfloat16><<<grid, 512, 0, s>>>(sp, reinterpret_cast<__nv_bfloat16*>(dst), n);
    else if (dst_dtype == DSB_DTYPE_F32) cast_kernel<S, float><<<grid, 512, 0, s>>>(sp, reinterpret_cast<float*>(dst), n);
    else return DSB_ERR_DTYPE;
    DSB_CHECK_LAUNCH();
}
extern "C" int dsb_cast(const void* src, int src_dtype, void* dst, int dst_dtype, long n, void* stream) {
    if (n <= 0) return 0;
    if ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) return DSB_ERR_ALIGN;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (src_dtype == DSB_DTYPE_F32) return cast_from<float>(src, dst, dst_dtype, n, s);
    if (src_dtype == DSB_DTYPE_F16) return cast_from<__half>(src, dst, dst_dtype, n, s);
    if (src_dtype == DSB_DTYPE_BF16) return cast_from<__nv_bfloat16>(src, dst, dst_dtype, n, s);
    return DSB_ERR_DTYPE;
}

extern "C" int dsb_scale(void* x, int dtype, long n, float scale, const float* scale_dev, void* stream) {
    if (n <= 0) return 0;
    if (reinterpret_cast<uintptr_t>(x) & 15) return DSB_ERR_ALIGN;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int grid = grid_for(n / 4 + 1, 512);
    if (dtype == DSB_DTYPE_F32) scale_kernel<float><<<grid, 512, 0, s>>>(reinterpret_cast<float*>(x), n, scale, scale_dev);
    else if (dtype == DSB_DTYPE_F16) scale_kernel<__half><<<grid, 512, 0, s>>>(reinterpret_cast<__half*>(x), n, scale, scale_dev);
    else if (dtype == DSB_DTYPE_BF16) scale_kernel<__nv_bfloat16><<<grid, 512, 0, s>>>(reinterpret_cast<__nv_bfloat16*>(x), n, scale, scale_dev);
    else return DSB_ERR_DTYPE;
    DSB_CHECK_LAUNCH();
}

extern "C" long dsb_sumsq_scratch_bytes(void) { return (long)SUMSQ_BLOCKS * (sizeof(float) + sizeof(int)); }

extern "C" int dsb_sumsq(const void* x, int dtype, long n, float* out, int* flag_out, int accumulate, void* scratch,
                         void* stream) {
    if (scratch == nullptr || out == nullptr) return DSB_ERR_ARG;
    if (reinterpret_cast<uintptr_t>(x) & 15) return DSB_ERR_ALIGN;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    float* partials = reinterpret_cast<float*>(scratch);
    int* flags = reinterpret_cast<int*>(partials + SUMSQ_BLOCKS);
    int grid = grid_for(n / 4 + 1, 512, 8);
    if (grid > SUMSQ_BLOCKS) grid = SUMSQ_BLOCKS;
    if (dtype == DSB_DTYPE_F32) sumsq_phase1<float><<<grid, 512, 0, s>>>(reinterpret_cast<const float*>(x), n, partials, flags);
    else if (dtype == DSB_DTYPE_F16) sumsq_phase1<__half><<<grid, 512, 0, s>>>(reinterpret_cast<const __half*>(x), n, partials, flags);
    else if (dtype == DSB_DTYPE_BF16) sumsq_phase1<__nv_bfloat16><<<grid, 512, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(x), n, partials, flags);
    else return DSB_ERR_DTYPE;
    sumsq_phase2<<<1, 512, 0, s>>>(partials, flags, grid, out, flag_out, accumulate);
    DSB_CHECK_LAUNCH();
}
