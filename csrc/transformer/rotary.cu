// Out-of-place, vectorised rotary embedding (GPT-NeoX half-split) with a precomputed cos/sin table.
// Reads q/k straight out of the fused QKV activation (arbitrary batch/seq/head strides, unit last stride) and writes
// rotated q/k into separately strided outputs (e.g. contiguous [B,S,H,D] buffers in forward, slices of the dQKV
// buffer in backward), so no clone / slice-backward / zero-fill passes are needed around attention.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"

namespace dsb {

__global__ void rotary_table_kernel(float* __restrict__ cs, float* __restrict__ sn, int S, int half, float log2_base,
                                    int rot) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= S * half) return;
    const int s = i / half, j = i % half;
    const float inv_freq = exp2f(-log2_base * (2.f * j) / rot);
    float a, b;
    sincosf(s * inv_freq, &a, &b);
    sn[i] = a;
    cs[i] = b;
}

struct RotStrides {
    long b, s, h;
};

// One CTA iteration = one (batch, position) token: b and s come from one division per token, the (head, vector) split
// inside the token is 32-bit arithmetic on small numbers. (The first version decomposed a flat 64-bit index with four
// long divisions per 16-byte vector and was instruction-bound: ncu showed 31 % DRAM throughput at 46 % SM throughput.)
template <typename T>
__global__ void __launch_bounds__(128) rotary_oop_kernel(const T* __restrict__ qi, const T* __restrict__ ki,
                                                         T* __restrict__ qo, T* __restrict__ ko, int B, int S, int H,
                                                         int D, int rot, RotStrides si, RotStrides so,
                                                         const float* __restrict__ cs, const float* __restrict__ sn,
                                                         float sign, const T* __restrict__ vi, T* __restrict__ vo,
                                                         RotStrides sv) {
    const int half = rot >> 1;
    const int hv = half >> 3;                  // 8-element vectors per rotated half
    const int pv = (D - rot) >> 3;             // pass-through vectors
    const int vv = vi != nullptr ? (D >> 3) : 0;   // optional third tensor copied verbatim (v / dv) in the same launch
    const int per_head = hv + pv;              // the verbatim copy runs as a second, divergence-free loop
    const int per_token = H * per_head;
    const int v_per_token = H * vv;
    const long tokens = (long)B * S;
    for (long tk = blockIdx.x; tk < tokens; tk += gridDim.x) {
        const int b = (int)(tk / S);
        const int s = (int)(tk - (long)b * S);
        const long ti = b * si.b + s * si.s;
        const long to = b * so.b + s * so.s;
        const float* cs_row = cs + (long)s * half;
        const float* sn_row = sn + (long)s * half;
        for (int e = threadIdx.x; e < per_token; e += blockDim.x) {
            const int h = e / per_head;
            const int v = e - h * per_head;
            const long oi = ti + h * si.h;
            const long oo = to + h * so.h;
            if (v < hv) {
                const int j = v * 8;
                // issue every load of this vector pair before the math
                const uint4 q0 = *reinterpret_cast<const uint4*>(qi + oi + j);
                const uint4 q1 = *reinterpret_cast<const uint4*>(qi + oi + j + half);
                const uint4 k0 = *reinterpret_cast<const uint4*>(ki + oi + j);
                const uint4 k1 = *reinterpret_cast<const uint4*>(ki + oi + j + half);
                float c8[8], s8[8];
                *reinterpret_cast<float4*>(c8) = __ldg(reinterpret_cast<const float4*>(cs_row + j));
                *reinterpret_cast<float4*>(c8 + 4) = __ldg(reinterpret_cast<const float4*>(cs_row + j + 4));
                *reinterpret_cast<float4*>(s8) = __ldg(reinterpret_cast<const float4*>(sn_row + j));
                *reinterpret_cast<float4*>(s8 + 4) = __ldg(reinterpret_cast<const float4*>(sn_row + j + 4));
#pragma unroll
                for (int w = 0; w < 2; ++w) {
                    T* out = w == 0 ? qo : ko;
                    float x0[8], x1[8], y0[8], y1[8];
                    unpack8<T>(w == 0 ? q0 : k0, x0);
                    unpack8<T>(w == 0 ? q1 : k1, x1);
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        const float sg = s8[t] * sign;
                        y0[t] = x0[t] * c8[t] - x1[t] * sg;
                        y1[t] = x1[t] * c8[t] + x0[t] * sg;
                    }
                    *reinterpret_cast<uint4*>(out + oo + j) = pack8<T>(y0);
                    *reinterpret_cast<uint4*>(out + oo + j + half) = pack8<T>(y1);
                }
            } else {
                const int j = rot + (v - hv) * 8;
                const uint4 a = *reinterpret_cast<const uint4*>(qi + oi + j);
                const uint4 c = *reinterpret_cast<const uint4*>(ki + oi + j);
                *reinterpret_cast<uint4*>(qo + oo + j) = a;
                *reinterpret_cast<uint4*>(ko + oo + j) = c;
            }
        }
        const long tv = b * sv.b + s * sv.s;
        for (int e = threadIdx.x; e < v_per_token; e += blockDim.x) {
            const int h = e / vv;
            const int j = (e - h * vv) * 8;
            *reinterpret_cast<uint4*>(vo + to + h * so.h + j) = *reinterpret_cast<const uint4*>(vi + tv + h * sv.h + j);
        }
    }
}

}  // namespace dsb

using namespace dsb;

extern "C" int dsb_rotary_table(float* cos_out, float* sin_out, int S, int rot_dim, float base, void* stream) {
    const int half = rot_dim / 2;
    const int n = S * half;
    if (n <= 0) return 0;
    rotary_table_kernel<<<(n + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(cos_out, sin_out, S, half,
                                                                                             log2f(base), rot_dim);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_rotary_oop_v(const void* q_in, const void* k_in, void* q_out, void* k_out, int B, int S, int H,
                                int D, int rot_dim, long in_sb, long in_ss, long in_sh, long out_sb, long out_ss,
                                long out_sh, const float* cos_tab, const float* sin_tab, int inverse, int dtype,
                                const void* v_in, void* v_out, long v_sb, long v_ss, long v_sh, void* stream);

extern "C" int dsb_rotary_oop(const void* q_in, const void* k_in, void* q_out, void* k_out, int B, int S, int H, int D,
                              int rot_dim, long in_sb, long in_ss, long in_sh, long out_sb, long out_ss, long out_sh,
                              const float* cos_tab, const float* sin_tab, int inverse, int dtype, void* stream) {
    return dsb_rotary_oop_v(q_in, k_in, q_out, k_out, B, S, H, D, rot_dim, in_sb, in_ss, in_sh, out_sb, out_ss, out_sh,
                            cos_tab, sin_tab, inverse, dtype, nullptr, nullptr, 0, 0, 0, stream);
}

// Same, plus an optional third tensor v_in [B,S,H,D] (strides v_s*) copied verbatim to v_out, which shares q_out's
// batch/seq/head strides (forward: nothing; backward: dV into its slice of the dQKV buffer).
extern "C" int dsb_rotary_oop_v(const void* q_in, const void* k_in, void* q_out, void* k_out, int B, int S, int H,
                                int D, int rot_dim, long in_sb, long in_ss, long in_sh, long out_sb, long out_ss,
                                long out_sh, const float* cos_tab, const float* sin_tab, int inverse, int dtype,
                                const void* v_in, void* v_out, long v_sb, long v_ss, long v_sh, void* stream) {
    if (rot_dim % 16 || rot_dim > D || D % 8 || (in_sb | in_ss | in_sh | out_sb | out_ss | out_sh) % 8)
        return DSB_ERR_SIZE;
    if (v_in && (!v_out || (v_sb | v_ss | v_sh) % 8)) return DSB_ERR_SIZE;
    const long total = (long)B * S * H * ((rot_dim / 2 + (D - rot_dim) + (v_in ? D : 0)) / 8);
    if (total <= 0) return 0;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long blocks = (long)B * S;                       // one token per CTA iteration
    if (blocks > (long)sms * 16) blocks = (long)sms * 16;
    RotStrides si{in_sb, in_ss, in_sh}, so{out_sb, out_ss, out_sh}, sv{v_sb, v_ss, v_sh};
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const float sign = inverse ? -1.f : 1.f;
    if (dtype == DSB_DTYPE_F16)
        rotary_oop_kernel<__half><<<(unsigned)blocks, 128, 0, s>>>((const __half*)q_in, (const __half*)k_in, (__half*)q_out, (__half*)k_out, B, S, H, D, rot_dim, si, so, cos_tab, sin_tab, sign, (const __half*)v_in, (__half*)v_out, sv);
    else if (dtype == DSB_DTYPE_BF16)
        rotary_oop_kernel<__nv_bfloat16><<<(unsigned)blocks, 128, 0, s>>>((const __nv_bfloat16*)q_in, (const __nv_bfloat16*)k_in, (__nv_bfloat16*)q_out, (__nv_bfloat16*)k_out, B, S, H, D, rot_dim, si, so, cos_tab, sin_tab, sign, (const __nv_bfloat16*)v_in, (__nv_bfloat16*)v_out, sv);
    else
        return DSB_ERR_DTYPE;
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}
