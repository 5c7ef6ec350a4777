// Shared fused-epilogue math for the tcgen05 GEMM kernels: operates on one 64-column chunk of one accumulator row.
#pragma once
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"

namespace dsb {

struct EpiParams {
    int M, N;
    int epilogue;
    float alpha;
    const void* bias;
    const void* aux;
    long ldaux, stride_aux;
    void* aux_out;
    long ldaux_out, stride_aux_out;
};

// v[64]: fp32 accumulator values of row `row`, columns [col0, col0 + 64) of batch b (already scaled by alpha).
template <typename T>
DSB_DEVICE void epilogue_apply(float (&v)[64], const EpiParams& p, int b, int row, bool row_ok, int col0) {
    const int epi = p.epilogue;
    if (p.bias != nullptr && epi != DSB_EPI_DGELU) {
        const T* bp = reinterpret_cast<const T*>(p.bias) + col0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (col0 + j * 8 < p.N) {
                uint4 u = __ldg(reinterpret_cast<const uint4*>(bp) + j);
                float f[8];
                unpack8<T>(u, f);
#pragma unroll
                for (int t = 0; t < 8; ++t) v[j * 8 + t] += f[t];
            }
        }
    }
    if (epi == DSB_EPI_BIAS_GELU) {
        if (p.aux_out != nullptr && row_ok) {
            T* op = reinterpret_cast<T*>(p.aux_out) + b * p.stride_aux_out + (long)row * p.ldaux_out + col0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (col0 + j * 8 < p.N) {
                    float f[8];
#pragma unroll
                    for (int t = 0; t < 8; ++t) f[t] = v[j * 8 + t];
                    *(reinterpret_cast<uint4*>(op) + j) = pack8<T>(f);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] = gelu_tanh(v[i]);
    } else if (epi == DSB_EPI_BIAS_RESIDUAL || epi == DSB_EPI_DGELU) {
        if (row_ok) {
            const T* ap = reinterpret_cast<const T*>(p.aux) + b * p.stride_aux + (long)row * p.ldaux + col0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (col0 + j * 8 < p.N) {
                    float f[8];
                    unpack8<T>(*(reinterpret_cast<const uint4*>(ap) + j), f);
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        if (epi == DSB_EPI_DGELU) v[j * 8 + t] *= gelu_tanh_grad(f[t]);
                        else v[j * 8 + t] += f[t];
                    }
                }
            }
        }
    }
}

// Write the 32 x 64 chunk owned by a warp into its 128-byte-swizzled staging buffer (row = lane).
template <typename T>
DSB_DEVICE void epilogue_stage(const float (&v)[64], uint8_t* sbuf, int lane) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float f[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) f[t] = v[j * 8 + t];
        *reinterpret_cast<uint4*>(sbuf + lane * 128 + ((j ^ (lane & 7)) << 4)) = pack8<T>(f);
    }
}

}  // namespace dsb
