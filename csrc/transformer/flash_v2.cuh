// Host-side launchers of the second-generation flash-attention kernels (flash_attn_v2.cu); called by the C entry
// points in flash_attn.cu. Tensor maps are built by the caller (make_map_bshd in flash_attn.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace dsb {
namespace fa2 {

// Attention dropout inside the kernels: element (b, h, query, key) of the probability matrix is kept when a 16-bit
// slice of lowbias32(seed_bh + query * C1 + (key >> 1) * C2) is >= thr16 (dropout probability thr16 / 65536), and the
// kept elements are scaled by inv_keep; forward and backward derive the same bits. thr16 == 0: no dropout.
struct Dropout {
    unsigned thr16;
    float inv_keep;
    unsigned seed;
};

// Head dims below 128 (multiples of 8): the tensor maps carry the true head dim, so the part of a 128-wide tile beyond it
// is zero-filled on load and clipped on store by TMA itself - the kernels work on 128 columns, half of them zeros, and no
// padded copy of q / k / v / out exists; only launch_prep and launch_dq_finish need to know D.
// Forward: tq / to with 128-row boxes, tk / tv with 64-row boxes; S must be a multiple of 256.
// sched (optional, device int32): block-sparse schedule, see flash_attn_v2.cu; kpm (optional): additive key mask
// [B, S] fp32, already multiplied by log2(e).
int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& to, float* lse,
               int B, int S, int H, float scale_log2, int causal, int bf16, const int* sched, int sched_hstride,
               int sched_pstride, const float* kpm, Dropout drop, cudaStream_t stream, int D = 128);

// lse2[bh, s] = lse * log2(e), delta_s[bh, s] = scale * sum_d dO * O  (both fp32, [B, H, S]).
int launch_prep(const void* dO, const void* O, const float* lse, float* lse2, float* delta_s, int B, int S, int H,
                const long* s_do, const long* s_o, float scale, int bf16, cudaStream_t stream, int D = 128);

// Backward: tq / tdo with 64-row boxes, tk / tv / tdk / tdv with 128-row boxes, tdq = map of the TRANSPOSED fp32
// accumulator [B, H, D, S] (dims S, D, H, B; box {32, 128, 1, 1}, 128-byte swizzle); S a multiple of 128.
int launch_bwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& tdo,
               const CUtensorMap& tdk, const CUtensorMap& tdv, const CUtensorMap& tdq, const float* lse2,
               const float* delta_s, int B, int S, int H, float scale, float scale_log2, int causal, int bf16,
               const float* rot_cs, const float* rot_sn, int rot, const int* sched, int sched_hstride, int sched_pstride,
               const float* kpm, Dropout drop, cudaStream_t stream, int D = 128);

// dq (16 bit, [B, S, H, 128] with strides s_dq) = transpose of the fp32 accumulator [B, H, 128, S].
// rot > 0: apply the inverse rotary embedding (tables [S, rot / 2] fp32) on the way, likewise for dK in launch_bwd.
int launch_dq_finish(const float* acc, void* dq, int B, int S, int H, const long* s_dq, int bf16, const float* rot_cs,
                     const float* rot_sn, int rot, cudaStream_t stream, int D = 128);

void set_debug_mode(int m);
void set_debug_timers(long long* p);   // 32 int64: phase totals of forward CTA (0, 0)

}  // namespace fa2
}  // namespace dsb
