// Error-compensated 1-bit all-reduce as ONE cooperative sm_100a kernel over NVLink peer memory (north-star item 3).
//
// Math (reference deepspeed/runtime/comm/nccl.py:47-186), W = data-parallel world, n = padded numel,
// c = n / W (server chunk):
//   worker : x = m + e_w ; s_w = |x|_2 / sqrt(n) ; e_w = x - s_w * sgn(x) ; bits = (x >= 0)
//   exchange: chunk r of the packed bits + s_w go to rank r            (all-to-all + all-gather in the reference)
//   server : y = (1/W) sum_w s_w * sgn_w + e_s ; s_s = |y|_2 / sqrt(c) ; e_s = y - s_s * sgn(y) ; bits = (y >= 0)
//   exchange: server bits + s_s go to every rank                        (2 x all-gather in the reference)
//   result : m[k*c + j] = s_s[k] * sgn_k[j]
//
// Instead of cupy packbits + all_to_all_single + 3 all_gathers, the kernel packs signs with warp ballots and writes
// the 32-bit words and the scales straight into the destination ranks' receive buffers (peer stores into a symmetric
// heap), separated by two in-kernel cross-GPU flag barriers; the 2-norms are folded deterministically between
// grid-wide syncs. One launch per tensor, no host synchronisation, no intermediate tensors.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace dsb {

constexpr int OB_MAX_RANKS = 16;
struct ObPtrs {
    uint32_t* recv_signs[OB_MAX_RANKS];   // per rank: W * (c/32) words, slot w = chunk sent by worker w
    float* recv_scales[OB_MAX_RANKS];     // per rank: W floats
    uint32_t* gath_signs[OB_MAX_RANKS];   // per rank: n/32 words, slot k = server k's chunk
    float* gath_scales[OB_MAX_RANKS];     // per rank: W floats
    uint32_t* flags[OB_MAX_RANKS];
};

DSB_DEVICE void ob_barrier(const ObPtrs& P, int W, int me, int slot, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x < W) {
        __threadfence_system();
        st_release_sys(P.flags[threadIdx.x] + (2 + slot) * 16 + me, epoch);
        const uint32_t* mine = P.flags[me] + (2 + slot) * 16 + threadIdx.x;
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {}
    }
    __syncthreads();
}

DSB_DEVICE float fold_partials(const float* partials, int nparts, float* sh) {
    float t = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) t += partials[i];
    return block_sum(t, sh);
}

// 4 sign bits per lane (elements 4q..4q+3 of a 32-element word, q = lane % 8) -> the 32-bit word, held by every lane
// of the 8-lane group afterwards.
DSB_DEVICE uint32_t pack_nibbles(uint32_t nib, int lane) {
    uint32_t v = nib << ((lane & 7) * 4);
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    v |= __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
}
DSB_DEVICE float sgn_sel(uint32_t bits, int t, float s) { return ((bits >> t) & 1u) ? s : -s; }

// All element loops run on float4 (16-byte accesses, two independent iterations in flight per thread in the pure
// streaming phases): the first version used 4-byte accesses and reached ~1/6 of HBM bandwidth
// (profiles/onebit_bench_w*.json history).
__global__ void __launch_bounds__(512) onebit_allreduce_kernel(float* __restrict__ m, float* __restrict__ e_w,
                                                               float* __restrict__ e_s, long n, int W, int me,
                                                               ObPtrs P, float* __restrict__ partials, uint32_t epoch) {
    cg::grid_group grid = cg::this_grid();
    __shared__ float sh[33];
    __shared__ float ssc[OB_MAX_RANKS];
    const long c = n / W;
    const long cwords = c >> 5;
    const long n4 = n >> 2, c4 = c >> 2;
    const long gtid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long gstride = (long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* ew4 = reinterpret_cast<float4*>(e_w);
    float4* es4 = reinterpret_cast<float4*>(e_s);

    // ---- phase A: x = m + e_w (kept in m), partial sum of squares
    float acc = 0.f;
    {
        long i = gtid;
        for (; i + gstride < n4; i += 2 * gstride) {
            float4 a0 = m4[i], b0 = ew4[i], a1 = m4[i + gstride], b1 = ew4[i + gstride];
            a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
            a1.x += b1.x; a1.y += b1.y; a1.z += b1.z; a1.w += b1.w;
            m4[i] = a0;
            m4[i + gstride] = a1;
            acc += a0.x * a0.x + a0.y * a0.y + a0.z * a0.z + a0.w * a0.w;
            acc += a1.x * a1.x + a1.y * a1.y + a1.z * a1.z + a1.w * a1.w;
        }
        if (i < n4) {
            float4 a0 = m4[i], b0 = ew4[i];
            a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
            m4[i] = a0;
            acc += a0.x * a0.x + a0.y * a0.y + a0.z * a0.z + a0.w * a0.w;
        }
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
    grid.sync();
    const float s_w = sqrtf(fold_partials(partials, gridDim.x, sh) / (float)n);
    grid.sync();   // partials are reused below

    // ---- phase B: worker error, sign packing, scatter chunk r to rank r (whole warps stay together: n % 32 == 0)
    for (long base = gtid - lane; base < n4; base += gstride) {
        const long i = base + lane;
        const bool valid = i < n4;
        uint32_t nib = 0;
        if (valid) {
            const float4 x = m4[i];
            nib = (x.x >= 0.f ? 1u : 0u) | (x.y >= 0.f ? 2u : 0u) | (x.z >= 0.f ? 4u : 0u) | (x.w >= 0.f ? 8u : 0u);
            float4 e;
            e.x = x.x - sgn_sel(nib, 0, s_w); e.y = x.y - sgn_sel(nib, 1, s_w);
            e.z = x.z - sgn_sel(nib, 2, s_w); e.w = x.w - sgn_sel(nib, 3, s_w);
            ew4[i] = e;
        }
        const uint32_t word = pack_nibbles(nib, lane);
        if (valid && (lane & 7) == 0) {
            const long w = i >> 3;
            const int dst = (int)(w / cwords);
            P.recv_signs[dst][(long)me * cwords + (w - (long)dst * cwords)] = word;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < W) P.recv_scales[threadIdx.x][me] = s_w;
    __threadfence_system();
    grid.sync();
    if (blockIdx.x == 0) ob_barrier(P, W, me, 0, epoch);
    grid.sync();

    // ---- phase C: server average of my chunk + server error feedback (y kept in m[me*c ..])
    float4* y4 = reinterpret_cast<float4*>(m + (long)me * c);
    const uint32_t* rs = P.recv_signs[me];
    if (threadIdx.x < W) ssc[threadIdx.x] = __ldcg(P.recv_scales[me] + threadIdx.x);
    __syncthreads();
    const float inv_w = 1.f / (float)W;
    acc = 0.f;
    for (long j = gtid; j < c4; j += gstride) {
        float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
        const int sh4 = (int)(j & 7) * 4;
        for (int w = 0; w < W; ++w) {
            const uint32_t bits = __ldcg(rs + (long)w * cwords + (j >> 3)) >> sh4;
            const float s = ssc[w];
            y.x += sgn_sel(bits, 0, s); y.y += sgn_sel(bits, 1, s);
            y.z += sgn_sel(bits, 2, s); y.w += sgn_sel(bits, 3, s);
        }
        const float4 e = es4[j];
        y.x = y.x * inv_w + e.x; y.y = y.y * inv_w + e.y; y.z = y.z * inv_w + e.z; y.w = y.w * inv_w + e.w;
        y4[j] = y;
        acc += y.x * y.x + y.y * y.y + y.z * y.z + y.w * y.w;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
    grid.sync();
    const float s_s = sqrtf(fold_partials(partials, gridDim.x, sh) / (float)c);

    // ---- phase D: server error, pack, broadcast my chunk's signs + scale to every rank
    for (long base = gtid - lane; base < c4; base += gstride) {
        const long j = base + lane;
        const bool valid = j < c4;
        uint32_t nib = 0;
        if (valid) {
            const float4 y = y4[j];
            nib = (y.x >= 0.f ? 1u : 0u) | (y.y >= 0.f ? 2u : 0u) | (y.z >= 0.f ? 4u : 0u) | (y.w >= 0.f ? 8u : 0u);
            float4 e;
            e.x = y.x - sgn_sel(nib, 0, s_s); e.y = y.y - sgn_sel(nib, 1, s_s);
            e.z = y.z - sgn_sel(nib, 2, s_s); e.w = y.w - sgn_sel(nib, 3, s_s);
            es4[j] = e;
        }
        const uint32_t word = pack_nibbles(nib, lane);
        if (valid && (lane & 7) == 0) {
            const long w = (long)me * cwords + (j >> 3);
            for (int r = 0; r < W; ++r) P.gath_signs[(me + r) % W][w] = word;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < W) P.gath_scales[threadIdx.x][me] = s_s;
    __threadfence_system();
    grid.sync();
    if (blockIdx.x == 0) ob_barrier(P, W, me, 1, epoch);
    grid.sync();

    // ---- phase E: decompress everything into m
    const uint32_t* gs = P.gath_signs[me];
    if (threadIdx.x < W) ssc[threadIdx.x] = __ldcg(P.gath_scales[me] + threadIdx.x);
    __syncthreads();
    {
        long i = gtid;
        for (; i + gstride < n4; i += 2 * gstride) {
            const long i1 = i + gstride;
            const uint32_t b0 = __ldcg(gs + (i >> 3)) >> ((int)(i & 7) * 4);
            const uint32_t b1 = __ldcg(gs + (i1 >> 3)) >> ((int)(i1 & 7) * 4);
            const float s0 = ssc[(int)(i / c4)], s1 = ssc[(int)(i1 / c4)];
            m4[i] = make_float4(sgn_sel(b0, 0, s0), sgn_sel(b0, 1, s0), sgn_sel(b0, 2, s0), sgn_sel(b0, 3, s0));
            m4[i1] = make_float4(sgn_sel(b1, 0, s1), sgn_sel(b1, 1, s1), sgn_sel(b1, 2, s1), sgn_sel(b1, 3, s1));
        }
        if (i < n4) {
            const uint32_t b0 = __ldcg(gs + (i >> 3)) >> ((int)(i & 7) * 4);
            const float s0 = ssc[(int)(i / c4)];
            m4[i] = make_float4(sgn_sel(b0, 0, s0), sgn_sel(b0, 1, s0), sgn_sel(b0, 2, s0), sgn_sel(b0, 3, s0));
        }
    }
}

}  // namespace dsb

using namespace dsb;

// Symmetric-heap layout per rank for one tensor size (bytes from `base`): see dsb_onebit_heap_bytes.
extern "C" long dsb_onebit_heap_bytes(long n, int W) {
    const long words = n / 32;
    return 2 * (words * 4 + 256) + 2 * 256;
}

extern "C" int dsb_onebit_allreduce(float* m, float* e_w, float* e_s, long n, int W, int me, void* const* heap_bases,
                                    long region_off, void* const* flag_bases, void* scratch, unsigned int epoch,
                                    void* stream) {
    if (W < 1 || W > OB_MAX_RANKS || n <= 0 || (n % (32L * W)) || !scratch) return DSB_ERR_ARG;
    const long words = n / 32;
    const long sign_bytes = words * 4 + 256;
    ObPtrs P;
    for (int r = 0; r < W; ++r) {
        char* b = reinterpret_cast<char*>(heap_bases[r]) + region_off;
        P.recv_signs[r] = reinterpret_cast<uint32_t*>(b);
        P.gath_signs[r] = reinterpret_cast<uint32_t*>(b + sign_bytes);
        P.recv_scales[r] = reinterpret_cast<float*>(b + 2 * sign_bytes);
        P.gath_scales[r] = reinterpret_cast<float*>(b + 2 * sign_bytes + 256);
        P.flags[r] = reinterpret_cast<uint32_t*>(flag_bases[r]);
    }
    float* partials = reinterpret_cast<float*>(scratch);
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, onebit_allreduce_kernel, 512, 0);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 2) per_sm = 2;
    void* args[] = {&m, &e_w, &e_s, &n, &W, &me, &P, &partials, &epoch};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)onebit_allreduce_kernel, dim3(sms * per_sm), dim3(512),
                                                args, 0, reinterpret_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : -(int)e;
}
