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

__global__ void __launch_bounds__(512) onebit_allreduce_kernel(float* __restrict__ m, float* __restrict__ e_w,
                                                               float* __restrict__ e_s, long n, int W, int me,
                                                               ObPtrs P, float* __restrict__ partials, uint32_t epoch) {
    cg::grid_group grid = cg::this_grid();
    __shared__ float sh[33];
    const long c = n / W;
    const long cwords = c >> 5;
    const long gtid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long gstride = (long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;

    // ---- phase A: x = m + e_w (kept in m), partial sum of squares
    float acc = 0.f;
    for (long i = gtid; i < n; i += gstride) {
        const float x = m[i] + e_w[i];
        m[i] = x;
        acc += x * x;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
    grid.sync();
    const float s_w = sqrtf(fold_partials(partials, gridDim.x, sh) / (float)n);
    grid.sync();   // partials are reused below

    // ---- phase B: worker error, sign packing (one ballot = 32 elements), scatter chunk r to rank r
    for (long i = gtid; i < n; i += gstride) {   // n is a multiple of 32 * W: whole warps stay together
        const float x = m[i];
        const bool pos = x >= 0.f;
        e_w[i] = x - (pos ? s_w : -s_w);
        const uint32_t word = __ballot_sync(0xffffffffu, pos);
        if (lane == 0) {
            const long w = i >> 5;
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
    float* ybuf = m + (long)me * c;
    const uint32_t* rs = P.recv_signs[me];
    const float* rsc = P.recv_scales[me];
    acc = 0.f;
    for (long j = gtid; j < c; j += gstride) {
        float y = 0.f;
        for (int w = 0; w < W; ++w) {
            const uint32_t word = __ldcg(rs + (long)w * cwords + (j >> 5));
            const float s = __ldcg(rsc + w);
            y += ((word >> (j & 31)) & 1u) ? s : -s;
        }
        y = y / (float)W + e_s[j];
        ybuf[j] = y;
        acc += y * y;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
    grid.sync();
    const float s_s = sqrtf(fold_partials(partials, gridDim.x, sh) / (float)c);

    // ---- phase D: server error, pack, broadcast my chunk's signs + scale to every rank
    for (long j = gtid; j < c; j += gstride) {
        const float y = ybuf[j];
        const bool pos = y >= 0.f;
        e_s[j] = y - (pos ? s_s : -s_s);
        const uint32_t word = __ballot_sync(0xffffffffu, pos);
        if (lane == 0) {
            const long w = (long)me * cwords + (j >> 5);
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
    const float* gsc = P.gath_scales[me];
    for (long i = gtid; i < n; i += gstride) {
        const uint32_t word = __ldcg(gs + (i >> 5));
        const float s = __ldcg(gsc + (int)(i / c));
        m[i] = ((word >> (i & 31)) & 1u) ? s : -s;
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
