// The ZeRO-1/2 data-parallel step of one gradient bucket as ONE cooperative kernel over the symmetric heap:
//
//   phase 1 (REDUCE)   cross-GPU barrier; this rank's slice of the bucket is reduced over all ranks - with NVLS one
//                      multimem.ld_reduce per 16 bytes (the switch adds every rank's copy in fp32 and returns the sum),
//                      otherwise R peer loads summed in registers - averaged, kept as fp32 in a local scratch, and its
//                      sum of squares + non-finite flag are folded over the grid;
//   phase X (EXCHANGE) CTA 0 trades (sum of squares, overflow) with every peer through mailboxes in the heap's flag
//                      block and derives the global norm, the combined unscale / clip factor and the skip flag on the
//                      device - no NCCL scalar all-reduce, no host round trip. Doubles as the barrier that tells every
//                      rank its gradient bucket has been read by all peers;
//   phase 2 (UPDATE)   Adam / AdamW on the fp32 master + moments and the cast; the low-precision slice is staged in
//                      shared memory and pushed into EVERY rank's parameter arena (the all-gather) by bulk copies
//                      (cp.async.bulk global <- shared on plain unicast peer addresses), and the local gradient
//                      bucket is cleared by bulk copies of a zero block (replaces the arena memset), so the SM's
//                      load / store path carries only the fp32 state; closing barrier.
//
// Why NVLS is opt-in for the reduce and unused for the gather: a reduce-SCATTER through the switch makes every rank
// send its whole arena out (the switch also fetches the reducing rank's own copy over NVLink), W / (W - 1) times the
// egress of peer loads, and a multicast store delivers W slices to every rank instead of W - 1. Measured: reduce
// phase 2.07 ms (multimem) vs 1.13 ms (peer loads) at W = 2, 1.98 vs 1.70 ms at W = 4 (profiles/fused_zero_bench_*).
// In-switch reduction pays for all-reduce-shaped exchanges, not for ZeRO's scatter / gather halves.
//
// `phases` selects what a launch does: REDUCE|EXCHANGE|UPDATE = the whole step (one bucket, no model-parallel group);
// REDUCE alone accumulates the local norm terms (several buckets, or a norm that must also be reduced over a
// model-parallel / pipeline group by the host) and UPDATE alone applies a factor computed elsewhere.
// Reference: engine.allreduce_bucket + zero/stage1.reduce_scatter_gradients + FP16 unscale / clip + stage1.step
// all_gather (deepspeed/runtime/zero/stage1.py:629-783, runtime/fp16/fused_optimizer.py:242-350).
#include "xgpu.cuh"
#include "../../include/dsb200.h"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace dsb {

enum { PH_REDUCE = 1, PH_UPDATE = 2, PH_EXCHANGE = 4 };

struct ZeroStepArgs {
    PeerPtrs grads, params;       // unicast base of every rank's gradient / parameter arena
    FlagPtrs flags;
    const char* mc_grads;         // multicast base of the gradient arena (nullptr: peer loads)
    char* mc_params;              // multicast base of the parameter arena (nullptr: peer stores)
    int R, me, phases, zero_grads;
    long slice_off;               // first element of this rank's slice inside the arenas
    long bucket_off, bucket_n;    // the whole bucket (all ranks' slices) inside the local gradient arena
    long n;                       // elements in the slice (multiple of 8)
    float avg;                    // 1 / R
    void* red;                    // [n] reduced gradient in the arena dtype (local scratch); nullptr: in place in
                                  // this rank's own slice of its gradient arena
    float *p32, *m, *v;           // fp32 master and moments of the slice
    float lr, beta1, beta2, eps, wd, bc1, bc2;
    int adamw;
    float loss_scale, clip;       // EXCHANGE: combined = loss_scale * max(1, (norm / loss_scale + 1e-6) / clip)
    float* sumsq;                 // [2] running local {sum of squares, non-finite count} (REDUCE adds to it)
    float* inv_scale;             // [1] 1 / combined (EXCHANGE writes, UPDATE reads)
    int* skip;                    // [1] overflow -> no update (EXCHANGE writes, UPDATE reads)
    float* host_out;              // [2] {global norm (scaled), overflow} for the host's bookkeeping (EXCHANGE writes)
    float* partials;              // 2 * gridDim.x floats
    uint32_t epoch;
};

// 1-D bulk copy shared -> global (any mapped address, local or a peer's), tracked by the thread's bulk group
DSB_DEVICE void bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
                 "r"(bytes)
                 : "memory");
}

DSB_DEVICE void stamp(const ZeroStepArgs& a, int slot) {
    // phase timestamps (ns) for tools / benches: CTA 0 only, 8 slots behind the partial sums
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        reinterpret_cast<unsigned long long*>(a.partials + 640)[slot] = t;
    }
}

template <typename T, bool MC>
__global__ void __launch_bounds__(512, 2) fused_zero_step_kernel(const __grid_constant__ ZeroStepArgs a) {
    cg::grid_group grid = cg::this_grid();
    stamp(a, 0);
    __shared__ float sh[33];
    const long stride = (long)gridDim.x * blockDim.x;
    const long tid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int R = a.R, me = a.me;

    if (a.phases & PH_REDUCE) {
        // every peer has finished writing its gradient arena (its kernel is stream-ordered after its backward)
        if (blockIdx.x == 0) xgpu_barrier(a.flags, R, me, 0, a.epoch);
        grid.sync();
        stamp(a, 1);
        const long nvec = a.n >> 3;
        float acc = 0.f;
        int bad = 0;
        // the averaged slice goes back, rounded to the arena dtype, into this rank's own slice of its gradient arena
        // (what an in-place NCCL reduce-scatter leaves there); no peer ever reads that slice
        T* own = a.red ? reinterpret_cast<T*>(a.red) : reinterpret_cast<T*>(a.grads.p[me]) + a.slice_off;
        auto finish = [&](long i, float (&s)[8]) {
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] *= a.avg;
            const uint4 packed = pack8<T>(s);
            unpack8<T>(packed, s);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                acc += s[k] * s[k];
                bad |= !isfinite(s[k]);
            }
            *reinterpret_cast<uint4*>(own + i * 8) = packed;
        };
        if (MC) {
            // four independent in-switch reductions in flight per thread
            const char* base = a.mc_grads + a.slice_off * (long)sizeof(T);
            for (long i0 = tid; i0 < nvec; i0 += 4 * stride) {
                uint4 raw[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const long i = i0 + u * stride;
                    if (i < nvec) raw[u] = multimem_ld_reduce_add8<T>(base + i * 16);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const long i = i0 + u * stride;
                    if (i < nvec) {
                        float s[8];
                        unpack8<T>(raw[u], s);
                        finish(i, s);
                    }
                }
            }
        } else {
            for (long i = tid; i < nvec; i += stride) {
                float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int j0 = 0; j0 < R; j0 += 4) {
                    uint4 raw[4];
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        if (j0 + jj < R) {
                            const int r = (me + j0 + jj) % R;
                            raw[jj] = ld_na_v4(reinterpret_cast<const T*>(a.grads.p[r]) + a.slice_off + i * 8);
                        }
                    }
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        if (j0 + jj < R) {
                            float f[8];
                            unpack8<T>(raw[jj], f);
#pragma unroll
                            for (int k = 0; k < 8; ++k) s[k] += f[k];
                        }
                    }
                }
                finish(i, s);
            }
        }
        acc = block_sum(acc, sh);
        bad = __syncthreads_or(bad);
        if (threadIdx.x == 0) {
            a.partials[blockIdx.x] = acc;
            a.partials[gridDim.x + blockIdx.x] = bad ? 1.f : 0.f;
        }
        grid.sync();
        stamp(a, 2);
        if (blockIdx.x == 0) {
            float t = 0.f, b = 0.f;
            for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
                t += a.partials[i];
                b += a.partials[gridDim.x + i];
            }
            t = block_sum(t, sh);
            b = block_sum(b, sh);
            if (threadIdx.x == 0) {
                a.sumsq[0] += t;
                a.sumsq[1] += (b > 0.f || !isfinite(t)) ? 1.f : 0.f;
            }
            __syncthreads();
        }
    }

    if (a.phases & PH_EXCHANGE) {
        if (blockIdx.x == 0) {
            __shared__ float tot[2];
            // a non-finite local sum must not poison the peers' arithmetic: it travels as the overflow count
            const float mine = a.sumsq[0], flag = a.sumsq[1];
            xgpu_sum2(a.flags, R, me, a.epoch, isfinite(mine) ? mine : 0.f, flag, tot);
            if (threadIdx.x == 0) {
                const float norm = sqrtf(fmaxf(tot[0], 0.f));
                const bool overflow = tot[1] > 0.f || !isfinite(norm);
                float combined = a.loss_scale;
                if (a.clip > 0.f) combined *= fmaxf((norm / a.loss_scale + 1e-6f) / a.clip, 1.f);
                *a.inv_scale = 1.f / combined;
                *a.skip = overflow ? 1 : 0;
                a.host_out[0] = norm;
                a.host_out[1] = overflow ? 1.f : 0.f;
                __threadfence();
            }
        }
        grid.sync();
    }
    stamp(a, 3);

    if (a.phases & PH_UPDATE) {
        // Stores that leave the SM's load/store path: the low-precision slice is staged in shared memory and pushed
        // to every rank's parameter arena (the all-gather) by bulk copies (cp.async.bulk global <- shared; peer
        // addresses are plain unicast mappings), and the gradient bucket is cleared by bulk copies of a zero block.
        // Loads / stores of the fp32 state then never queue behind NVLink.
        constexpr int TILE = 512;                              // vectors (of 8 elements) per CTA iteration
        __shared__ alignas(128) uint4 stage[2][TILE];
        __shared__ alignas(128) uint4 zeros[TILE];
        zeros[threadIdx.x] = make_uint4(0, 0, 0, 0);
        fence_proxy_async_smem();
        __syncthreads();
        const bool skip = *reinterpret_cast<const volatile int*>(a.skip) != 0;
        const float inv = *reinterpret_cast<const volatile float*>(a.inv_scale);
        const long nvec = a.n >> 3;
        const long ntiles = (nvec + TILE - 1) / TILE;
        // Clearing the local bucket (every peer has read it: EXCHANGE, or the host's all-reduce, follows REDUCE). This
        // rank's own slice holds the reduced gradient until the Adam loop has read it, so it is cleared tile by tile
        // behind that loop; the rest of the bucket - vectors [0, own_v0) and [own_v0 + nvec, bucket) - is one virtual
        // range of `zvec` vectors cut into tiles and spread over the rounds of all CTAs.
        char* gz = reinterpret_cast<char*>(reinterpret_cast<T*>(a.grads.p[me]) + a.bucket_off);
        T* own = reinterpret_cast<T*>(a.grads.p[me]) + a.slice_off;
        const T* redsrc = a.red ? reinterpret_cast<const T*>(a.red) : own;
        const long own_v0 = (a.slice_off - a.bucket_off) >> 3;
        const long zvec = a.zero_grads ? (a.bucket_n >> 3) - nvec : 0;
        const long ztiles = (zvec + TILE - 1) / TILE;
        const long z_per_iter = ntiles > 0 ? (ztiles + ntiles - 1) / ntiles : 0;
        auto zero_span = [&](long u0, long len) {               // virtual vectors [u0, u0 + len), one side of the slice
            const long v0 = u0 < own_v0 ? u0 : u0 + nvec;
            bulk_store(gz + v0 * 16, zeros, (uint32_t)(len * 16));
        };
        auto zero_tiles = [&](long first, long count) {          // thread 0 only
            for (long z = first; z < first + count && z < ztiles; ++z) {
                const long u0 = z * TILE;
                const long len = zvec - u0 < TILE ? zvec - u0 : TILE;
                if (u0 < own_v0 && u0 + len > own_v0) {           // straddles the own slice
                    zero_span(u0, own_v0 - u0);
                    zero_span(own_v0, u0 + len - own_v0);
                } else {
                    zero_span(u0, len);
                }
            }
        };
        int it = 0;
        for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
            const int buf = it & 1;
            if (threadIdx.x == 0) tma_store_wait_read<1>();      // the copies that read stage[buf] two rounds ago
            __syncthreads();
            const long i = tile * TILE + threadIdx.x;
            if (!skip && i < nvec) {
                float g[8], p[8], mm[8], vv[8];
                unpack8<T>(*reinterpret_cast<const uint4*>(redsrc + i * 8), g);
                *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(a.p32 + i * 8);
                *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(a.p32 + i * 8 + 4);
                *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(a.m + i * 8);
                *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(a.m + i * 8 + 4);
                *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(a.v + i * 8);
                *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(a.v + i * 8 + 4);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    float gr = g[k] * inv;
                    if (!a.adamw) gr += a.wd * p[k];
                    mm[k] = a.beta1 * mm[k] + (1.f - a.beta1) * gr;
                    vv[k] = a.beta2 * vv[k] + (1.f - a.beta2) * gr * gr;
                    float upd = (mm[k] / a.bc1) / (sqrtf(vv[k] / a.bc2) + a.eps);
                    if (a.adamw) upd += a.wd * p[k];
                    p[k] -= a.lr * upd;
                }
                *reinterpret_cast<float4*>(a.p32 + i * 8) = *reinterpret_cast<float4*>(p);
                *reinterpret_cast<float4*>(a.p32 + i * 8 + 4) = *reinterpret_cast<float4*>(p + 4);
                *reinterpret_cast<float4*>(a.m + i * 8) = *reinterpret_cast<float4*>(mm);
                *reinterpret_cast<float4*>(a.m + i * 8 + 4) = *reinterpret_cast<float4*>(mm + 4);
                *reinterpret_cast<float4*>(a.v + i * 8) = *reinterpret_cast<float4*>(vv);
                *reinterpret_cast<float4*>(a.v + i * 8 + 4) = *reinterpret_cast<float4*>(vv + 4);
                stage[buf][threadIdx.x] = pack8<T>(p);
            }
            fence_proxy_async_smem();
            __syncthreads();
            if (threadIdx.x == 0) {
                if (!skip) {
                    const long v0 = tile * TILE;
                    const uint32_t bytes = (uint32_t)((nvec - v0 < TILE ? nvec - v0 : TILE) * 16);
                    const long off = (a.slice_off + v0 * 8) * (long)sizeof(T);
                    for (int j = 0; j < R; ++j) {
                        const int r = (me + j) % R;
                        bulk_store(reinterpret_cast<char*>(a.params.p[r]) + off, stage[buf], bytes);
                    }
                }
                if (a.zero_grads) {                               // the Adam loop is done with this tile of the slice
                    const long v0 = tile * TILE;
                    bulk_store(own + v0 * 8, zeros, (uint32_t)((nvec - v0 < TILE ? nvec - v0 : TILE) * 16));
                }
                zero_tiles((long)it * gridDim.x * z_per_iter + (long)blockIdx.x * z_per_iter, z_per_iter);
                tma_store_commit();
            }
        }
        // zero tiles not covered by the rounds above (CTAs with fewer rounds, skip with few tiles, ...)
        if (threadIdx.x == 0) {
            const long rounds = (ntiles + gridDim.x - 1) / gridDim.x;
            const long done = rounds * gridDim.x * z_per_iter;       // tiles [0, done) were assigned to some round
            // rounds a CTA did not run (its tile index fell past ntiles) are picked up here
            for (long r2 = it; r2 < rounds; ++r2)
                zero_tiles(r2 * gridDim.x * z_per_iter + (long)blockIdx.x * z_per_iter, z_per_iter);
            for (long z = done + blockIdx.x; z < ztiles; z += gridDim.x) zero_tiles(z, 1);
            tma_store_commit();
            tma_store_wait<0>();                                     // all my bulk writes have been performed
        }
        __threadfence_system();
        grid.sync();
        stamp(a, 4);
        // nobody leaves before every peer's stores into my arena are globally visible
        if (blockIdx.x == 0) xgpu_barrier(a.flags, R, me, 1, a.epoch);
        stamp(a, 5);
    }
}

static int step_grid(const void* kern) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 512, 0);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 2) per_sm = 2;
    return sms * per_sm;
}

}  // namespace dsb

using namespace dsb;

// `max_ctas` caps the grid (0 = fill the GPU): a reduce phase launched on a side stream while backward is still
// running should leave most SMs to the GEMMs - it is NVLink-bound and may sit in its entry barrier for a while.
extern "C" int dsb_fused_zero_step_grid(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs,
                                        const void* mc_grads, void* mc_params, int R, int me, int phases,
                                        int zero_grads, long slice_off, long bucket_off, long bucket_n, long n,
                                        int dtype, void* red, float* p32, float* m, float* v, float lr, float beta1,
                                        float beta2, float eps, float weight_decay, int adamw_mode,
                                        int bias_correction, int step, float loss_scale, float clip, float* sumsq2,
                                        float* inv_scale, int* skip, float* host_out2, void* scratch,
                                        unsigned int epoch, int max_ctas, void* stream) {
    if (R < 1 || R > MAX_RANKS || me < 0 || me >= R || (n % 8) || (slice_off % 8) || (bucket_off % 8) ||
        (bucket_n % 8) || !scratch || !(phases & 7))
        return DSB_ERR_ARG;
    if (dtype != DSB_DTYPE_BF16 && dtype != DSB_DTYPE_F16) return DSB_ERR_DTYPE;
    if ((phases & PH_EXCHANGE) && (!host_out2 || !inv_scale || !skip || !sumsq2)) return DSB_ERR_ARG;
    ZeroStepArgs a;
    for (int i = 0; i < MAX_RANKS; ++i) {
        a.grads.p[i] = grad_ptrs[i < R ? i : 0];
        a.params.p[i] = param_ptrs[i < R ? i : 0];
        a.flags.p[i] = reinterpret_cast<uint32_t*>(flag_ptrs[i < R ? i : 0]);
    }
    const bool mc = mc_grads != nullptr && mc_params != nullptr;
    a.mc_grads = reinterpret_cast<const char*>(mc_grads);
    a.mc_params = reinterpret_cast<char*>(mc_params);
    a.R = R; a.me = me; a.phases = phases; a.zero_grads = zero_grads;
    a.slice_off = slice_off; a.bucket_off = bucket_off; a.bucket_n = bucket_n; a.n = n;
    a.avg = 1.f / (float)R;
    a.red = red; a.p32 = p32; a.m = m; a.v = v;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay; a.adamw = adamw_mode;
    a.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
    a.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
    a.loss_scale = loss_scale; a.clip = clip;
    a.sumsq = sumsq2; a.inv_scale = inv_scale; a.skip = skip; a.host_out = host_out2;
    a.partials = reinterpret_cast<float*>(scratch);
    a.epoch = epoch;
    const void* kern;
    if (dtype == DSB_DTYPE_BF16)
        kern = mc ? (const void*)fused_zero_step_kernel<__nv_bfloat16, true>
                  : (const void*)fused_zero_step_kernel<__nv_bfloat16, false>;
    else
        kern = mc ? (const void*)fused_zero_step_kernel<__half, true> : (const void*)fused_zero_step_kernel<__half, false>;
    void* args[] = {&a};
    int grid = step_grid(kern);
    if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
    cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(512), args, 0,
                                                reinterpret_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_fused_zero_step(void* const* grad_ptrs, void* const* param_ptrs, void* const* flag_ptrs,
                                   const void* mc_grads, void* mc_params, int R, int me, int phases, int zero_grads,
                                   long slice_off, long bucket_off, long bucket_n, long n, int dtype, void* red,
                                   float* p32, float* m, float* v, float lr, float beta1, float beta2, float eps,
                                   float weight_decay, int adamw_mode, int bias_correction, int step, float loss_scale,
                                   float clip, float* sumsq2, float* inv_scale, int* skip, float* host_out2,
                                   void* scratch, unsigned int epoch, void* stream) {
    return dsb_fused_zero_step_grid(grad_ptrs, param_ptrs, flag_ptrs, mc_grads, mc_params, R, me, phases, zero_grads,
                                    slice_off, bucket_off, bucket_n, n, dtype, red, p32, m, v, lr, beta1, beta2, eps,
                                    weight_decay, adamw_mode, bias_correction, step, loss_scale, clip, sumsq2, inv_scale,
                                    skip, host_out2, scratch, epoch, 0, stream);
}
