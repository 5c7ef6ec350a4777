// Fused ZeRO-1/2 data-parallel step over NVLink / NVSwitch peer memory (north-star item 1).
//
// The gradient and parameter arenas of every rank of a data-parallel group live in a symmetric heap (cudaMalloc +
// CUDA IPC, see deeperspeed_b200/parallel/symm.py), so each GPU can address every peer's arena directly. Two kernels
// replace {NCCL reduce-scatter, norm pass, unscale pass, Adam, fp32->fp16 copy, NCCL all-gather}:
//
//   fused_rs_norm   : cross-GPU flag barrier -> each rank PULLS its slice of the gradient bucket from all peers with
//                     16-byte peer loads (ld.relaxed.sys over NVLink), sums in fp32, averages, writes the fp32
//                     reduced gradient and emits sum-of-squares + non-finite flag (deterministic last-block fold).
//   fused_adam_ag   : Adam/AdamW on fp32 master/m/v with the device-side inverse scale + skip flag, casts to the
//                     low-precision dtype and PUSHES the updated slice into every peer's parameter arena with 16-byte
//                     peer stores (this is the all-gather), then a closing cross-GPU barrier.
//
// Between the two, the host issues one 2-float NCCL all-reduce (norm + overflow across dp and pipeline stages).
// Cross-GPU synchronisation: monotonically increasing epoch flags in each rank's heap, st.release.sys /
// ld.acquire.sys; grids are launched cooperatively (all CTAs resident) because CTAs wait on CTA 0.
// Replaces reference engine.allreduce_bucket + zero/stage1.reduce_scatter_gradients + stage1.step all_gather
// (deepspeed/runtime/engine.py:1323-1395, zero/stage1.py:629-783).
#include "xgpu.cuh"
#include "../../include/dsb200.h"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace dsb {

struct AdamH {
    float lr, beta1, beta2, eps, wd, bc1, bc2;
    int adamw;
};

template <typename T>
__global__ void __launch_bounds__(512) fused_rs_norm_kernel(PeerPtrs grads, FlagPtrs flags, int R, int me,
                                                            long slice_off, long n, float scale,
                                                            float* __restrict__ red32, float* __restrict__ sumsq_out,
                                                            int* __restrict__ flag_out, float* __restrict__ partials,
                                                            unsigned int* __restrict__ counter, uint32_t epoch) {
    cg::grid_group grid = cg::this_grid();
    __shared__ float sh[33];
    // every peer has finished writing its gradient arena (its kernel is stream-ordered after its backward)
    if (blockIdx.x == 0) xgpu_barrier(flags, R, me, 0, epoch);
    grid.sync();

    const long nvec = n >> 3;
    const long stride = (long)gridDim.x * blockDim.x;
    float acc = 0.f;
    int bad = 0;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // peers in batches of 4: issue the 16-byte peer loads first (independent requests in flight), then reduce
        for (int j0 = 0; j0 < R; j0 += 4) {
            uint4 raw[4];
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                if (j0 + jj < R) {
                    const int r = (me + j0 + jj) % R;
                    raw[jj] = ld_na_v4(reinterpret_cast<const T*>(grads.p[r]) + slice_off + i * 8);
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
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            s[k] *= scale;
            acc += s[k] * s[k];
            bad |= !isfinite(s[k]);
        }
        *reinterpret_cast<float4*>(red32 + i * 8) = make_float4(s[0], s[1], s[2], s[3]);
        *reinterpret_cast<float4*>(red32 + i * 8 + 4) = make_float4(s[4], s[5], s[6], s[7]);
    }
    acc = block_sum(acc, sh);
    bad = __syncthreads_or(bad);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = acc;
        partials[gridDim.x + blockIdx.x] = bad ? 1.f : 0.f;
    }
    grid.sync();
    if (blockIdx.x == 0) {
        float t = 0.f, b = 0.f;
        for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
            t += partials[i];
            b += partials[gridDim.x + i];
        }
        t = block_sum(t, sh);
        b = block_sum(b, sh);
        if (threadIdx.x == 0) {
            *sumsq_out += t;
            if (b > 0.f || !isfinite(t)) *flag_out = 1;
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(512) fused_adam_ag_kernel(const float* __restrict__ red32, float* __restrict__ p32,
                                                            float* __restrict__ m, float* __restrict__ v,
                                                            PeerPtrs params, FlagPtrs flags, int R, int me,
                                                            long slice_off, long n, AdamH h,
                                                            const float* __restrict__ inv_scale_dev,
                                                            const int* __restrict__ skip_dev, uint32_t epoch) {
    cg::grid_group grid = cg::this_grid();
    const bool skip = skip_dev != nullptr && *skip_dev != 0;
    const float inv = inv_scale_dev != nullptr ? *inv_scale_dev : 1.f;
    if (!skip) {
        const long nvec = n >> 3;
        const long stride = (long)gridDim.x * blockDim.x;
        for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
            float g[8], p[8], mm[8], vv[8];
            *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(red32 + i * 8);
            *reinterpret_cast<float4*>(g + 4) = *reinterpret_cast<const float4*>(red32 + i * 8 + 4);
            *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(p32 + i * 8);
            *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(p32 + i * 8 + 4);
            *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(m + i * 8);
            *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(m + i * 8 + 4);
            *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(v + i * 8);
            *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(v + i * 8 + 4);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float gr = g[k] * inv;
                if (!h.adamw) gr += h.wd * p[k];
                mm[k] = h.beta1 * mm[k] + (1.f - h.beta1) * gr;
                vv[k] = h.beta2 * vv[k] + (1.f - h.beta2) * gr * gr;
                float upd = (mm[k] / h.bc1) / (sqrtf(vv[k] / h.bc2) + h.eps);
                if (h.adamw) upd += h.wd * p[k];
                p[k] -= h.lr * upd;
            }
            *reinterpret_cast<float4*>(p32 + i * 8) = *reinterpret_cast<float4*>(p);
            *reinterpret_cast<float4*>(p32 + i * 8 + 4) = *reinterpret_cast<float4*>(p + 4);
            *reinterpret_cast<float4*>(m + i * 8) = *reinterpret_cast<float4*>(mm);
            *reinterpret_cast<float4*>(m + i * 8 + 4) = *reinterpret_cast<float4*>(mm + 4);
            *reinterpret_cast<float4*>(v + i * 8) = *reinterpret_cast<float4*>(vv);
            *reinterpret_cast<float4*>(v + i * 8 + 4) = *reinterpret_cast<float4*>(vv + 4);
            const uint4 lp = pack8<T>(p);
            // the all-gather: push the updated slice into every rank's parameter arena (own copy first)
#pragma unroll 4
            for (int j = 0; j < R; ++j) {
                const int r = (me + j) % R;
                st_na_v4(reinterpret_cast<T*>(params.p[r]) + slice_off + i * 8, lp);
            }
        }
    }
    __threadfence_system();
    grid.sync();
    // nobody leaves before every peer's stores into my arena are globally visible
    if (blockIdx.x == 0) xgpu_barrier(flags, R, me, 1, epoch);
}

// ---- LAMB variant ------------------------------------------------------------------------------------------------
// One launch per parameter group covers every bucket slice this rank owns (local index l -> bucket l / sub, arena
// offset (bucket * R + me) * sub + l % sub). LAMB's trust ratio is per TENSOR while a tensor may be spread over
// several ranks' slices, so the kernel has three phases separated by cross-GPU barriers:
//   1. moments + update u (kept in place of the reduced gradient) + per-tensor partial sums of w^2 and u^2, folded
//      per CTA in shared-memory bins and added to this rank's norm table in the symmetric heap;
//   2. after a cross-GPU barrier every CTA reads the R norm tables over NVLink for the tensors it touches and
//      forms coeff = clamp(|w| / |u|, min, max);
//   3. p -= step_size * coeff * u on the fp32 master, cast, and push into every peer's parameter arena.
// Replaces reference csrc/lamb/fused_lamb_cuda_kernel.cu:186-310 + the stage-1 all_gather (zero/stage1.py:763-773);
// the reference itself applies LAMB per flat sub-partition under ZeRO, i.e. with a per-partition ratio.
struct LambH {
    float beta1, beta2, eps, wd, max_coeff, min_coeff, step_size;
    int eps_inside;
};

DSB_DEVICE int find_seg(const long* __restrict__ seg_start, int nseg, long e) {
    int lo = 0, hi = nseg - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (seg_start[mid] <= e) lo = mid; else hi = mid - 1;
    }
    return lo;
}

constexpr int LAMB_BINS = 256;

template <typename T>
__global__ void __launch_bounds__(512) fused_lamb_ag_kernel(float* __restrict__ u32, float* __restrict__ p32,
                                                            float* __restrict__ m, float* __restrict__ v,
                                                            PeerPtrs params, PeerPtrs norms, FlagPtrs flags, int R,
                                                            int me, long sub, long n,
                                                            const long* __restrict__ seg_start,
                                                            const int* __restrict__ seg_tid, int nseg, int ntensors,
                                                            LambH h, const float* __restrict__ inv_scale_dev,
                                                            const int* __restrict__ skip_dev,
                                                            float* __restrict__ coeff_out, uint32_t epoch) {
    cg::grid_group grid = cg::this_grid();
    __shared__ float bin_w[LAMB_BINS], bin_u[LAMB_BINS], bin_c[LAMB_BINS];
    const bool skip = skip_dev != nullptr && *skip_dev != 0;
    const float inv = inv_scale_dev != nullptr ? *inv_scale_dev : 1.f;
    float* my_norms = reinterpret_cast<float*>(norms.p[me]);
    // the previous launch's closing barrier guarantees that no peer still reads this table
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * ntensors; i += gridDim.x * blockDim.x) my_norms[i] = 0.f;
    for (int i = threadIdx.x; i < LAMB_BINS; i += blockDim.x) {
        bin_w[i] = 0.f;
        bin_u[i] = 0.f;
        bin_c[i] = 1.f;
    }
    grid.sync();

    // contiguous range of 8-element vectors per CTA, so a CTA crosses few tensor boundaries
    const long nvec = n >> 3;
    const long per_block = (nvec + gridDim.x - 1) / gridDim.x;
    const long v0 = (long)blockIdx.x * per_block;
    const long v1 = v0 + per_block < nvec ? v0 + per_block : nvec;
    const bool have = v0 < v1;
    const int seg0 = have ? find_seg(seg_start, nseg, v0 * 8) : 0;
    const int seg1 = have ? find_seg(seg_start, nseg, (v1 - 1) * 8) : -1;

    if (!skip && have) {
        int cur = -1;
        long next_start = 0;
        float w2 = 0.f, u2 = 0.f;
        for (long i = v0 + threadIdx.x; i < v1; i += blockDim.x) {
            const long e = i * 8;
            if (cur < 0 || e >= next_start) {
                if (cur >= 0) {
                    const int rel = cur - seg0;
                    if (rel < LAMB_BINS) {
                        atomicAdd(&bin_w[rel], w2);
                        atomicAdd(&bin_u[rel], u2);
                    } else {
                        atomicAdd(&my_norms[2 * seg_tid[cur]], w2);
                        atomicAdd(&my_norms[2 * seg_tid[cur] + 1], u2);
                    }
                    w2 = u2 = 0.f;
                }
                cur = find_seg(seg_start, nseg, e);
                next_start = seg_start[cur + 1];
            }
            float g[8], p[8], mm[8], vv[8];
            *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(u32 + e);
            *reinterpret_cast<float4*>(g + 4) = *reinterpret_cast<const float4*>(u32 + e + 4);
            *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(p32 + e);
            *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(p32 + e + 4);
            *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(m + e);
            *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(m + e + 4);
            *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(v + e);
            *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(v + e + 4);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float gr = g[k] * inv;
                mm[k] = h.beta1 * mm[k] + (1.f - h.beta1) * gr;
                vv[k] = h.beta2 * vv[k] + (1.f - h.beta2) * gr * gr;
                const float denom = h.eps_inside ? sqrtf(vv[k] + h.eps) : sqrtf(vv[k]) + h.eps;
                g[k] = mm[k] / denom + h.wd * p[k];
                w2 += p[k] * p[k];
                u2 += g[k] * g[k];
            }
            *reinterpret_cast<float4*>(u32 + e) = *reinterpret_cast<float4*>(g);
            *reinterpret_cast<float4*>(u32 + e + 4) = *reinterpret_cast<float4*>(g + 4);
            *reinterpret_cast<float4*>(m + e) = *reinterpret_cast<float4*>(mm);
            *reinterpret_cast<float4*>(m + e + 4) = *reinterpret_cast<float4*>(mm + 4);
            *reinterpret_cast<float4*>(v + e) = *reinterpret_cast<float4*>(vv);
            *reinterpret_cast<float4*>(v + e + 4) = *reinterpret_cast<float4*>(vv + 4);
        }
        if (cur >= 0) {
            const int rel = cur - seg0;
            if (rel < LAMB_BINS) {
                atomicAdd(&bin_w[rel], w2);
                atomicAdd(&bin_u[rel], u2);
            } else {
                atomicAdd(&my_norms[2 * seg_tid[cur]], w2);
                atomicAdd(&my_norms[2 * seg_tid[cur] + 1], u2);
            }
        }
    }
    __syncthreads();
    for (int b = threadIdx.x; b < LAMB_BINS && seg0 + b <= seg1; b += blockDim.x) {
        if (bin_w[b] != 0.f || bin_u[b] != 0.f) {
            atomicAdd(&my_norms[2 * seg_tid[seg0 + b]], bin_w[b]);
            atomicAdd(&my_norms[2 * seg_tid[seg0 + b] + 1], bin_u[b]);
        }
    }
    __threadfence_system();
    grid.sync();
    // every rank's norm table is complete and visible before anybody reads it
    if (blockIdx.x == 0) xgpu_barrier(flags, R, me, 2, epoch);
    grid.sync();

    auto coeff_for = [&](int tid) -> float {
        float w = 0.f, u = 0.f;
        for (int r = 0; r < R; ++r) {
            const uint32_t* t = reinterpret_cast<const uint32_t*>(norms.p[r]) + 2 * tid;
            w += __uint_as_float(ld_relaxed_sys(t));
            u += __uint_as_float(ld_relaxed_sys(t + 1));
        }
        float c = 1.f;
        if (w > 0.f && u > 0.f) c = fminf(fmaxf(sqrtf(w) / sqrtf(u), h.min_coeff), h.max_coeff);
        return c;
    };
    for (int b = threadIdx.x; b < LAMB_BINS && seg0 + b <= seg1; b += blockDim.x) {
        const int tid = seg_tid[seg0 + b];
        const float c = coeff_for(tid);
        bin_c[b] = c;
        if (coeff_out != nullptr && !skip) coeff_out[tid] = c;
    }
    __syncthreads();

    if (!skip && have) {
        int cur = -1;
        long next_start = 0;
        float c = 1.f;
        for (long i = v0 + threadIdx.x; i < v1; i += blockDim.x) {
            const long e = i * 8;
            if (cur < 0 || e >= next_start) {
                cur = find_seg(seg_start, nseg, e);
                next_start = seg_start[cur + 1];
                c = (cur - seg0 < LAMB_BINS) ? bin_c[cur - seg0] : coeff_for(seg_tid[cur]);
            }
            float u[8], p[8];
            *reinterpret_cast<float4*>(u) = *reinterpret_cast<const float4*>(u32 + e);
            *reinterpret_cast<float4*>(u + 4) = *reinterpret_cast<const float4*>(u32 + e + 4);
            *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(p32 + e);
            *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(p32 + e + 4);
            const float s = h.step_size * c;
#pragma unroll
            for (int k = 0; k < 8; ++k) p[k] -= s * u[k];
            *reinterpret_cast<float4*>(p32 + e) = *reinterpret_cast<float4*>(p);
            *reinterpret_cast<float4*>(p32 + e + 4) = *reinterpret_cast<float4*>(p + 4);
            const uint4 lp = pack8<T>(p);
            const long arena_off = ((e / sub) * R + me) * sub + e % sub;
#pragma unroll 4
            for (int j = 0; j < R; ++j) {
                const int r = (me + j) % R;
                st_na_v4(reinterpret_cast<T*>(params.p[r]) + arena_off, lp);
            }
        }
    }
    __threadfence_system();
    grid.sync();
    // nobody leaves before every peer's stores into my arena are visible and every peer is done with my norm table
    if (blockIdx.x == 0) xgpu_barrier(flags, R, me, 1, epoch);
}

static int coop_grid(const void* kern, int threads) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, 0);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 2) per_sm = 2;
    return sms * per_sm;
}

}  // namespace dsb

using namespace dsb;

extern "C" long dsb_fused_zero_scratch_bytes(void) { return (2L * 148 * 4 + 16) * sizeof(float); }

extern "C" int dsb_fused_rs_norm(void* const* grad_ptrs, void* const* flag_ptrs, int R, int me, long slice_off, long n,
                                 int dtype, float scale, float* red32, float* sumsq_out, int* flag_out, void* scratch,
                                 unsigned int epoch, void* stream) {
    if (R < 1 || R > MAX_RANKS || (n % 8) || (slice_off % 8) || !scratch) return DSB_ERR_ARG;
    PeerPtrs g;
    FlagPtrs f;
    for (int i = 0; i < R; ++i) {
        g.p[i] = grad_ptrs[i];
        f.p[i] = reinterpret_cast<uint32_t*>(flag_ptrs[i]);
    }
    float* partials = reinterpret_cast<float*>(scratch);
    unsigned int* counter = reinterpret_cast<unsigned int*>(partials + 2 * 148 * 4);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    void* args[] = {&g, &f, &R, &me, &slice_off, &n, &scale, &red32, &sumsq_out, &flag_out, &partials, &counter, &epoch};
    const void* kern = dtype == DSB_DTYPE_BF16 ? (const void*)fused_rs_norm_kernel<__nv_bfloat16>
                                               : (const void*)fused_rs_norm_kernel<__half>;
    if (dtype != DSB_DTYPE_BF16 && dtype != DSB_DTYPE_F16) return DSB_ERR_DTYPE;
    cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(coop_grid(kern, 512)), dim3(512), args, 0, s);
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_fused_adam_ag(const float* red32, float* p32, float* m, float* v, void* const* param_ptrs,
                                 void* const* flag_ptrs, int R, int me, long slice_off, long n, int dtype, float lr,
                                 float beta1, float beta2, float eps, float weight_decay, int adamw_mode,
                                 int bias_correction, int step, const float* inv_scale_dev, const int* skip_dev,
                                 unsigned int epoch, void* stream) {
    if (R < 1 || R > MAX_RANKS || (n % 8) || (slice_off % 8)) return DSB_ERR_ARG;
    PeerPtrs pp;
    FlagPtrs f;
    for (int i = 0; i < R; ++i) {
        pp.p[i] = param_ptrs[i];
        f.p[i] = reinterpret_cast<uint32_t*>(flag_ptrs[i]);
    }
    AdamH h;
    h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.wd = weight_decay; h.adamw = adamw_mode;
    h.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
    h.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    void* args[] = {&red32, &p32, &m, &v, &pp, &f, &R, &me, &slice_off, &n, &h, &inv_scale_dev, &skip_dev, &epoch};
    if (dtype != DSB_DTYPE_BF16 && dtype != DSB_DTYPE_F16) return DSB_ERR_DTYPE;
    const void* kern = dtype == DSB_DTYPE_BF16 ? (const void*)fused_adam_ag_kernel<__nv_bfloat16>
                                               : (const void*)fused_adam_ag_kernel<__half>;
    cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(coop_grid(kern, 512)), dim3(512), args, 0, s);
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_fused_lamb_ag(float* upd32, float* p32, float* m, float* v, void* const* param_ptrs,
                                 void* const* norm_ptrs, void* const* flag_ptrs, int R, int me, long sub, long n,
                                 const long* seg_start, const int* seg_tid, int nseg, int ntensors, int dtype, float lr,
                                 float beta1, float beta2, float eps, float weight_decay, float max_coeff,
                                 float min_coeff, int eps_inside_sqrt, int bias_correction, int step,
                                 const float* inv_scale_dev, const int* skip_dev, float* coeff_out,
                                 unsigned int epoch, void* stream) {
    if (R < 1 || R > MAX_RANKS || sub <= 0 || (sub % 8) || (n % sub) || nseg < 1 || ntensors < 1 || !seg_start ||
        !seg_tid)
        return DSB_ERR_ARG;
    if (dtype != DSB_DTYPE_BF16 && dtype != DSB_DTYPE_F16) return DSB_ERR_DTYPE;
    PeerPtrs pp, nn;
    FlagPtrs f;
    for (int i = 0; i < R; ++i) {
        pp.p[i] = param_ptrs[i];
        nn.p[i] = norm_ptrs[i];
        f.p[i] = reinterpret_cast<uint32_t*>(flag_ptrs[i]);
    }
    LambH h;
    h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.wd = weight_decay;
    h.max_coeff = max_coeff; h.min_coeff = min_coeff; h.eps_inside = eps_inside_sqrt;
    h.step_size = bias_correction ? lr * sqrtf(1.f - powf(beta2, (float)step)) / (1.f - powf(beta1, (float)step)) : lr;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    void* args[] = {&upd32, &p32, &m, &v, &pp, &nn, &f, &R, &me, &sub, &n, &seg_start, &seg_tid, &nseg, &ntensors,
                    &h, &inv_scale_dev, &skip_dev, &coeff_out, &epoch};
    const void* kern = dtype == DSB_DTYPE_BF16 ? (const void*)fused_lamb_ag_kernel<__nv_bfloat16>
                                               : (const void*)fused_lamb_ag_kernel<__half>;
    cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(coop_grid(kern, 512)), dim3(512), args, 0, s);
    return e == cudaSuccess ? 0 : -(int)e;
}

// ---- symmetric heap plumbing (CUDA IPC)------------------------------------------------------------------------
extern "C" int dsb_symm_alloc(long bytes, void** out) {
    cudaError_t e = cudaMalloc(out, bytes);
    if (e != cudaSuccess) return -(int)e;
    e = cudaMemset(*out, 0, bytes);
    return e == cudaSuccess ? 0 : -(int)e;
}
extern "C" int dsb_symm_free(void* p) { return -(int)cudaFree(p); }
extern "C" int dsb_ipc_get_handle(void* devptr, void* handle64) {
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, devptr);
    if (e != cudaSuccess) return -(int)e;
    memcpy(handle64, &h, sizeof(h));
    return 0;
}
extern "C" int dsb_ipc_open(const void* handle64, void** out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    cudaError_t e = cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess);
    return e == cudaSuccess ? 0 : -(int)e;
}
extern "C" int dsb_ipc_close(void* p) { return -(int)cudaIpcCloseMemHandle(p); }
