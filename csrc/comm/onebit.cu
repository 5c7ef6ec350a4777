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

// phase timestamps (ns) for benches: CTA 0 writes %globaltimer at the phase boundaries behind the partial sums
DSB_DEVICE void ob_stamp(float* partials, int slot) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        reinterpret_cast<unsigned long long*>(partials + 1024)[slot] = t;
    }
}

// Optional fusion with 1-bit Adam's compression stage (reference runtime/fp16/onebit/adam.py:200-260): with `g` the
// momentum is formed on the way in (m <- beta1 * m + (1 - beta1) * g), with `p` the parameter update runs on the way
// out (p <- p - lr * (m / (sqrt(v) + eps) + wd * p)) - the averaged momentum is decompressed straight into the update
// instead of being written, re-read and combined by separate elementwise passes.
struct ObAdam {
    const float* g;      // local gradient (nullptr: m already holds the momentum to exchange)
    float* p;            // fp32 parameters to update (nullptr: exchange only)
    const float* v;      // frozen second moment
    float beta1, lr, eps, wd;
};

// All element loops run on float4 (16-byte accesses, two independent iterations in flight per thread in the pure
// streaming phases): the first version used 4-byte accesses and reached ~1/6 of HBM bandwidth
// (profiles/onebit_bench_w*.json history). `m` holds n_real <= n elements (n = padded size of the error buffers); the
// compensated values live in the error buffers between the phases (x in e_w, y in e_s), so m is read once and
// written once.
__global__ void __launch_bounds__(512) onebit_allreduce_kernel(float* __restrict__ m, float* __restrict__ e_w,
                                                               float* __restrict__ e_s, long n, long n_real, int W,
                                                               int me, ObPtrs P, ObAdam A,
                                                               float* __restrict__ partials, uint32_t epoch) {
    cg::grid_group grid = cg::this_grid();
    __shared__ float sh[33];
    __shared__ float ssc[OB_MAX_RANKS];
    const long c = n / W;
    const long cwords = c >> 5;
    const bool vec4 = (cwords & 3) == 0;      // four consecutive sign words never straddle two servers' chunks
    const long n4 = n >> 2, c4 = c >> 2, r4 = n_real >> 2;
    const long gtid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long gstride = (long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* ew4 = reinterpret_cast<float4*>(e_w);
    float4* es4 = reinterpret_cast<float4*>(e_s);
    const float4* g4 = reinterpret_cast<const float4*>(A.g);

    ob_stamp(partials, 0);
    // ---- phase A: x = momentum + e_w (kept in e_w), partial sum of squares
    float acc = 0.f;
    for (long i = gtid; i < n4; i += gstride) {
        float4 x = ew4[i];
        if (i < r4) {
            float4 a = m4[i];
            if (g4 != nullptr) {
                const float4 g = g4[i];
                const float b1 = A.beta1, b2 = 1.f - A.beta1;
                a.x = b1 * a.x + b2 * g.x; a.y = b1 * a.y + b2 * g.y; a.z = b1 * a.z + b2 * g.z; a.w = b1 * a.w + b2 * g.w;
            }
            x.x += a.x; x.y += a.y; x.z += a.z; x.w += a.w;
        }
        ew4[i] = x;
        acc += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
    grid.sync();
    const float s_w = sqrtf(fold_partials(partials, gridDim.x, sh) / (float)n);
    grid.sync();   // partials are reused below
    ob_stamp(partials, 1);

    // ---- phase B: worker error, sign packing, scatter chunk r to rank r (whole warps stay together: n % 32 == 0).
    // Two vectors per thread and trip, both loads issued before anything is stored (the loop is latency-bound otherwise).
    auto quantise_w = [&](long i, bool valid, const float4& x) {
        uint32_t nib = 0;
        if (valid) {
            nib = (x.x >= 0.f ? 1u : 0u) | (x.y >= 0.f ? 2u : 0u) | (x.z >= 0.f ? 4u : 0u) | (x.w >= 0.f ? 8u : 0u);
            float4 e;
            e.x = x.x - sgn_sel(nib, 0, s_w); e.y = x.y - sgn_sel(nib, 1, s_w);
            e.z = x.z - sgn_sel(nib, 2, s_w); e.w = x.w - sgn_sel(nib, 3, s_w);
            ew4[i] = e;
        }
        const uint32_t word = pack_nibbles(nib, lane);
        // (per-word stores here: gathering the four words into one 16-byte store measured slower in this phase)
        if (valid && (lane & 7) == 0) {
            const long w = i >> 3;
            const int dst = (int)(w / cwords);
            P.recv_signs[dst][(long)me * cwords + (w - (long)dst * cwords)] = word;
        }
    };
    for (long base = gtid - lane; base < n4; base += 2 * gstride) {
        const long i0 = base + lane, i1 = i0 + gstride;
        const bool v0 = i0 < n4, v1 = i1 < n4;
        const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 x0 = v0 ? ew4[i0] : zero4;
        const float4 x1 = v1 ? ew4[i1] : zero4;
        quantise_w(i0, v0, x0);
        quantise_w(i1, v1, x1);
    }
    if (blockIdx.x == 0 && threadIdx.x < W) P.recv_scales[threadIdx.x][me] = s_w;
    __threadfence_system();
    grid.sync();
    ob_stamp(partials, 2);
    if (blockIdx.x == 0) ob_barrier(P, W, me, 0, epoch);
    grid.sync();
    ob_stamp(partials, 3);

    // ---- phase C: server average of my chunk + server error feedback (y kept in e_s)
    const uint32_t* rs = P.recv_signs[me];
    if (threadIdx.x < W) ssc[threadIdx.x] = __ldcg(P.recv_scales[me] + threadIdx.x);
    __syncthreads();
    const float inv_w = 1.f / (float)W;
    acc = 0.f;
    auto average = [&](long j, const float4& e) {
        float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
        const int sh4 = (int)(j & 7) * 4;
        for (int w = 0; w < W; ++w) {
            const uint32_t bits = __ldcg(rs + (long)w * cwords + (j >> 3)) >> sh4;
            const float s = ssc[w];
            y.x += sgn_sel(bits, 0, s); y.y += sgn_sel(bits, 1, s);
            y.z += sgn_sel(bits, 2, s); y.w += sgn_sel(bits, 3, s);
        }
        y.x = y.x * inv_w + e.x; y.y = y.y * inv_w + e.y; y.z = y.z * inv_w + e.z; y.w = y.w * inv_w + e.w;
        es4[j] = y;
        acc += y.x * y.x + y.y * y.y + y.z * y.z + y.w * y.w;
    };
    {
        long j = gtid;
        for (; j + gstride < c4; j += 2 * gstride) {
            const float4 e0 = es4[j], e1 = es4[j + gstride];
            average(j, e0);
            average(j + gstride, e1);
        }
        if (j < c4) average(j, es4[j]);
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
    grid.sync();
    const float s_s = sqrtf(fold_partials(partials, gridDim.x, sh) / (float)c);

    // ---- phase D: server error, pack, broadcast my chunk's signs + scale to every rank (two vectors per trip)
    auto quantise_s = [&](long j, bool valid, const float4& y, long jw, bool valid_warp) {
        uint32_t nib = 0;
        if (valid) {
            nib = (y.x >= 0.f ? 1u : 0u) | (y.y >= 0.f ? 2u : 0u) | (y.z >= 0.f ? 4u : 0u) | (y.w >= 0.f ? 8u : 0u);
            float4 e;
            e.x = y.x - sgn_sel(nib, 0, s_s); e.y = y.y - sgn_sel(nib, 1, s_s);
            e.z = y.z - sgn_sel(nib, 2, s_s); e.w = y.w - sgn_sel(nib, 3, s_s);
            es4[j] = e;
        }
        const uint32_t word = pack_nibbles(nib, lane);
        if (vec4) {
            uint4 wv;
            wv.x = __shfl_sync(0xffffffffu, word, 0);
            wv.y = __shfl_sync(0xffffffffu, word, 8);
            wv.z = __shfl_sync(0xffffffffu, word, 16);
            wv.w = __shfl_sync(0xffffffffu, word, 24);
            if (lane < W && valid_warp) {      // one lane per destination rank
                const long w = (long)me * cwords + (jw >> 3);
                *reinterpret_cast<uint4*>(P.gath_signs[(me + lane) % W] + w) = wv;
            }
        } else if (valid && (lane & 7) == 0) {
            const long w = (long)me * cwords + (j >> 3);
            for (int r = 0; r < W; ++r) P.gath_signs[(me + r) % W][w] = word;
        }
    };
    for (long base = gtid - lane; base < c4; base += 2 * gstride) {
        const long j0 = base + lane, j1 = j0 + gstride;
        const bool v0 = j0 < c4, v1 = j1 < c4;
        const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 y0 = v0 ? es4[j0] : zero4;
        const float4 y1 = v1 ? es4[j1] : zero4;
        quantise_s(j0, v0, y0, base, base < c4);
        quantise_s(j1, v1, y1, base + gstride, base + gstride < c4);
    }
    if (blockIdx.x == 0 && threadIdx.x < W) P.gath_scales[threadIdx.x][me] = s_s;
    __threadfence_system();
    grid.sync();
    ob_stamp(partials, 4);
    if (blockIdx.x == 0) ob_barrier(P, W, me, 1, epoch);
    grid.sync();
    ob_stamp(partials, 5);

    // ---- phase E: decompress the averaged momentum into m (and, fused, apply the parameter update)
    const uint32_t* gs = P.gath_signs[me];
    if (threadIdx.x < W) ssc[threadIdx.x] = __ldcg(P.gath_scales[me] + threadIdx.x);
    __syncthreads();
    float4* p4 = reinterpret_cast<float4*>(A.p);
    const float4* v4 = reinterpret_cast<const float4*>(A.v);
    // chunk by chunk (one scale per chunk: no 64-bit division per vector), two vectors in flight per thread
    auto emit = [&](long i, float s0) {
        const uint32_t b0 = __ldcg(gs + (i >> 3)) >> ((int)(i & 7) * 4);
        const float4 mm = make_float4(sgn_sel(b0, 0, s0), sgn_sel(b0, 1, s0), sgn_sel(b0, 2, s0), sgn_sel(b0, 3, s0));
        m4[i] = mm;
        if (p4 != nullptr) {
            const float4 v = v4[i];
            float4 p = p4[i];
            p.x -= A.lr * (mm.x / (sqrtf(v.x) + A.eps) + A.wd * p.x);
            p.y -= A.lr * (mm.y / (sqrtf(v.y) + A.eps) + A.wd * p.y);
            p.z -= A.lr * (mm.z / (sqrtf(v.z) + A.eps) + A.wd * p.z);
            p.w -= A.lr * (mm.w / (sqrtf(v.w) + A.eps) + A.wd * p.w);
            p4[i] = p;
        }
    };
    for (int k = 0; k < W; ++k) {
        const float s0 = ssc[k];
        const long base = (long)k * c4;
        const long stop = base + c4 < r4 ? base + c4 : r4;      // m may be shorter than the padded error buffers
        long i = base + gtid;
        for (; i + gstride < stop; i += 2 * gstride) {
            emit(i, s0);
            emit(i + gstride, s0);
        }
        if (i < stop) emit(i, s0);
    }
    ob_stamp(partials, 6);
}

}  // namespace dsb

using namespace dsb;

// Symmetric-heap layout per rank for one tensor size (bytes from `base`): see dsb_onebit_heap_bytes.
extern "C" long dsb_onebit_heap_bytes(long n, int W) {
    const long words = n / 32;
    return 2 * (words * 4 + 256) + 2 * 256;
}

static int onebit_launch(float* m, float* e_w, float* e_s, long n, long n_real, int W, int me,
                         void* const* heap_bases, long region_off, void* const* flag_bases, void* scratch,
                         unsigned int epoch, const ObAdam& adam, void* stream) {
    if (W < 1 || W > OB_MAX_RANKS || n <= 0 || (n % (32L * W)) || n_real <= 0 || n_real > n || (n_real % 4) || !scratch)
        return DSB_ERR_ARG;
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
    ObAdam A = adam;
    void* args[] = {&m, &e_w, &e_s, &n, &n_real, &W, &me, &P, &A, &partials, &epoch};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)onebit_allreduce_kernel, dim3(sms * per_sm), dim3(512),
                                                args, 0, reinterpret_cast<cudaStream_t>(stream));
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_onebit_allreduce(float* m, float* e_w, float* e_s, long n, int W, int me, void* const* heap_bases,
                                    long region_off, void* const* flag_bases, void* scratch, unsigned int epoch,
                                    void* stream) {
    ObAdam none = {nullptr, nullptr, nullptr, 0.f, 0.f, 0.f, 0.f};
    return onebit_launch(m, e_w, e_s, n, n, W, me, heap_bases, region_off, flag_bases, scratch, epoch, none, stream);
}

// 1-bit Adam compression-stage step in one kernel: momentum from the local gradient, compressed exchange, parameter
// update. m / g / p / v hold n_real elements (multiple of 4), the error buffers the padded n.
extern "C" int dsb_onebit_adam_step(float* m, const float* g, float* p, const float* v, float* e_w, float* e_s, long n,
                                    long n_real, int W, int me, void* const* heap_bases, long region_off,
                                    void* const* flag_bases, void* scratch, unsigned int epoch, float beta1, float lr,
                                    float eps, float weight_decay, void* stream) {
    if (!g || !p || !v) return DSB_ERR_ARG;
    ObAdam a = {g, p, v, beta1, lr, eps, weight_decay};
    return onebit_launch(m, e_w, e_s, n, n_real, W, me, heap_bases, region_off, flag_bases, scratch, epoch, a, stream);
}
