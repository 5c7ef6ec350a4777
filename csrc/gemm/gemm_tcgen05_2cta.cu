// CTA-pair (cta_group::2) tcgen05 GEMM: two SMs of a TPC cooperate on one 256 x 256 output tile.
//
// Why: with one CTA per tile the tensor core must read A (128 x 16) and B (256 x 16) from shared memory for every
// 128-cycle UMMA while TMA refills the same shared memory, which exceeds the 128 B/clk shared-memory port (ncu:
// sm__pipe_tensor_cycles_active 74 %, profiles/gemm_v1_ncu.txt). In pair mode each CTA stages only HALF of B (128 rows)
// plus its own 128 rows of A, the pair MMA (M = 256, N = 256) reads both halves from the two SMs, so per-SM shared
// memory traffic drops by a third and the L2 -> SM operand traffic by a third as well.
//
//   cluster (2,1,1); CTA rank r owns accumulator rows [m0 + 128 r, m0 + 128 r + 128) x 256 columns in its own TMEM
//   warp 0 lane 0 (both CTAs): TMA producer, loads credited to CTA 0's full barrier (cta_group::2 TMA)
//   warp 1 lane 0 (CTA 0 only): issues tcgen05.mma.cta_group::2, commits (multicast) to both CTAs' empty barriers
//   warp 2 (both): pair TMEM allocation; warps 4..7 (both): epilogue of the CTA's own 128 rows
#include "../common/ptx_2sm.cuh"
#include "gemm_epilogue.cuh"
#include "../common/host_ctx.h"
#include <type_traits>

namespace dsb {
namespace g2 {

constexpr int BM = 128;          // rows per CTA (pair tile = 256)
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int EPI_COLS = 64;
constexpr int NUM_THREADS = 256;
constexpr int A_BYTES = BM * BK * 2;
constexpr int EPI_BYTES = 4 * 2 * 32 * 128;

// BN = pair-tile columns (256, or 128 when 256-wide tiles would leave the last wave mostly empty); each CTA stages
// BN / 2 rows of B.
template <int BN_>
struct Cfg {
    static constexpr int BN = BN_;
    static constexpr int BNH = BN / 2;
    static constexpr int B_BYTES = BNH * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;       // 32 KiB (BN 256) / 24 KiB (BN 128) per CTA
    static constexpr int STAGES = BN == 256 ? 6 : 8;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 256 + 1024;
    static constexpr int TMEM_COLS = 2 * BN;
};

// Work units. The first `full_tiles` units are whole output tiles. When the tile count leaves the last wave of CTA
// pairs mostly idle (256 tiles on 74 pairs = 3.46 waves), the remaining `tail_tiles` tiles are each cut into `split`
// K-slices of `kpp` k-blocks: slices 0..split-2 ("contributors") park their fp32 accumulators in a workspace and raise
// a flag, the last slice ("owner") adds them in its epilogue and writes the tile. Units are numbered
// [full tiles][all slices 0][all slices 1]...[owners] and every pair walks its units in increasing order, so each
// contributor precedes every owner on every pair and never waits itself: no deadlock with all pairs resident.
struct Params {
    EpiParams e;
    int K, batch;
    int a_mn, b_mn;
    int full_tiles, tail_tiles, split, kpp;
    float* ws;      // [tail_tiles][split - 1][2 CTAs][4 warps] regions of BN x 32 floats
    int* flags;     // [tail_tiles][2][4], zero between launches
};

enum { ROLE_FULL = 0, ROLE_CONTRIB = 1, ROLE_OWNER = 2 };

DSB_DEVICE void red_release_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
DSB_DEVICE int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <typename T, int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmD, const Params p) {
    using C = Cfg<BN>;
    constexpr int BNH = C::BNH, B_BYTES = C::B_BYTES, STAGE_BYTES = C::STAGE_BYTES, STAGES = C::STAGES,
                  TMEM_COLS = C::TMEM_COLS;
    (void)B_BYTES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_epi = smem + STAGES * STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_epi + EPI_BYTES);
    uint64_t* full_bar = bars;                 // [STAGES]  used in CTA 0
    uint64_t* empty_bar = bars + STAGES;       // [STAGES]  per CTA (multicast commit)
    uint64_t* tfull_bar = bars + 2 * STAGES;   // [2]       per CTA (multicast commit)
    uint64_t* tempty_bar = tfull_bar + 2;      // [2]       used in CTA 0, 8 arrivals (4 epilogue warps x 2 CTAs)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform role index
    const int lane = threadIdx.x & 31;
    const uint32_t cta = cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1;
    const int num_clusters = gridDim.x >> 1;

    const int M = p.e.M, N = p.e.N;
    const int m_blocks = (M + 2 * BM - 1) / (2 * BM);
    const int n_blocks = (N + BN - 1) / BN;
    const int tiles_per_batch = m_blocks * n_blocks;
    const int num_tiles = tiles_per_batch * p.batch;
    const int k_blocks = (p.K + BK - 1) / BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmD);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], 8);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_2sm<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();      // both CTAs' barriers are initialised before any cross-CTA traffic
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    auto decode = [&](int tile, int& b, int& mb, int& nb) {
        b = tile / tiles_per_batch;
        int t = tile - b * tiles_per_batch;
        const int GROUP = 4;
        int per_group = GROUP * n_blocks;
        int g = t / per_group;
        int first_m = g * GROUP;
        int gsz = min(m_blocks - first_m, GROUP);
        int r = t - g * per_group;
        mb = first_m + (r % gsz);
        nb = r / gsz;
    };

    const int num_units = p.full_tiles + p.tail_tiles * p.split;
    (void)num_tiles;
    auto unit_of = [&](int u, int& tile, int& kb0, int& kb1, int& role, int& tt, int& piece) {
        if (u < p.full_tiles) {
            tile = u; kb0 = 0; kb1 = k_blocks; role = ROLE_FULL; tt = 0; piece = 0;
        } else {
            const int v = u - p.full_tiles;
            piece = v / p.tail_tiles;
            tt = v - piece * p.tail_tiles;
            tile = p.full_tiles + tt;
            kb0 = piece * p.kpp;
            kb1 = min(kb0 + p.kpp, k_blocks);
            role = p.split == 1 ? ROLE_FULL : (piece == p.split - 1 ? ROLE_OWNER : ROLE_CONTRIB);
        }
    };

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (both CTAs)
        // elect.sync (not lane == 0): ptxas then knows exactly one thread is active and emits the uniform-datapath
        // UTMALDG / UTCHMMA directly instead of a per-active-lane serialisation loop around each of them.
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int u = cluster_id; u < num_units; u += num_clusters) {
                int tile, kb0, kb1, role, tt, piece, b, mb, nb;
                unit_of(u, tile, kb0, kb1, role, tt, piece);
                decode(tile, b, mb, nb);
                const int m0 = mb * 2 * BM + (int)cta * BM;      // this CTA's A rows
                const int n0 = nb * BN + (int)cta * BNH;         // this CTA's half of B
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    if (cta == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
                    if (!p.a_mn) {
                        tma_load_3d_2sm(sa, &tmA, &full_bar[stage], kb * BK, m0, b);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)
                            tma_load_3d_2sm(sa + j * (BK * 128), &tmA, &full_bar[stage], m0 + j * 64, kb * BK, b);
                    }
                    if (!p.b_mn) {
                        tma_load_3d_2sm(sb, &tmB, &full_bar[stage], kb * BK, n0, b);
                    } else {
#pragma unroll
                        for (int j = 0; j < BNH / 64; ++j)
                            tma_load_3d_2sm(sb + j * (BK * 128), &tmB, &full_bar[stage], n0 + j * 64, kb * BK, b);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ pair MMA issuer (CTA 0, one thread)
        if (cta == 0 && elect_one()) {
            const uint32_t idesc = make_idesc_f16(2 * BM, BN, std::is_same<T, __nv_bfloat16>::value, p.a_mn, p.b_mn);
            // smem matrix descriptor = {hi: SBO 1024 B | version | SWIZZLE_128B, lo: address >> 4 | LBO >> 4 << 16};
            // stage and k-step advance only the address field, so the loop carries two 32-bit adds per MMA.
            constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t a_lo0 = ((smem_u32(smem) & 0x3FFFF) >> 4) | ((p.a_mn ? (BK * 128u) >> 4 : 1u) << 16);
            const uint32_t b_lo0 = (((smem_u32(smem) + A_BYTES) & 0x3FFFF) >> 4) | ((p.b_mn ? (BK * 128u) >> 4 : 1u) << 16);
            const uint32_t a_kstep = p.a_mn ? (2048u >> 4) : (32u >> 4);
            const uint32_t b_kstep = p.b_mn ? (2048u >> 4) : (32u >> 4);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int u = cluster_id; u < num_units; u += num_clusters, ++it) {
                int tile, kb0, kb1, role, tt, piece;
                unit_of(u, tile, kb0, kb1, role, tt, piece);
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(&tempty_bar[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_lo = a_lo0 + stage * (STAGE_BYTES >> 4);
                    const uint32_t b_lo = b_lo0 + stage * (STAGE_BYTES >> 4);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)
                        umma_f16_2sm_lohi(d_tmem, a_lo + k * a_kstep, b_lo + k * b_kstep, DESC_HI, idesc,
                                          ((kb - kb0) | k) != 0);
                    umma_commit_2sm(&empty_bar[stage], 0x3);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit_2sm(&tfull_bar[as], 0x3);
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue (own 128 rows, 256 columns)
        const int q = warp & 3;
        uint8_t* my_epi = smem_epi + q * (2 * 32 * 128);
        int it = 0;
        int buf = 0;
        for (int u = cluster_id; u < num_units; u += num_clusters, ++it) {
            int tile, kb0, kb1, role, tt, piece, b, mb, nb;
            unit_of(u, tile, kb0, kb1, role, tt, piece);
            decode(tile, b, mb, nb);
            const int m0 = mb * 2 * BM + (int)cta * BM, n0 = nb * BN;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            mbar_wait(&tfull_bar[as], aphase);
            tc_fence_after();
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < M;
            // split-K bookkeeping: one flag and (split - 1) workspace regions per (tail tile, CTA, epilogue warp)
            int* flag = p.flags + ((tt * 2 + (int)cta) * 4 + q);
            constexpr int REGION_F4 = BN * 8;   // BN x 32 floats as float4
            float4* ws4 = reinterpret_cast<float4*>(p.ws) +
                          (size_t)(((tt * (p.split - 1)) * 2 + (int)cta) * 4 + q) * REGION_F4 + lane;
            const size_t piece_stride = (size_t)2 * 4 * REGION_F4;
            if (role == ROLE_OWNER) {
                if (lane == 0) {
                    while (ld_acquire_gpu(flag) < p.split - 1) {}
                }
                __syncwarp();
            }
#pragma unroll 1
            for (int c = 0; c < BN / EPI_COLS; ++c) {
                uint32_t r0[32], r1[32];
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + c * EPI_COLS;
                tmem_ld_32x32b_x32(taddr, r0);
                tmem_ld_32x32b_x32(taddr + 32, r1);
                tmem_ld_wait();
                if (c == BN / EPI_COLS - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(&tempty_bar[as], 0);   // hand the stage back to CTA 0's MMA warp
                }
                const int col0 = n0 + c * EPI_COLS;
                if (col0 >= N || m0 + q * 32 >= M) continue;
                if (role == ROLE_CONTRIB) {
                    // park the raw fp32 accumulators: [chunk][16 float4][lane] so a warp writes 512 contiguous bytes
                    float4* dst = ws4 + piece * piece_stride + (size_t)c * 16 * 32;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        dst[j * 32] = make_float4(__uint_as_float(r0[4 * j]), __uint_as_float(r0[4 * j + 1]),
                                                  __uint_as_float(r0[4 * j + 2]), __uint_as_float(r0[4 * j + 3]));
                        dst[(8 + j) * 32] = make_float4(__uint_as_float(r1[4 * j]), __uint_as_float(r1[4 * j + 1]),
                                                        __uint_as_float(r1[4 * j + 2]), __uint_as_float(r1[4 * j + 3]));
                    }
                    continue;
                }
                float v[64];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    v[i] = __uint_as_float(r0[i]);
                    v[32 + i] = __uint_as_float(r1[i]);
                }
                if (role == ROLE_OWNER) {
                    for (int pc = 0; pc < p.split - 1; ++pc) {
                        const float4* src = ws4 + pc * piece_stride + (size_t)c * 16 * 32;
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float4 t = __ldcg(src + j * 32);
                            v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
                        }
                    }
                }
#pragma unroll
                for (int i = 0; i < 64; ++i) v[i] *= p.e.alpha;
                epilogue_apply<T>(v, p.e, b, row, row_ok, col0);
                uint8_t* sbuf = my_epi + buf * (32 * 128);
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
                epilogue_stage<T>(v, sbuf, lane);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_store_3d(&tmD, sbuf, col0, m0 + q * 32, b);
                    tma_store_commit();
                }
                buf ^= 1;
            }
            if (role == ROLE_CONTRIB) {
                __threadfence();
                __syncwarp();
                if (lane == 0) red_release_add(flag, 1);
            } else if (role == ROLE_OWNER) {
                __syncwarp();
                if (lane == 0) *flag = 0;   // ready for the next launch (stream order separates the launches)
            }
        }
        if (lane == 0) tma_store_wait<0>();
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();      // the peer may still be reading this CTA's shared memory / signalling its barriers
    tc_fence_after();
    if (warp == 2) tmem_dealloc_2sm<TMEM_COLS>(tmem_base);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}
static int make_map(CUtensorMap* m, int is_bf16, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1,
                    uint64_t s2, uint32_t b0, uint32_t b1) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return -1;
    cuuint64_t dims[3] = {d0, d1, d2};
    cuuint64_t strides[2] = {s1 * 2, s2 * 2};
    cuuint32_t box[3] = {b0, b1, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3,
                     const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -(int)r - 1000;
}

template <typename T, int BN>
static int launch(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& d, const Params& p, int sms,
                  cudaStream_t stream) {
    static bool attr_set = false;
    constexpr int SMEM_BYTES = Cfg<BN>::SMEM_BYTES;
    auto kern = gemm_pair_kernel<T, BN>;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        if (e != cudaSuccess) return -(int)e;
        attr_set = true;
    }
    const long units = (long)p.full_tiles + (long)p.tail_tiles * p.split;
    long clusters = sms / 2;
    if (units < clusters) clusters = units;
    kern<<<(unsigned)(2 * clusters), NUM_THREADS, SMEM_BYTES, stream>>>(a, b, d, p);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace g2
}  // namespace dsb

// ---- split-K workspace: one per stream that launches a split GEMM (launches on one stream are ordered, so a region
// is never written by two launches at once), grown on demand.
namespace {
struct SplitWs {
    cudaStream_t stream;
    float* ws;
    size_t ws_bytes;
    int* flags;
    int flag_count;
};
constexpr int MAX_WS = 8;
SplitWs g_ws[MAX_WS];
int g_ws_used = 0;
int g_split_mode = -1;   // -1 = choose per shape, 1 = never split, 2..4 = force (test hook)

SplitWs* get_ws(cudaStream_t stream, size_t bytes, int flag_count) {
    SplitWs* w = nullptr;
    for (int i = 0; i < g_ws_used; ++i)
        if (g_ws[i].stream == stream) w = &g_ws[i];
    if (!w) {
        if (g_ws_used == MAX_WS) return nullptr;
        w = &g_ws[g_ws_used++];
        *w = SplitWs{stream, nullptr, 0, nullptr, 0};
    }
    if (w->ws_bytes < bytes || w->flag_count < flag_count) {
        // growing means cudaMalloc / cudaFree / a stream sync: none of that is legal while the stream is being
        // captured into a CUDA graph -> the caller falls back to whole tiles for this launch
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) {
            cudaGetLastError();
            return nullptr;
        }
    }
    if (w->ws_bytes < bytes) {
        if (w->ws) { cudaStreamSynchronize(stream); cudaFree(w->ws); }
        if (cudaMalloc(&w->ws, bytes) != cudaSuccess) { w->ws = nullptr; w->ws_bytes = 0; return nullptr; }
        w->ws_bytes = bytes;
    }
    if (w->flag_count < flag_count) {
        if (w->flags) { cudaStreamSynchronize(stream); cudaFree(w->flags); }
        if (cudaMalloc(&w->flags, flag_count * sizeof(int)) != cudaSuccess) { w->flags = nullptr; w->flag_count = 0; return nullptr; }
        cudaMemsetAsync(w->flags, 0, flag_count * sizeof(int), stream);
        w->flag_count = flag_count;
    }
    return w;
}
}  // namespace

// Test / tuning hook: -1 = choose the K split of the tail wave per shape, 1 = never split, 2..4 = force that split.
extern "C" int dsb_gemm_pair_force_split(int split) {
    if (split != -1 && (split < 1 || split > 4)) return DSB_ERR_ARG;
    g_split_mode = split;
    return 0;
}

static int g_force_bn = 0;
// Test / tuning hook: 0 = choose per shape, 128 or 256 = force that pair-tile width.
extern "C" int dsb_gemm_pair_force_bn(int bn) {
    if (bn != 0 && bn != 128 && bn != 256) return DSB_ERR_ARG;
    g_force_bn = bn;
    return 0;
}

// Same contract as dsb_gemm (include/dsb200.h); 256 x 256 (or 256 x 128) pair tiles. Best for M >= 256, N >= 128.
extern "C" int dsb_gemm_pair_ex(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda,
                                long ldb, long ldd, long strideA, long strideB, long strideD, int a_mn_major,
                                int b_mn_major, int dtype, int epilogue, float alpha, const void* bias,
                                const void* aux, long ldaux, long stride_aux, void* aux_out, long ldaux_out,
                                long stride_aux_out, int split_mode, void* stream_);

extern "C" int dsb_gemm_pair(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb,
                             long ldd, long strideA, long strideB, long strideD, int a_mn_major, int b_mn_major,
                             int dtype, int epilogue, float alpha, const void* bias, const void* aux, long ldaux,
                             long stride_aux, void* aux_out, long ldaux_out, long stride_aux_out, void* stream_) {
    return dsb_gemm_pair_ex(A, B, D, M, N, K, batch, lda, ldb, ldd, strideA, strideB, strideD, a_mn_major, b_mn_major,
                            dtype, epilogue, alpha, bias, aux, ldaux, stride_aux, aux_out, ldaux_out, stride_aux_out,
                            g_split_mode, stream_);
}

// split_mode: -1 = choose the tail-wave K split per shape, 1 = whole tiles only, 2..4 = that split (if K allows).
extern "C" int dsb_gemm_pair_ex(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda,
                                long ldb, long ldd, long strideA, long strideB, long strideD, int a_mn_major,
                                int b_mn_major, int dtype, int epilogue, float alpha, const void* bias,
                                const void* aux, long ldaux, long stride_aux, void* aux_out, long ldaux_out,
                                long stride_aux_out, int split_mode, void* stream_) {
    using namespace dsb;
    using namespace dsb::g2;
    ensure_context_on_this_thread();
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (M <= 0 || N <= 0 || K <= 0 || batch <= 0) return 0;
    if ((lda % 8) || (ldb % 8) || (ldd % 8) || (N % 8)) return DSB_ERR_ALIGN;
    if ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(D)) & 15)
        return DSB_ERR_ALIGN;
    if (batch > 1 && ((strideA % 8) || (strideB % 8) || (strideD % 8))) return DSB_ERR_ALIGN;
    if ((epilogue == DSB_EPI_BIAS_RESIDUAL || epilogue == DSB_EPI_DGELU) && (!aux || (ldaux % 8))) return DSB_ERR_ARG;
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int bf = dtype == DSB_DTYPE_BF16;
    if (batch == 1) {
        strideA = (long)(a_mn_major ? K : M) * lda;
        strideB = (long)(b_mn_major ? K : N) * ldb;
        strideD = (long)M * ldd;
    }
    // 256-wide pair tiles; 128-wide only for narrow outputs. (Measured: picking 128 to even out the last wave of the
    // K = 8192, N = 2048 shapes loses more to the extra operand traffic per FLOP than it wins: 1290 vs 1475 TFLOP/s.)
    const int BN = g_force_bn ? g_force_bn : (N <= 128 ? 128 : 256);
    const int BNH = BN / 2;
    CUtensorMap ta, tb, td;
    int rc;
    if (!a_mn_major) rc = make_map(&ta, bf, A, K, M, batch, lda, strideA, BK, BM);
    else rc = make_map(&ta, bf, A, M, K, batch, lda, strideA, 64, BK);
    if (rc) return rc;
    if (!b_mn_major) rc = make_map(&tb, bf, B, K, N, batch, ldb, strideB, BK, BNH);
    else rc = make_map(&tb, bf, B, N, K, batch, ldb, strideB, 64, BK);
    if (rc) return rc;
    rc = make_map(&td, bf, D, N, M, batch, ldd, strideD, EPI_COLS, 32);
    if (rc) return rc;
    Params p;
    p.e.M = M; p.e.N = N; p.e.epilogue = epilogue; p.e.alpha = alpha;
    p.e.bias = bias; p.e.aux = aux; p.e.ldaux = ldaux; p.e.stride_aux = stride_aux;
    p.e.aux_out = aux_out; p.e.ldaux_out = ldaux_out; p.e.stride_aux_out = stride_aux_out;
    p.K = K; p.batch = batch; p.a_mn = a_mn_major; p.b_mn = b_mn_major;
    // tail-wave K split (see Params): pick the split that minimises the tail's duration in units of one whole tile
    const int tiles = ((M + 2 * BM - 1) / (2 * BM)) * ((N + BN - 1) / BN) * batch;
    const int k_blocks = (K + BK - 1) / BK;
    const int pairs = sms / 2;
    const int tail = tiles >= pairs ? tiles % pairs : tiles;
    int split = 1;
    if (tail > 0 && BN == 256) {
        double best = 1.0;
        for (int sp = 2; sp <= 4; ++sp) {
            const int kpp = (k_blocks + sp - 1) / sp;
            if (kpp < 8 || (sp - 1) * kpp >= k_blocks) continue;
            const double cost = (double)((tail * sp + pairs - 1) / pairs) / sp + 0.03 * (sp - 1);
            if (cost < best - 1e-9) { best = cost; split = sp; }
        }
        if (split_mode >= 1) {
            split = split_mode;
            const int kpp = (k_blocks + split - 1) / split;
            if ((split - 1) * kpp >= k_blocks) split = 1;
        }
    }
    p.split = split;
    p.kpp = (k_blocks + split - 1) / split;
    p.tail_tiles = split > 1 ? tail : 0;
    p.full_tiles = tiles - p.tail_tiles;
    p.ws = nullptr;
    p.flags = nullptr;
    if (split > 1) {
        SplitWs* w = get_ws(stream, (size_t)tail * (split - 1) * 8 * BN * 32 * sizeof(float), tail * 8);
        if (!w) {   // no workspace: fall back to whole tiles
            p.split = 1; p.kpp = k_blocks; p.tail_tiles = 0; p.full_tiles = tiles;
        } else {
            p.ws = w->ws;
            p.flags = w->flags;
        }
    }
    if (BN == 256) {
        if (bf) return launch<__nv_bfloat16, 256>(ta, tb, td, p, sms, stream);
        return launch<__half, 256>(ta, tb, td, p, sms, stream);
    }
    if (bf) return launch<__nv_bfloat16, 128>(ta, tb, td, p, sms, stream);
    return launch<__half, 128>(ta, tb, td, p, sms, stream);
}
