// Persistent warp-specialised tcgen05 GEMM for sm_100a with fused epilogues.
//
//   D[b] = epilogue(alpha * op(A[b]) * op(B[b])^T)      A: M x K, B: N x K (each K-major or MN-major), D: M x N
//
// Design (B200-first, no library GEMM on this path):
//   * one CTA per SM, static round-robin tile schedule with grouped rasterisation (L2 reuse of B tiles)
//   * warp 0 lane 0 : TMA producer (cp.async.bulk.tensor, 128-byte swizzle) into a multi-stage smem ring
//   * warp 1 lane 0 : issues tcgen05.mma (UMMA 128 x BN x 16, fp32 accumulate in TMEM) and tcgen05.commit
//   * warp 2        : TMEM allocator (2 accumulator stages so epilogue(i) overlaps mainloop(i+1))
//   * warps 4..7    : epilogue; tcgen05.ld -> registers -> fused math -> swizzled smem -> TMA store
//   * operands may be K-major ([rows, K] row-major) or MN-major ([K, rows] row-major), which is what the
//     forward (NT), dgrad (NN) and wgrad (TN) products of a Linear layer need without any transposed copy.
//
// Replaces the reference's cublasGemmEx / cublasGemmStridedBatchedEx call sites and the SIMT epilogue kernels
// that followed them (reference: csrc/transformer/cublas_wrappers.cu:16-164, gelu_kernels.cu:61-128,
// dropout_kernels.cu:386-465, general_kernels.cu:89-182).
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"
#include "../common/host_ctx.h"
#include <stdio.h>
#include <type_traits>

namespace dsb {

constexpr int BM = 128;
constexpr int BK = 64;           // 64 x 16-bit = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int EPI_COLS = 64;     // epilogue chunk width (one 128-byte swizzled row of D)
constexpr int NUM_THREADS = 256;

struct GemmParams {
    int M, N, K, batch;
    int a_mn, b_mn;
    int epilogue;
    float alpha;
    const void* bias;
    const void* aux;
    long ldaux, stride_aux;
    void* aux_out;
    long ldaux_out, stride_aux_out;
    int rows_per_rank;   // gather variant: rows of B owned by each rank
    // overlapped all-gather variant (ag_overlap != 0): warps 2-3 of every CTA copy the remote shards panel by panel
    // (one panel = the BN rows of B one n-block needs) into `gathered`, the producer loads a remote panel from there
    // once panel_ready[nb] shows that every CTA has delivered its share; own panels come straight from the own shard.
    // ag_overlap == 2: B is MN-major ([K, N], the dgrad of a row-sharded weight) and the shards split K. A panel is
    // 256 rows of B = a K range; every tile needs all panels, so the k loop starts at the own shard's range and
    // walks the remote ranges in the order the copiers deliver them.
    int ag_overlap, me, nranks;
    const void* peer_shards[8];
    void* gathered;
    int* panel_ready;
    // GEMM -> reduce-scatter variant (rs_mode != 0, wgrad of a ZeRO-3 linear): rows of D are owned by ranks
    // (rows_per_rank each). Every output tile is TMA-stored into slot `me` of the OWNER's staging buffer
    // ([nranks, rows_per_rank, N], peer-mapped; tmBs.m[owner] for remote owners, tmD for the own one), remote owners
    // first so that the transfers run under the remaining tiles. After a grid + cross-GPU barrier (flag blocks at
    // peer_shards[r], slot 3) all CTAs sum the nranks slots of the local staging into rs_out, scaled by rs_scale.
    int rs_mode;
    unsigned int rs_epoch;
    float rs_scale;
    void* rs_out;       // [rows_per_rank, N]
    const void* rs_stage;   // local staging base
    int* rs_sync;       // [2] zeroed by the host: arrive counter, go flag
};

template <int BN> struct GemmCfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (BN == 256) ? 4 : 6;
    static constexpr int EPI_BYTES = 4 /*warps*/ * 2 /*bufs*/ * 32 * 128;
    static constexpr int BAR_BYTES = 256;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024 /*align slack*/;
    static constexpr int TMEM_COLS = 2 * BN;  // 512 or 256, power of two
};

// NMAPS > 1: the B operand (K-major weight [N, K]) is row-sharded over NMAPS ranks; each rank's shard has its own
// tensor map (peer-mapped memory), and the producer picks the map of the rank that owns the tile's rows. This is the
// fused all-gather + GEMM of ZeRO-3: weight tiles travel over NVLink straight into shared memory, no gathered copy.
template <int NMAPS> struct BMapSet {
    CUtensorMap m[NMAPS];
};

template <typename T, int BN, int NMAPS>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ BMapSet<NMAPS> tmBs,
                    const __grid_constant__ CUtensorMap tmD, const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_epi = smem + Cfg::STAGES * Cfg::STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_epi + Cfg::EPI_BYTES);
    uint64_t* full_bar = bars;                      // [STAGES]
    uint64_t* empty_bar = bars + Cfg::STAGES;       // [STAGES]
    uint64_t* tfull_bar = bars + 2 * Cfg::STAGES;   // [2]
    uint64_t* tempty_bar = tfull_bar + 2;           // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform role index
    const int lane = threadIdx.x & 31;

    const int m_blocks = (p.M + BM - 1) / BM;
    const int n_blocks = (p.N + BN - 1) / BN;
    const int tiles_per_batch = m_blocks * n_blocks;
    const int num_tiles = tiles_per_batch * p.batch;
    const int k_blocks = (p.K + BK - 1) / BK;
    // K-sharded gather (ag_overlap == 2): two passes over the tile list, each tile on the same CTA both times. Pass 0
    // multiplies the own shard's K range (nothing to wait for), pass 1 the remote ranges and adds the pass-0 result
    // in its epilogue, so the remote panels are first needed half-way through the kernel.
    // With a short own range (< 32 k-blocks) a pass's mainloop no longer covers the epilogue and the second trip
    // through D costs more than the first wave's wait for the first panels, so such shapes run in one pass.
    const bool ksplit = NMAPS > 1 && p.ag_overlap == 2;
    const int kb_own = ksplit ? p.rows_per_rank / BK : 0;
    const int n_pass = (ksplit && p.nranks > 1 && kb_own >= 32) ? 2 : 1;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmBs.m[0]);
        tma_prefetch_desc(&tmD);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < Cfg::STAGES; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], 4);
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // tile -> (batch, m_blk, n_blk) with groups of 8 M-blocks sweeping N (keeps a B panel hot in L2)
    auto decode = [&](int tile, int& b, int& mb, int& nb) {
        if constexpr (NMAPS > 1) {
            if (p.ag_overlap == 1) {
                // panel-major order starting at the own shard: remote panels are needed as late as possible
                b = 0;
                const int seq = tile / m_blocks;
                mb = tile - seq * m_blocks;
                nb = (seq + p.me * (p.rows_per_rank / BN)) % n_blocks;
                return;
            }
            if (p.rs_mode) {
                // owner-major order: (me + 1) % R first, own rows last
                b = 0;
                const int mbpr = p.rows_per_rank / BM;
                const int per_owner = mbpr * n_blocks;
                const int seq = tile / per_owner;
                const int r = tile - seq * per_owner;
                const int owner = (p.me + 1 + seq) % p.nranks;
                mb = owner * mbpr + r % mbpr;
                nb = r / mbpr;
                return;
            }
        }
        b = tile / tiles_per_batch;
        int t = tile - b * tiles_per_batch;
        const int GROUP = 8;
        int per_group = GROUP * n_blocks;
        int g = t / per_group;
        int first_m = g * GROUP;
        int gsz = min(m_blocks - first_m, GROUP);
        int r = t - g * per_group;
        mb = first_m + (r % gsz);
        nb = r / gsz;
    };

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        // elect.sync rather than lane == 0: ptxas then knows exactly one thread is active and emits UTMALDG / UTCHMMA
        // directly instead of a per-active-lane serialisation loop around every one of them (137 -> ~50 SASS
        // instructions per k-block in the MMA loop; the issuing thread was the bottleneck, profiles/gemm_ncu_v2.md).
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            int ready_seq = -1;   // K-sharded gather: last remote panel (copier order) this CTA has seen complete
            const int last_panel = ksplit ? ((p.me * kb_own + k_blocks - 1) % k_blocks) >> 2 : 0;
            for (int pass = 0; pass < n_pass; ++pass) {
            const int kbi0 = (n_pass == 2 && pass == 1) ? kb_own : 0;
            const int kbi1 = (n_pass == 2 && pass == 0) ? kb_own : k_blocks;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int b, mb, nb;
                decode(tile, b, mb, nb);
                const int m0 = mb * BM, n0 = nb * BN;
                const int kb_rot = ksplit ? p.me * kb_own : 0;
                for (int kbi = kbi0; kbi < kbi1; ++kbi) {
                    // K-sharded gather: own K range first, then the remote ranges in copier order
                    int kb = kbi + kb_rot;
                    if (kb >= k_blocks) kb -= k_blocks;
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * Cfg::STAGE_BYTES;
                    uint8_t* sb = sa + Cfg::A_BYTES;
                    mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
                    if (!p.a_mn) {
                        tma_load_3d(sa, &tmA, &full_bar[stage], kb * BK, m0, b);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)
                            tma_load_3d(sa + j * (BK * 128), &tmA, &full_bar[stage], m0 + j * 64, kb * BK, b);
                    }
                    const int bslot = (NMAPS > 1 && p.rs_mode) ? p.me : 0;   // rs mode: B is local, its map sits in slot me
                    if (!p.b_mn) {
                        if (NMAPS == 1) {
                            tma_load_3d(sb, &tmBs.m[0], &full_bar[stage], kb * BK, n0, b);
                        } else if (p.rs_mode) {
                            tma_load_3d(sb, &tmBs.m[bslot], &full_bar[stage], kb * BK, n0, b);
                        } else if (!p.ag_overlap) {
                            const int owner = n0 / p.rows_per_rank;
                            tma_load_3d(sb, &tmBs.m[owner], &full_bar[stage], kb * BK, n0 - owner * p.rows_per_rank, b);
                        } else {
                            const int owner = n0 / p.rows_per_rank;
                            if (owner == p.me) {
                                tma_load_3d(sb, &tmBs.m[1], &full_bar[stage], kb * BK, n0 - owner * p.rows_per_rank, b);
                            } else {
                                if (kb == 0) {   // the panel was written by generic-proxy stores of all CTAs
                                    while (ld_acquire_gpu_s32(p.panel_ready + nb) < (int)gridDim.x) {}
                                    fence_proxy_async_all();
                                }
                                tma_load_3d(sb, &tmBs.m[0], &full_bar[stage], kb * BK, n0, b);
                            }
                        }
                    } else if (ksplit) {
                        const int krow = kb * BK;
                        if (kbi < kb_own) {   // the rotation puts the own shard's range first
#pragma unroll
                            for (int j = 0; j < BN / 64; ++j)
                                tma_load_3d(sb + j * (BK * 128), &tmBs.m[1], &full_bar[stage], n0 + j * 64,
                                            krow - p.me * p.rows_per_rank, 0);
                        } else {
                            // Panels become ready in copier order and stay ready, so a panel is polled (an L2 round
                            // trip plus a proxy fence on the producer's critical path) only the first time this CTA
                            // gets that far; with a poll at every panel start of every tile the producer could not
                            // keep the ring full (dgrad ran at 0.31 ms instead of 0.20 ms).
                            const int seq = (kbi - kb_own) >> 2;   // remote panel index in copier order (4 k-blocks each)
                            if (seq > ready_seq) {
                                while (ld_acquire_gpu_s32(p.panel_ready + (krow >> 8)) < (int)gridDim.x) {}
                                ready_seq = seq;
                                if (ld_acquire_gpu_s32(p.panel_ready + last_panel) >= (int)gridDim.x) ready_seq = 1 << 30;
                                fence_proxy_async_all();
                            }
#pragma unroll
                            for (int j = 0; j < BN / 64; ++j)
                                tma_load_3d(sb + j * (BK * 128), &tmBs.m[0], &full_bar[stage], n0 + j * 64, krow, 0);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_3d(sb + j * (BK * 128), &tmBs.m[bslot], &full_bar[stage], n0 + j * 64, kb * BK, b);
                    }
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
            }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (single thread)
        if (elect_one()) {
            const uint32_t idesc = make_idesc_f16(BM, BN, sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value,
                                                  p.a_mn, p.b_mn);
            // smem matrix descriptor = {hi: SBO 1024 B | version | SWIZZLE_128B, lo: address >> 4 | LBO >> 4 << 16};
            // K-major: k-step = 16 elements (32 B) inside the swizzled 128-B row; MN-major: 16 k-rows (2048 B),
            // LBO = next 64-element MN chunk. Stage and k-step advance only the address field.
            constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t a_lo0 = ((smem_u32(smem) & 0x3FFFF) >> 4) | ((p.a_mn ? (BK * 128u) >> 4 : 1u) << 16);
            const uint32_t b_lo0 =
                (((smem_u32(smem) + Cfg::A_BYTES) & 0x3FFFF) >> 4) | ((p.b_mn ? (BK * 128u) >> 4 : 1u) << 16);
            const uint32_t a_kstep = p.a_mn ? (2048u >> 4) : (32u >> 4);
            const uint32_t b_kstep = p.b_mn ? (2048u >> 4) : (32u >> 4);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int pass = 0; pass < n_pass; ++pass) {
            const int kcount = n_pass == 2 ? (pass == 0 ? kb_own : k_blocks - kb_own) : k_blocks;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(&tempty_bar[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = 0; kb < kcount; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_lo = a_lo0 + stage * (Cfg::STAGE_BYTES >> 4);
                    const uint32_t b_lo = b_lo0 + stage * (Cfg::STAGE_BYTES >> 4);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)
                        umma_f16_lohi(d_tmem, a_lo + k * a_kstep, b_lo + k * b_kstep, DESC_HI, idesc, (kb | k) != 0);
                    umma_commit(&empty_bar[stage]);
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(&tfull_bar[as]);
            }
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue (4 warps, 32 rows each)
        const int q = warp & 3;  // TMEM lane quarter this warp may touch
        uint8_t* my_epi = smem_epi + q * (2 * 32 * 128);
        int it = 0;
        int buf = 0;
        for (int pass = 0; pass < n_pass; ++pass) {
        if (pass == 1) {   // pass 1 reads (generic proxy) what this warp's pass-0 TMA stores wrote
            if (lane == 0) {
                tma_store_wait<0>();
                fence_proxy_async_all();
            }
            __syncwarp();
        }
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            int b, mb, nb;
            decode(tile, b, mb, nb);
            const int m0 = mb * BM, n0 = nb * BN;
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            mbar_wait(&tfull_bar[as], aphase);
            tc_fence_after();
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < p.M;
#pragma unroll 1
            for (int c = 0; c < BN / EPI_COLS; ++c) {
                uint32_t r0[32], r1[32];
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + c * EPI_COLS;
                tmem_ld_32x32b_x32(taddr, r0);
                tmem_ld_32x32b_x32(taddr + 32, r1);
                tmem_ld_wait();
                if (c == BN / EPI_COLS - 1) {
                    // accumulator stage fully read -> hand it back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[as]);
                }
                const int col0 = n0 + c * EPI_COLS;
                if (col0 >= p.N || m0 + q * 32 >= p.M) continue;  // whole box out of range (warp-uniform)

                float v[64];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    v[i] = __uint_as_float(r0[i]) * p.alpha;
                    v[32 + i] = __uint_as_float(r1[i]) * p.alpha;
                }
                const int epi = pass == 1 ? DSB_EPI_BIAS_RESIDUAL : p.epilogue;   // pass 1: aux = D of pass 0
                if (p.bias != nullptr && epi != DSB_EPI_DGELU) {
                    const T* bp = reinterpret_cast<const T*>(p.bias) + col0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (col0 + j * 8 < p.N) {
                            uint4 u = __ldg(reinterpret_cast<const uint4*>(bp) + j);
                            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                float2 f = Cvt<T>::unpack2(w[t]);
                                v[j * 8 + t * 2] += f.x;
                                v[j * 8 + t * 2 + 1] += f.y;
                            }
                        }
                    }
                }
                if (epi == DSB_EPI_BIAS_GELU) {
                    if (p.aux_out != nullptr && row_ok) {
                        T* op = reinterpret_cast<T*>(p.aux_out) + b * p.stride_aux_out + (long)row * p.ldaux_out + col0;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            if (col0 + j * 8 < p.N) {
                                uint4 u;
                                u.x = Cvt<T>::pack2(v[j * 8 + 0], v[j * 8 + 1]);
                                u.y = Cvt<T>::pack2(v[j * 8 + 2], v[j * 8 + 3]);
                                u.z = Cvt<T>::pack2(v[j * 8 + 4], v[j * 8 + 5]);
                                u.w = Cvt<T>::pack2(v[j * 8 + 6], v[j * 8 + 7]);
                                *(reinterpret_cast<uint4*>(op) + j) = u;
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
                                uint4 u = *(reinterpret_cast<const uint4*>(ap) + j);
                                const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                                for (int t = 0; t < 4; ++t) {
                                    float2 f = Cvt<T>::unpack2(w[t]);
                                    if (epi == DSB_EPI_DGELU) {
                                        v[j * 8 + t * 2] *= gelu_tanh_grad(f.x);
                                        v[j * 8 + t * 2 + 1] *= gelu_tanh_grad(f.y);
                                    } else {
                                        v[j * 8 + t * 2] += f.x;
                                        v[j * 8 + t * 2 + 1] += f.y;
                                    }
                                }
                            }
                        }
                    }
                }
                // stage the 32 x 64 chunk in swizzled smem, then one TMA store per warp
                uint8_t* sbuf = my_epi + buf * (32 * 128);
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint4 u;
                    u.x = Cvt<T>::pack2(v[j * 8 + 0], v[j * 8 + 1]);
                    u.y = Cvt<T>::pack2(v[j * 8 + 2], v[j * 8 + 3]);
                    u.z = Cvt<T>::pack2(v[j * 8 + 4], v[j * 8 + 5]);
                    u.w = Cvt<T>::pack2(v[j * 8 + 6], v[j * 8 + 7]);
                    *reinterpret_cast<uint4*>(sbuf + lane * 128 + ((j ^ (lane & 7)) << 4)) = u;
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (NMAPS > 1) {
                        if (p.rs_mode) {
                            const int owner = m0 / p.rows_per_rank;
                            const int srow = p.me * p.rows_per_rank + (m0 - owner * p.rows_per_rank) + q * 32;
                            tma_store_3d(owner == p.me ? &tmD : &tmBs.m[owner], sbuf, col0, srow, 0);
                        } else {
                            tma_store_3d(&tmD, sbuf, col0, m0 + q * 32, b);
                        }
                    } else {
                        tma_store_3d(&tmD, sbuf, col0, m0 + q * 32, b);
                    }
                    tma_store_commit();
                }
                buf ^= 1;
            }
        }
        }
        if (lane == 0) {
            tma_store_wait<0>();
            if constexpr (NMAPS > 1) {
                if (p.rs_mode) {   // the tile stores (async proxy, possibly into a peer) precede the barrier below
                    fence_proxy_async_all();
                    __threadfence_system();
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ warps 2-3: all-gather copier (overlap mode)
        if constexpr (NMAPS > 1) {
            if (p.ag_overlap && p.nranks > 1) {
                const int t = threadIdx.x - 64;                       // 0..63
                // a panel = 256 rows of the sharded operand; its rows are K long (mode 1) or N long (mode 2)
                const int ppr = p.rows_per_rank / 256;                // panels per rank
                const long panel_vecs = 256L * (p.ag_overlap == 2 ? p.N : p.K) / 8;   // shards are dense
                const int n_panels = ppr * p.nranks;
                const long cta_stride = (long)gridDim.x * 64;
                uint4* gath = reinterpret_cast<uint4*>(p.gathered);
                // remote panels in the order the tiles will need them, two panels per round so that every thread keeps
                // up to 16 independent 16-byte peer loads in flight (NVLink round trips are ~4 us: with 4 in flight the
                // copy ran at ~100 GB/s and the W = 4 case stalled the tensor pipe)
                constexpr int PG = 2, UNR = 16;
                const int pv = (int)panel_vecs;
                const int cstride = (int)cta_stride;
                (void)gath;
                for (int seq0 = ppr; seq0 < n_panels; seq0 += PG) {
                    const int np = min(PG, n_panels - seq0);
                    const int total = np * pv;
                    for (int i0 = blockIdx.x * 64 + t; i0 < total; i0 += UNR * cstride) {
                        uint4 v[UNR];
#pragma unroll
                        for (int u = 0; u < UNR; ++u) {
                            const int i = i0 + u * cstride;
                            if (i < total) {
                                const int pi = i / pv, off = i - pi * pv;
                                const int nb = (seq0 + pi + p.me * ppr) % n_panels;
                                const int owner = nb / ppr;
                                v[u] = ld_na_v4(reinterpret_cast<const uint4*>(p.peer_shards[owner]) +
                                                (long)(nb - owner * ppr) * panel_vecs + off);
                            }
                        }
#pragma unroll
                        for (int u = 0; u < UNR; ++u) {
                            const int i = i0 + u * cstride;
                            if (i < total) {
                                const int pi = i / pv, off = i - pi * pv;
                                const int nb = (seq0 + pi + p.me * ppr) % n_panels;
                                reinterpret_cast<uint4*>(p.gathered)[(long)nb * panel_vecs + off] = v[u];
                            }
                        }
                    }
                    __threadfence();
                    named_bar_sync(1, 64);
                    if (t < np) red_release_gpu_add_s32(p.panel_ready + (seq0 + t + p.me * ppr) % n_panels, 1);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 2) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);

    if constexpr (NMAPS > 1) {
        if (p.rs_mode) {
            // every CTA of every rank has delivered its tiles -> reduce the slots that landed here
            if (threadIdx.x == 0) {
                __threadfence_system();
                red_release_gpu_add_s32(p.rs_sync, 1);
            }
            if (blockIdx.x == 0) {
                if (threadIdx.x == 0)
                    while (ld_acquire_gpu_s32(p.rs_sync) < (int)gridDim.x) {}
                __syncthreads();
                if ((int)threadIdx.x < p.nranks) {
                    __threadfence_system();
                    uint32_t* theirs = reinterpret_cast<uint32_t*>(const_cast<void*>(p.peer_shards[threadIdx.x]));
                    st_release_sys(theirs + 3 * 16 + p.me, p.rs_epoch);
                    const uint32_t* mine = reinterpret_cast<const uint32_t*>(p.peer_shards[p.me]) + 3 * 16 + threadIdx.x;
                    while ((int32_t)(ld_acquire_sys(mine) - p.rs_epoch) < 0) {}
                }
                __syncthreads();
                if (threadIdx.x == 0) red_release_gpu_add_s32(p.rs_sync + 1, 1);
            }
            if (threadIdx.x == 0)
                while (ld_acquire_gpu_s32(p.rs_sync + 1) < 1) {}
            __syncthreads();
            const long nvec = (long)p.rows_per_rank * p.N / 8;
            const uint4* st = reinterpret_cast<const uint4*>(p.rs_stage);
            uint4* out = reinterpret_cast<uint4*>(p.rs_out);
            // one CTA per SM and 256 threads: keep UN x nranks 16-byte loads in flight per thread, otherwise the slot
            // sum runs at a fraction of the HBM bandwidth (it was ~40 us of the 75 us this kernel spent beyond its GEMM)
            constexpr int UN = 4;
            const long gstride = (long)gridDim.x * NUM_THREADS;
            for (long i0 = (long)blockIdx.x * NUM_THREADS + threadIdx.x; i0 < nvec; i0 += UN * gstride) {
                float acc[UN][8];
#pragma unroll
                for (int u = 0; u < UN; ++u)
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[u][k] = 0.f;
                for (int r0 = 0; r0 < p.nranks; r0 += 2) {
                    uint4 raw[UN][2];
#pragma unroll
                    for (int u = 0; u < UN; ++u)
#pragma unroll
                        for (int rr = 0; rr < 2; ++rr) {
                            const long i = i0 + u * gstride;
                            if (i < nvec && r0 + rr < p.nranks) raw[u][rr] = ld_na_v4(st + (long)(r0 + rr) * nvec + i);
                        }
#pragma unroll
                    for (int u = 0; u < UN; ++u)
#pragma unroll
                        for (int rr = 0; rr < 2; ++rr) {
                            if (i0 + u * gstride < nvec && r0 + rr < p.nranks) {
                                float f[8];
                                unpack8<T>(raw[u][rr], f);
#pragma unroll
                                for (int k = 0; k < 8; ++k) acc[u][k] += f[k];
                            }
                        }
                }
#pragma unroll
                for (int u = 0; u < UN; ++u) {
                    const long i = i0 + u * gstride;
                    if (i < nvec) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) acc[u][k] *= p.rs_scale;
                        out[i] = pack8<T>(acc[u]);
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------- host side
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

template <typename T, int BN, int NMAPS = 1>
static int launch(const CUtensorMap& a, const BMapSet<NMAPS>& b, const CUtensorMap& d, const GemmParams& p, int sms,
                  cudaStream_t stream) {
    using Cfg = GemmCfg<BN>;
    static bool attr_set = false;
    auto kern = gemm_tcgen05_kernel<T, BN, NMAPS>;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return -(int)e;
        attr_set = true;
    }
    const int m_blocks = (p.M + BM - 1) / BM, n_blocks = (p.N + BN - 1) / BN;
    const long tiles = (long)m_blocks * n_blocks * p.batch;
    const int grid = (int)(tiles < sms ? tiles : sms);
    if (p.ag_overlap) {
        // the TMA producers wait for panels copied by OTHER CTAs of this grid: every CTA must be resident. A
        // cooperative launch guarantees that or fails (cudaErrorCooperativeLaunchTooLarge) instead of hanging when
        // SMs are held by another kernel's context limit.
        void* args[] = {const_cast<CUtensorMap*>(&a), const_cast<BMapSet<NMAPS>*>(&b), const_cast<CUtensorMap*>(&d),
                        const_cast<GemmParams*>(&p)};
        cudaError_t e = cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(NUM_THREADS), args,
                                                    Cfg::SMEM_BYTES, stream);
        return e == cudaSuccess ? 0 : -(int)e;
    }
    kern<<<grid, NUM_THREADS, Cfg::SMEM_BYTES, stream>>>(a, b, d, p);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace dsb

extern "C" int dsb_gemm(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb,
                        long ldd, long strideA, long strideB, long strideD, int a_mn_major, int b_mn_major, int dtype,
                        int epilogue, float alpha, const void* bias, const void* aux, long ldaux, long stride_aux,
                        void* aux_out, long ldaux_out, long stride_aux_out, void* stream_) {
    using namespace dsb;
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
    const int bn = (N > 128) ? 256 : 128;
    const int bf = dtype == DSB_DTYPE_BF16;
    if (batch == 1) {
        strideA = (long)(a_mn_major ? K : M) * lda;
        strideB = (long)(b_mn_major ? K : N) * ldb;
        strideD = (long)M * ldd;
    }
    CUtensorMap ta, td;
    BMapSet<1> tbs;
    CUtensorMap& tb = tbs.m[0];
    int rc;
    if (!a_mn_major) rc = make_map(&ta, bf, A, K, M, batch, lda, strideA, BK, BM);
    else rc = make_map(&ta, bf, A, M, K, batch, lda, strideA, 64, BK);
    if (rc) return rc;
    if (!b_mn_major) rc = make_map(&tb, bf, B, K, N, batch, ldb, strideB, BK, bn);
    else rc = make_map(&tb, bf, B, N, K, batch, ldb, strideB, 64, BK);
    if (rc) return rc;
    rc = make_map(&td, bf, D, N, M, batch, ldd, strideD, EPI_COLS, 32);
    if (rc) return rc;
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.batch = batch;
    p.a_mn = a_mn_major; p.b_mn = b_mn_major;
    p.epilogue = epilogue; p.alpha = alpha;
    p.bias = bias; p.aux = aux; p.ldaux = ldaux; p.stride_aux = stride_aux;
    p.aux_out = aux_out; p.ldaux_out = ldaux_out; p.stride_aux_out = stride_aux_out;
    p.rows_per_rank = N;
    if (bf) {
        return bn == 256 ? launch<__nv_bfloat16, 256>(ta, tbs, td, p, sms, stream)
                         : launch<__nv_bfloat16, 128>(ta, tbs, td, p, sms, stream);
    }
    return bn == 256 ? launch<__half, 256>(ta, tbs, td, p, sms, stream)
                     : launch<__half, 128>(ta, tbs, td, p, sms, stream);
}

// Fused all-gather + GEMM (ZeRO-3): D = epi(A * W^T) where W [N, K] is row-sharded over `nranks` peers;
// shard_ptrs[r] points at rank r's [N / nranks, K] slice (peer-mapped device memory, leading dimension ldb).
extern "C" int dsb_gemm_gather_b(const void* A, void* const* shard_ptrs, int nranks, void* D, int M, int N, int K,
                                 long lda, long ldb, long ldd, int dtype, int epilogue, float alpha, const void* bias,
                                 const void* aux, long ldaux, void* stream_) {
    using namespace dsb;
    ensure_context_on_this_thread();
    constexpr int MAXR = 8;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (M <= 0 || N <= 0 || K <= 0) return 0;
    if (nranks < 1 || nranks > MAXR || (N % nranks)) return DSB_ERR_ARG;
    const int rows = N / nranks;
    if ((rows % 256) || (lda % 8) || (ldb % 8) || (ldd % 8)) return DSB_ERR_SIZE;
    if ((epilogue == DSB_EPI_BIAS_RESIDUAL || epilogue == DSB_EPI_DGELU) && (!aux || (ldaux % 8))) return DSB_ERR_ARG;
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int bf = dtype == DSB_DTYPE_BF16;
    CUtensorMap ta, td;
    BMapSet<MAXR> tbs;
    int rc = make_map(&ta, bf, A, K, M, 1, lda, (long)M * lda, BK, BM);
    if (rc) return rc;
    for (int r = 0; r < MAXR; ++r) {
        const void* base = shard_ptrs[r < nranks ? r : 0];
        rc = make_map(&tbs.m[r], bf, base, K, rows, 1, ldb, (long)rows * ldb, BK, 256);
        if (rc) return rc;
    }
    rc = make_map(&td, bf, D, N, M, 1, ldd, (long)M * ldd, EPI_COLS, 32);
    if (rc) return rc;
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.batch = 1;
    p.a_mn = 0; p.b_mn = 0;
    p.epilogue = epilogue; p.alpha = alpha;
    p.bias = bias; p.aux = aux; p.ldaux = ldaux; p.stride_aux = 0;
    p.aux_out = nullptr; p.ldaux_out = 0; p.stride_aux_out = 0;
    p.rows_per_rank = rows;
    p.ag_overlap = 0; p.me = 0; p.nranks = nranks; p.gathered = nullptr; p.panel_ready = nullptr;
    p.rs_mode = 0;
    for (int r = 0; r < MAXR; ++r) p.peer_shards[r] = nullptr;
    if (bf) return launch<__nv_bfloat16, 256, MAXR>(ta, tbs, td, p, sms, stream);
    return launch<__half, 256, MAXR>(ta, tbs, td, p, sms, stream);
}

// Overlapped all-gather + GEMM: D = epi(alpha * A @ gather(B)^T) where B [N, K] is row-sharded (dense shards,
// N / nranks rows each, a multiple of 256) over `nranks` peers. Inside ONE kernel, two warps of every CTA stream the
// remote shards over NVLink into `gathered` ([N, K], local scratch) panel by panel while the tensor pipe works on the
// panels that are already local (own shard first); a panel's tiles start as soon as its flag in `panel_flags`
// (N / 256 ints, zeroed here) shows every CTA has delivered its share. Each weight byte crosses NVLink once
// (dsb_gemm_gather_b re-reads remote tiles for every M block). Callers order the peers' shard writes before the launch.
extern "C" int dsb_gemm_ag_overlap(const void* A, void* const* shard_ptrs, int nranks, int me, void* gathered,
                                   int* panel_flags, void* D, int M, int N, int K, long lda, long ldd, int dtype,
                                   int epilogue, float alpha, const void* bias, const void* aux, long ldaux,
                                   void* stream_) {
    using namespace dsb;
    ensure_context_on_this_thread();
    constexpr int MAXR = 8;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (M <= 0 || N <= 0 || K <= 0) return 0;
    if (nranks < 1 || nranks > MAXR || me < 0 || me >= nranks || (N % nranks) || !gathered || !panel_flags)
        return DSB_ERR_ARG;
    const int rows = N / nranks;
    if ((rows % 256) || (lda % 8) || (K % 8) || (ldd % 8)) return DSB_ERR_SIZE;
    if ((epilogue == DSB_EPI_BIAS_RESIDUAL || epilogue == DSB_EPI_DGELU) && (!aux || (ldaux % 8))) return DSB_ERR_ARG;
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int bf = dtype == DSB_DTYPE_BF16;
    CUtensorMap ta, td;
    BMapSet<MAXR> tbs;
    int rc = make_map(&ta, bf, A, K, M, 1, lda, (long)M * lda, BK, BM);
    if (rc) return rc;
    rc = make_map(&tbs.m[0], bf, gathered, K, N, 1, K, (long)N * K, BK, 256);
    if (rc) return rc;
    rc = make_map(&tbs.m[1], bf, shard_ptrs[me], K, rows, 1, K, (long)rows * K, BK, 256);
    if (rc) return rc;
    for (int r = 2; r < MAXR; ++r) tbs.m[r] = tbs.m[0];
    rc = make_map(&td, bf, D, N, M, 1, ldd, (long)M * ldd, EPI_COLS, 32);
    if (rc) return rc;
    cudaError_t e = cudaMemsetAsync(panel_flags, 0, sizeof(int) * (N / 256), stream);
    if (e != cudaSuccess) return -(int)e;
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.batch = 1;
    p.a_mn = 0; p.b_mn = 0;
    p.epilogue = epilogue; p.alpha = alpha;
    p.bias = bias; p.aux = aux; p.ldaux = ldaux; p.stride_aux = 0;
    p.aux_out = nullptr; p.ldaux_out = 0; p.stride_aux_out = 0;
    p.rows_per_rank = rows;
    p.ag_overlap = 1; p.me = me; p.nranks = nranks; p.gathered = gathered; p.panel_ready = panel_flags;
    p.rs_mode = 0;
    for (int r = 0; r < MAXR; ++r) p.peer_shards[r] = shard_ptrs[r < nranks ? r : 0];
    if (bf) return launch<__nv_bfloat16, 256, MAXR>(ta, tbs, td, p, sms, stream);
    return launch<__half, 256, MAXR>(ta, tbs, td, p, sms, stream);
}

// Overlapped all-gather + GEMM for the DGRAD of a row-sharded weight: D [M, N] = alpha * A [M, K] * W where
// W [K, N] (nn.Linear weight: K = out features, N = in features) is row-sharded over `nranks` peers (dense shards,
// K / nranks rows each, a multiple of 256). The shards split the REDUCTION dimension, so every output tile needs all of
// them: the k loop of every tile starts with the own shard's K range and then walks the remote ranges in the order
// in which two copier warps per CTA stream them (256-row panels) into `gathered` ([K, N], local scratch);
// `panel_flags`: K / 256 ints, zeroed here. Each weight byte crosses NVLink once.
extern "C" int dsb_gemm_ag_overlap_k(const void* A, void* const* shard_ptrs, int nranks, int me, void* gathered,
                                     int* panel_flags, void* D, int M, int N, int K, long lda, long ldd, int dtype,
                                     float alpha, void* stream_) {
    using namespace dsb;
    ensure_context_on_this_thread();
    constexpr int MAXR = 8;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (M <= 0 || N <= 0 || K <= 0) return 0;
    if (nranks < 1 || nranks > MAXR || me < 0 || me >= nranks || (K % nranks) || !gathered || !panel_flags)
        return DSB_ERR_ARG;
    const int rows = K / nranks;
    if ((rows % 256) || (lda % 8) || (N % 8) || (ldd % 8)) return DSB_ERR_SIZE;
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int bf = dtype == DSB_DTYPE_BF16;
    CUtensorMap ta, td;
    BMapSet<MAXR> tbs;
    int rc = make_map(&ta, bf, A, K, M, 1, lda, (long)M * lda, BK, BM);
    if (rc) return rc;
    rc = make_map(&tbs.m[0], bf, gathered, N, K, 1, N, (long)K * N, 64, BK);
    if (rc) return rc;
    rc = make_map(&tbs.m[1], bf, shard_ptrs[me], N, rows, 1, N, (long)rows * N, 64, BK);
    if (rc) return rc;
    for (int r = 2; r < MAXR; ++r) tbs.m[r] = tbs.m[0];
    rc = make_map(&td, bf, D, N, M, 1, ldd, (long)M * ldd, EPI_COLS, 32);
    if (rc) return rc;
    cudaError_t e = cudaMemsetAsync(panel_flags, 0, sizeof(int) * (K / 256), stream);
    if (e != cudaSuccess) return -(int)e;
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.batch = 1;
    p.a_mn = 0; p.b_mn = 1;
    p.epilogue = DSB_EPI_NONE; p.alpha = alpha;
    p.bias = nullptr; p.aux = D; p.ldaux = ldd; p.stride_aux = 0;   // pass 1 accumulates onto pass 0's D
    p.aux_out = nullptr; p.ldaux_out = 0; p.stride_aux_out = 0;
    p.rows_per_rank = rows;
    p.ag_overlap = 2; p.me = me; p.nranks = nranks; p.gathered = gathered; p.panel_ready = panel_flags;
    p.rs_mode = 0;
    for (int r = 0; r < MAXR; ++r) p.peer_shards[r] = shard_ptrs[r < nranks ? r : 0];
    if (bf) return launch<__nv_bfloat16, 256, MAXR>(ta, tbs, td, p, sms, stream);
    return launch<__half, 256, MAXR>(ta, tbs, td, p, sms, stream);
}

// GEMM + reduce-scatter in ONE kernel (ZeRO-3 weight gradient): D = alpha * op(A) op(B)^T with D [M, N] row-owned by
// `nranks` ranks (M / nranks rows each, a multiple of 128). Every rank computes its full local product; each output
// tile is TMA-stored straight into the owner's staging buffer over NVLink while later tiles are still in the tensor
// pipe (remote owners first); after a cross-GPU barrier every rank sums the nranks slots that landed in ITS staging
// into `out` ([M / nranks, N]) times `scale`. `stage_ptrs[r]`: rank r's staging buffer ([nranks, M / nranks, N] of
// the GEMM dtype, peer-mapped; callers alternate between two such buffers from call to call so a rank may still be
// reducing the previous call's slots). `flag_bases[r]`: rank r's symmetric-heap flag block. `sync2`: two ints of
// local scratch. Launched cooperatively: CTAs wait for each other.
extern "C" int dsb_gemm_reduce_scatter(const void* A, const void* B, void* out, void* const* stage_ptrs,
                                       void* const* flag_bases, int nranks, int me, int* sync2, int M, int N, int K,
                                       long lda, long ldb, int a_mn_major, int b_mn_major, int dtype, float alpha,
                                       float scale, unsigned int epoch, void* stream_) {
    using namespace dsb;
    ensure_context_on_this_thread();
    constexpr int MAXR = 8;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (M <= 0 || N <= 0 || K <= 0) return 0;
    if (nranks < 1 || nranks > MAXR || me < 0 || me >= nranks || (M % nranks) || !sync2 || !out) return DSB_ERR_ARG;
    const int rows = M / nranks;
    if ((rows % BM) || (lda % 8) || (ldb % 8) || (N % 8)) return DSB_ERR_SIZE;
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const int bf = dtype == DSB_DTYPE_BF16;
    CUtensorMap ta, td;
    BMapSet<MAXR> tbs;
    int rc;
    if (!a_mn_major) rc = make_map(&ta, bf, A, K, M, 1, lda, (long)M * lda, BK, BM);
    else rc = make_map(&ta, bf, A, M, K, 1, lda, (long)K * lda, 64, BK);
    if (rc) return rc;
    for (int r = 0; r < MAXR; ++r) {
        if (r == me || r >= nranks) {
            if (!b_mn_major) rc = make_map(&tbs.m[r], bf, B, K, N, 1, ldb, (long)N * ldb, BK, 256);
            else rc = make_map(&tbs.m[r], bf, B, N, K, 1, ldb, (long)K * ldb, 64, BK);
        } else {
            rc = make_map(&tbs.m[r], bf, stage_ptrs[r], N, (uint64_t)nranks * rows, 1, N, (long)nranks * rows * N,
                          EPI_COLS, 32);
        }
        if (rc) return rc;
    }
    rc = make_map(&td, bf, stage_ptrs[me], N, (uint64_t)nranks * rows, 1, N, (long)nranks * rows * N, EPI_COLS, 32);
    if (rc) return rc;
    cudaError_t e = cudaMemsetAsync(sync2, 0, 2 * sizeof(int), stream);
    if (e != cudaSuccess) return -(int)e;
    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.batch = 1;
    p.a_mn = a_mn_major; p.b_mn = b_mn_major;
    p.epilogue = DSB_EPI_NONE; p.alpha = alpha;
    p.bias = nullptr; p.aux = nullptr; p.ldaux = 0; p.stride_aux = 0;
    p.aux_out = nullptr; p.ldaux_out = 0; p.stride_aux_out = 0;
    p.rows_per_rank = rows;
    p.ag_overlap = 0; p.me = me; p.nranks = nranks; p.gathered = nullptr; p.panel_ready = nullptr;
    for (int r = 0; r < MAXR; ++r) p.peer_shards[r] = flag_bases[r < nranks ? r : 0];
    p.rs_mode = 1; p.rs_epoch = epoch; p.rs_scale = scale; p.rs_out = out; p.rs_stage = stage_ptrs[me];
    p.rs_sync = sync2;
    using Cfg = GemmCfg<256>;
    const void* kern = bf ? (const void*)gemm_tcgen05_kernel<__nv_bfloat16, 256, MAXR>
                          : (const void*)gemm_tcgen05_kernel<__half, 256, MAXR>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
    if (e != cudaSuccess) return -(int)e;
    const long tiles = (long)(M / BM) * ((N + 255) / 256);
    const int grid = (int)(tiles < sms ? tiles : sms);
    void* args[] = {&ta, &tbs, &td, &p};
    e = cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(NUM_THREADS), args, Cfg::SMEM_BYTES, stream);
    return e == cudaSuccess ? 0 : -(int)e;
}
