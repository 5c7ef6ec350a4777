// This file holds (1) the C entry points of ALL flash-attention kernels (tensor maps, argument checks, variant
// dispatch; the default kernels are the second generation in flash_attn_v2.cu) and (2) the FIRST-generation kernels,
// kept selectable for comparison (dsb_flash_attn_fwd_variant 64 / 128, dsb_flash_attn_bwd_variant 1):
//
// Causal / full flash-attention forward on tcgen05 (head dim 128, fp16 / bf16), one CTA per 128 query rows of one
// (batch, head).
//
//   warp 0 (one elected lane): TMA producer. Q tile once, then K_j / V_j tiles (128 keys each) into a 2-stage ring.
//   warp 1 (one elected lane): issues S_j = Q K_j^T (UMMA 128 x 128 x 16, both operands K-major) into one of two
//                              TMEM S buffers and O += P_j V_j (A = P from shared memory, B = V read MN-major from the
//                              tile exactly as TMA delivered it). S_{j+1} is issued before P_j is awaited, so the
//                              tensor pipe computes the next scores while the softmax warps work on the current ones.
//   warp 2: TMEM allocation (512 columns: S0 | S1 | O).
//   warps 4-7: online softmax, thread t owns query row t (TMEM lane t). Two trips over the 128 scores of the row
//              (max, then exp2 / sum / fp16 P written into 128-byte-swizzled shared memory), rescale of the O
//              accumulator through tcgen05.ld / tcgen05.st, final 1/l scaling, O staged in shared memory and stored
//              with TMA, log-sum-exp written for the backward pass.
//
// Q, K, V, O are addressed as [B, S, H, D] with arbitrary batch / sequence / head strides (unit last stride), i.e. the
// q / k buffers the rotary kernel writes and the V slice of the fused QKV activation are read in place and the
// context comes out in the [B, S, H * D] layout the output projection consumes.
// Reference: the score / softmax / context chain of csrc/transformer/ds_transformer_cuda.cpp:186-214 (strided batch
// GEMMs + attn_softmax + dropout), which materialises the [B, H, S, S] probabilities; this kernel never does.
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"
#include "../common/host_ctx.h"
#include "flash_v2.cuh"
#include <type_traits>

namespace dsb {
namespace fa {

constexpr int BQ = 128;
constexpr int BKV = 128;
constexpr int HD = 128;
constexpr int TILE_BYTES = 128 * HD * 2;       // 32 KiB: two 64-column chunks of [128 rows x 128 B]
constexpr int CHUNK_BYTES = 128 * 128;         // 16 KiB
constexpr int KV_STAGES = 2;
constexpr int OFF_Q = 0;
constexpr int OFF_K = OFF_Q + TILE_BYTES;
constexpr int OFF_V = OFF_K + KV_STAGES * TILE_BYTES;
constexpr int OFF_P = OFF_V + KV_STAGES * TILE_BYTES;
constexpr int OFF_BAR = OFF_P + TILE_BYTES;
constexpr int OFF_XCH = OFF_BAR + 256;
constexpr int SMEM_BYTES = OFF_XCH + 2048 + 1024 /*align slack*/;
constexpr int TMEM_COLS = 512;
constexpr int TM_S = 0;      // S buffers at columns 0 and 128
constexpr int TM_O = 256;

DSB_DEVICE void tma_load_4d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2),
        "r"(c3)
        : "memory");
}
DSB_DEVICE void tma_store_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                     reinterpret_cast<uint64_t>(m)),
                 "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
DSB_DEVICE void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
DSB_DEVICE void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
DSB_DEVICE float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <typename T>
__global__ void __launch_bounds__(384, 1)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                 float* __restrict__ lse, int S, int H, float scale_log2, int causal) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem + OFF_Q;
    uint8_t* sK = smem + OFF_K;
    uint8_t* sV = smem + OFF_V;
    uint8_t* sP = smem + OFF_P;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* q_full = bars;            // [1]
    uint64_t* kv_full = bars + 1;       // [2]
    uint64_t* kv_empty = bars + 3;      // [2]
    uint64_t* s_full = bars + 5;        // [2]
    uint64_t* s_free = bars + 7;        // [2]
    uint64_t* p_ready = bars + 9;       // [1]
    uint64_t* pv_done = bars + 10;      // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
    float* xch = reinterpret_cast<float*>(smem + OFF_XCH);   // [2 parities][2 halves][128 rows] partial row maxima / sums

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int qb = gridDim.x - 1 - blockIdx.x;   // long (late) query blocks first
    const int b = blockIdx.y / H, h = blockIdx.y - b * H;
    const int q0 = qb * BQ;
    const int nkv = causal ? qb + 1 : S / BKV;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmO);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(q_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&kv_full[i], 1);
            mbar_init(&kv_empty[i], 1);
            mbar_init(&s_full[i], 1);
            mbar_init(&s_free[i], 8);
        }
        mbar_init(p_ready, 8);
        mbar_init(pv_done, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (elect_one()) {
            mbar_arrive_expect_tx(q_full, TILE_BYTES);
            tma_load_4d(sQ, &tmQ, q_full, 0, h, q0, b);
            tma_load_4d(sQ + CHUNK_BYTES, &tmQ, q_full, 64, h, q0, b);
            for (int j = 0; j < nkv; ++j) {
                const int st = j & 1;
                if (j >= 2) mbar_wait(&kv_empty[st], ((j >> 1) - 1) & 1);
                mbar_arrive_expect_tx(&kv_full[st], 2 * TILE_BYTES);
                uint8_t* k = sK + st * TILE_BYTES;
                uint8_t* v = sV + st * TILE_BYTES;
                tma_load_4d(k, &tmK, &kv_full[st], 0, h, j * BKV, b);
                tma_load_4d(k + CHUNK_BYTES, &tmK, &kv_full[st], 64, h, j * BKV, b);
                tma_load_4d(v, &tmV, &kv_full[st], 0, h, j * BKV, b);
                tma_load_4d(v + CHUNK_BYTES, &tmV, &kv_full[st], 64, h, j * BKV, b);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (elect_one()) {
            constexpr int is_bf16 = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
            const uint32_t idesc_s = make_idesc_f16(BQ, BKV, is_bf16, 0, 0);
            const uint32_t idesc_pv = make_idesc_f16(BQ, HD, is_bf16, 0, 1);
            constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t q_lo = ((smem_u32(sQ) & 0x3FFFF) >> 4) | (1u << 16);
            const uint32_t p_lo = ((smem_u32(sP) & 0x3FFFF) >> 4) | (1u << 16);
            auto issue_s = [&](int j) {
                const uint32_t k_lo = (((smem_u32(sK) + (j & 1) * TILE_BYTES) & 0x3FFFF) >> 4) | (1u << 16);
                const uint32_t d = tmem_base + TM_S + (j & 1) * BKV;
#pragma unroll
                for (int k = 0; k < HD / 16; ++k) {
                    const uint32_t off = (k >> 2) * (CHUNK_BYTES >> 4) + (k & 3) * (32u >> 4);
                    umma_f16_lohi(d, q_lo + off, k_lo + off, DESC_HI, idesc_s, k != 0);
                }
                umma_commit(&s_full[j & 1]);
            };
            auto issue_pv = [&](int j) {
                // B = V tile read MN-major: 64-element d chunks CHUNK_BYTES apart (LBO), 16 key rows per k-step
                const uint32_t v_lo = (((smem_u32(sV) + (j & 1) * TILE_BYTES) & 0x3FFFF) >> 4) |
                                      ((uint32_t)(CHUNK_BYTES >> 4) << 16);
#pragma unroll
                for (int k = 0; k < BKV / 16; ++k) {
                    const uint32_t a_off = (k >> 2) * (CHUNK_BYTES >> 4) + (k & 3) * (32u >> 4);
                    umma_f16_lohi(tmem_base + TM_O, p_lo + a_off, v_lo + k * (2048u >> 4), DESC_HI, idesc_pv,
                                  (j | k) != 0);
                }
                umma_commit(pv_done);
                umma_commit(&kv_empty[j & 1]);
            };
            mbar_wait(q_full, 0);
            mbar_wait(&kv_full[0], 0);
            tc_fence_after();
            issue_s(0);
            for (int j = 0; j < nkv; ++j) {
                if (j + 1 < nkv) {
                    const int n = j + 1;
                    mbar_wait(&kv_full[n & 1], (n >> 1) & 1);
                    if (n >= 2) mbar_wait(&s_free[n & 1], ((n >> 1) - 1) & 1);
                    tc_fence_after();
                    issue_s(n);
                }
                mbar_wait(p_ready, j & 1);
                tc_fence_after();
                issue_pv(j);
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ softmax / correction / epilogue
        // Eight warps: warp w owns TMEM lane quarter w % 4 (32 query rows) and one half of the 128 columns, so a row
        // is shared by two threads; they exchange their partial row maximum through shared memory once per key block
        // and their partial row sums once at the end (both use the same running maximum).
        const int q = warp & 3;
        const int half = (warp - 4) >> 2;
        const int c_lo = half * 2;
        const int row = q * 32 + lane;
        const uint32_t lane_field = static_cast<uint32_t>(q * 32) << 16;
        uint8_t* p_row = sP + row * 128;
        const int sw = row & 7;
        float m_run = -INFINITY, l_run = 0.f;
        for (int j = 0; j < nkv; ++j) {
            const int buf = j & 1;
            const bool diag = causal && j == nkv - 1;
            mbar_wait(&s_full[buf], (j >> 1) & 1);
            tc_fence_after();
            const uint32_t t_s = tmem_base + lane_field + TM_S + buf * BKV;
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = c_lo; c < c_lo + 2; ++c) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(t_s + c * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float v = __uint_as_float(r[i]);
                    if (diag && c * 32 + i > row) v = -INFINITY;
                    mx = fmaxf(mx, v);
                }
            }
            float* xm = xch + (j & 1) * 256;
            xm[half * 128 + row] = mx;
            named_bar_sync(2, 256);
            mx = fmaxf(mx, xm[(half ^ 1) * 128 + row]);
            const float m_new = fmaxf(m_run, mx * scale_log2);
            const float alpha = ex2(m_run - m_new);
            if (j > 0) {
                // O and the P buffer are free once the previous P V product has completed
                mbar_wait(pv_done, (j - 1) & 1);
                tc_fence_after();
                const uint32_t t_o = tmem_base + lane_field + TM_O;
                if (!__all_sync(0xffffffffu, alpha == 1.f)) {   // the running maxima of these 32 rows moved
#pragma unroll 1
                    for (int c = c_lo; c < c_lo + 2; ++c) {
                        uint32_t r[32];
                        tmem_ld_32x32b_x32(t_o + c * 32, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        tmem_st_32x32b_x32(t_o + c * 32, r);
                    }
                    tmem_st_wait();
                }
            }
            float rs = 0.f;
#pragma unroll 1
            for (int c = c_lo; c < c_lo + 2; ++c) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(t_s + c * 32, r);
                tmem_ld_wait();
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float p0 = ex2(__uint_as_float(r[i]) * scale_log2 - m_new);
                    float p1 = ex2(__uint_as_float(r[i + 1]) * scale_log2 - m_new);
                    if (diag && c * 32 + i > row) p0 = 0.f;
                    if (diag && c * 32 + i + 1 > row) p1 = 0.f;
                    rs += p0 + p1;
                    pk[i >> 1] = Cvt<T>::pack2(p0, p1);
                }
                uint8_t* dst = p_row + (c >> 1) * CHUNK_BYTES;
                const int u0 = (c & 1) * 4;
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    *reinterpret_cast<uint4*>(dst + (((u0 + u) ^ sw) << 4)) =
                        make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
            }
            l_run = l_run * alpha + rs;
            m_run = m_new;
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&s_free[buf]);
                mbar_arrive(p_ready);
            }
        }
        // epilogue: O / l -> shared (same swizzled layout as P) -> TMA store; log-sum-exp for the backward pass
        mbar_wait(pv_done, (nkv - 1) & 1);
        tc_fence_after();
        float* xl = xch + (nkv & 1) * 256;
        xl[half * 128 + row] = l_run;
        named_bar_sync(2, 256);
        l_run += xl[(half ^ 1) * 128 + row];
        const float inv_l = 1.f / l_run;
        const uint32_t t_o = tmem_base + lane_field + TM_O;
#pragma unroll 1
        for (int c = c_lo; c < c_lo + 2; ++c) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(t_o + c * 32, r);
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2)
                pk[i >> 1] = Cvt<T>::pack2(__uint_as_float(r[i]) * inv_l, __uint_as_float(r[i + 1]) * inv_l);
            uint8_t* dst = p_row + (c >> 1) * CHUNK_BYTES;
            const int u0 = (c & 1) * 4;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                *reinterpret_cast<uint4*>(dst + (((u0 + u) ^ sw) << 4)) =
                    make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
        }
        if (lse != nullptr && half == 0)
            lse[((long)blockIdx.y) * S + q0 + row] = (m_run + log2f(l_run)) * 0.6931471805599453f;
        fence_proxy_async_smem();
        named_bar_sync(1, 256);
        if (warp == 4 && lane == 0) {
            tma_store_4d(&tmO, sP, 0, h, q0, b);
            tma_store_4d(&tmO, sP + CHUNK_BYTES, 64, h, q0, b);
            tma_store_commit();
            tma_store_wait<0>();
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ forward, 64-key blocks
// Same algorithm with 64 keys per step: 112 KiB of shared memory and 256 TMEM columns per CTA, so TWO CTAs share an SM
// and one CTA's S -> softmax -> P -> PV chain runs in the shadow of the other's (with one CTA per SM that chain is
// latency-bound: about 2 us per 128-key block against 0.55 us of tensor time).
constexpr int K64 = 64;
constexpr int KV64_TILE = K64 * HD * 2;          // 16 KiB: two 64-column chunks of [64 rows x 128 B]
constexpr int KV64_CHUNK = K64 * 128;            // 8 KiB
constexpr int P64_BYTES = BQ * K64 * 2;          // 16 KiB: one chunk of [128 rows x 128 B]
constexpr int F_OFF_Q = 0;
constexpr int F_OFF_K = F_OFF_Q + TILE_BYTES;
constexpr int F_OFF_V = F_OFF_K + 2 * KV64_TILE;
constexpr int F_OFF_P = F_OFF_V + 2 * KV64_TILE;
constexpr int F_OFF_BAR = F_OFF_P + P64_BYTES;
constexpr int F_SMEM_BYTES = F_OFF_BAR + 256 + 512;   // 115456 B: two CTAs fit in the 228 KiB of an SM
constexpr int F_TMEM_COLS = 256;
constexpr int F_TM_O = 128;                      // S buffers at columns 0 and 64

// One 64-key block of the online softmax for one row: maximum, rescale factor, probabilities packed to 16 bit and
// their sum. MASK = false is the code for blocks no row of the warp has to mask (all but the last two of a causal
// row block): without the per-score compares it is about half the instructions.
template <typename T, bool MASK>
DSB_DEVICE void softmax_block64(const uint32_t (&r0)[32], const uint32_t (&r1)[32], int lim, float scale_log2,
                                float m_run, float& m_new, float& alpha, float& rs, uint32_t (&pk)[32]) {
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        if (!MASK || i <= lim) mx = fmaxf(mx, __uint_as_float(r0[i]));
        if (!MASK || 32 + i <= lim) mx = fmaxf(mx, __uint_as_float(r1[i]));
    }
    m_new = fmaxf(m_run, mx * scale_log2);
    alpha = ex2(m_run - m_new);
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
        float a0 = ex2(__uint_as_float(r0[i]) * scale_log2 - m_new);
        float a1 = ex2(__uint_as_float(r0[i + 1]) * scale_log2 - m_new);
        float b0 = ex2(__uint_as_float(r1[i]) * scale_log2 - m_new);
        float b1 = ex2(__uint_as_float(r1[i + 1]) * scale_log2 - m_new);
        if (MASK) {
            if (i > lim) a0 = 0.f;
            if (i + 1 > lim) a1 = 0.f;
            if (32 + i > lim) b0 = 0.f;
            if (33 + i > lim) b1 = 0.f;
        }
        s0 += a0 + a1;
        s1 += b0 + b1;
        pk[i >> 1] = Cvt<T>::pack2(a0, a1);
        pk[16 + (i >> 1)] = Cvt<T>::pack2(b0, b1);
    }
    rs = s0 + s1;
}

template <typename T>
__global__ void __launch_bounds__(256, 2)
flash_fwd64_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                   float* __restrict__ lse, int S, int H, float scale_log2, int causal, long long* dbg) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    if (smem - smem_raw > 512) __trap();   // the launch reserves only 512 bytes of alignment slack
    uint8_t* sQ = smem + F_OFF_Q;
    uint8_t* sK = smem + F_OFF_K;
    uint8_t* sV = smem + F_OFF_V;
    uint8_t* sP = smem + F_OFF_P;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + F_OFF_BAR);
    uint64_t* q_full = bars;
    uint64_t* k_full = bars + 1;        // K and V have their own rings: a K tile is free as soon as its S product
    uint64_t* k_empty = bars + 3;       // has completed, one block before the matching V tile
    uint64_t* v_full = bars + 11;
    uint64_t* v_empty = bars + 13;
    uint64_t* s_full = bars + 5;
    uint64_t* s_free = bars + 7;
    uint64_t* p_ready = bars + 9;
    uint64_t* pv_done = bars + 10;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int qb = gridDim.x - 1 - blockIdx.x;
    const int b = blockIdx.y / H, h = blockIdx.y - b * H;
    const int q0 = qb * BQ;
    const int nkv = causal ? (q0 + BQ) / K64 : S / K64;
    const bool tm = dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0;
    long long t0 = 0, acc[6] = {0, 0, 0, 0, 0, 0};
#define TICK(n) if (tm) { long long t1 = clock64(); acc[n] += t1 - t0; t0 = t1; }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmO);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(q_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&k_full[i], 1);
            mbar_init(&k_empty[i], 1);
            mbar_init(&v_full[i], 1);
            mbar_init(&v_empty[i], 1);
            mbar_init(&s_full[i], 1);
            mbar_init(&s_free[i], 4);
        }
        mbar_init(p_ready, 4);
        mbar_init(pv_done, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<F_TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (elect_one()) {
            mbar_arrive_expect_tx(q_full, TILE_BYTES);
            tma_load_4d(sQ, &tmQ, q_full, 0, h, q0, b);
            tma_load_4d(sQ + CHUNK_BYTES, &tmQ, q_full, 64, h, q0, b);
            // With one ring for both, tile j + 2 could only be requested after P V of block j and the ~1800-cycle TMA
            // round trip sat on the critical path of every other block (measured with the clock64 timers below: the
            // MMA thread waited 1790 cycles per block for its operands).
            for (int j = 0; j < nkv; ++j) {
                const int st = j & 1;
                if (j >= 2) mbar_wait(&k_empty[st], ((j >> 1) - 1) & 1);
                mbar_arrive_expect_tx(&k_full[st], KV64_TILE);
                uint8_t* kk = sK + st * KV64_TILE;
                tma_load_4d(kk, &tmK, &k_full[st], 0, h, j * K64, b);
                tma_load_4d(kk + KV64_CHUNK, &tmK, &k_full[st], 64, h, j * K64, b);
            }
        }
    } else if (warp == 3) {
        if (elect_one()) {
            for (int j = 0; j < nkv; ++j) {
                const int st = j & 1;
                if (j >= 2) mbar_wait(&v_empty[st], ((j >> 1) - 1) & 1);
                mbar_arrive_expect_tx(&v_full[st], KV64_TILE);
                uint8_t* v = sV + st * KV64_TILE;
                tma_load_4d(v, &tmV, &v_full[st], 0, h, j * K64, b);
                tma_load_4d(v + KV64_CHUNK, &tmV, &v_full[st], 64, h, j * K64, b);
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr int is_bf16 = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
            const uint32_t idesc_s = make_idesc_f16(BQ, K64, is_bf16, 0, 0);
            const uint32_t idesc_pv = make_idesc_f16(BQ, HD, is_bf16, 0, 1);
            constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t q_lo = ((smem_u32(sQ) & 0x3FFFF) >> 4) | (1u << 16);
            const uint32_t p_lo = ((smem_u32(sP) & 0x3FFFF) >> 4) | (1u << 16);
            auto issue_s = [&](int j) {
                const uint32_t k_lo = (((smem_u32(sK) + (j & 1) * KV64_TILE) & 0x3FFFF) >> 4) | (1u << 16);
                const uint32_t d = tmem_base + (j & 1) * K64;
#pragma unroll
                for (int k = 0; k < HD / 16; ++k)
                    umma_f16_lohi(d, q_lo + (k >> 2) * (CHUNK_BYTES >> 4) + (k & 3) * 2,
                                  k_lo + (k >> 2) * (KV64_CHUNK >> 4) + (k & 3) * 2, DESC_HI, idesc_s, k != 0);
                umma_commit(&s_full[j & 1]);
                umma_commit(&k_empty[j & 1]);
            };
            auto issue_pv = [&](int j) {
                const uint32_t v_lo = (((smem_u32(sV) + (j & 1) * KV64_TILE) & 0x3FFFF) >> 4) |
                                      ((uint32_t)(KV64_CHUNK >> 4) << 16);
#pragma unroll
                for (int k = 0; k < K64 / 16; ++k)
                    umma_f16_lohi(tmem_base + F_TM_O, p_lo + k * 2, v_lo + k * (2048u >> 4), DESC_HI, idesc_pv,
                                  (j | k) != 0);
                umma_commit(pv_done);
                umma_commit(&v_empty[j & 1]);
            };
            mbar_wait(q_full, 0);
            mbar_wait(&k_full[0], 0);
            tc_fence_after();
            issue_s(0);
            t0 = clock64();
            for (int j = 0; j < nkv; ++j) {
                if (j + 1 < nkv) {
                    const int n = j + 1;
                    mbar_wait(&k_full[n & 1], (n >> 1) & 1);
                    if (n >= 2) mbar_wait(&s_free[n & 1], ((n >> 1) - 1) & 1);
                    tc_fence_after();
                    TICK(0)
                    issue_s(n);
                    TICK(1)
                }
                mbar_wait(p_ready, j & 1);
                mbar_wait(&v_full[j & 1], (j >> 1) & 1);
                tc_fence_after();
                TICK(2)
                issue_pv(j);
                TICK(3)
            }
            if (tm) { for (int i = 0; i < 4; ++i) dbg[8 + i] = acc[i]; dbg[15] = nkv; }
        }
    } else if (warp >= 4) {
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const uint32_t lane_field = static_cast<uint32_t>(q * 32) << 16;
        const int sw = row & 7;
        float m_run = -INFINITY, l_run = 0.f;
        t0 = clock64();
        for (int j = 0; j < nkv; ++j) {
            const int buf = j & 1;
            const int lim = causal ? q0 + row - j * K64 : K64;   // columns <= lim of this block are visible
            const bool edge = lim < K64 - 1;
            mbar_wait(&s_full[buf], (j >> 1) & 1);
            tc_fence_after();
            TICK(0)
            const uint32_t t_s = tmem_base + lane_field + buf * K64;
            uint32_t r0[32], r1[32];
            tmem_ld_32x32b_x32(t_s, r0);
            tmem_ld_32x32b_x32(t_s + 32, r1);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_free[buf]);   // the scores are in registers: the buffer can be refilled
            TICK(1)
            float m_new, alpha, rs;
            uint32_t pk[32];
            if (__any_sync(0xffffffffu, edge))
                softmax_block64<T, true>(r0, r1, lim, scale_log2, m_run, m_new, alpha, rs, pk);
            else
                softmax_block64<T, false>(r0, r1, lim, scale_log2, m_run, m_new, alpha, rs, pk);
            TICK(2)
            if (j > 0) {
                mbar_wait(pv_done, (j - 1) & 1);   // O and the P buffer are free
                tc_fence_after();
                if (!__all_sync(0xffffffffu, alpha == 1.f)) {
                    const uint32_t t_o = tmem_base + lane_field + F_TM_O;
#pragma unroll 1
                    for (int c = 0; c < 4; ++c) {
                        uint32_t r[32];
                        tmem_ld_32x32b_x32(t_o + c * 32, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        tmem_st_32x32b_x32(t_o + c * 32, r);
                    }
                    tmem_st_wait();
                }
            }
            TICK(3)
            uint8_t* dst = sP + row * 128;
#pragma unroll
            for (int u = 0; u < 8; ++u)
                *reinterpret_cast<uint4*>(dst + ((u ^ sw) << 4)) =
                    make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
            l_run = l_run * alpha + rs;
            m_run = m_new;
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_ready);
            TICK(4)
        }
        if (tm && warp == 4 && lane == 0) for (int i = 0; i < 5; ++i) dbg[i] = acc[i];
        mbar_wait(pv_done, (nkv - 1) & 1);
        tc_fence_after();
        const float inv_l = 1.f / l_run;
        const uint32_t t_o = tmem_base + lane_field + F_TM_O;
        uint8_t* o_row = sQ + row * 128;   // the Q tile is no longer needed: stage O there
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(t_o + c * 32, r);
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2)
                pk[i >> 1] = Cvt<T>::pack2(__uint_as_float(r[i]) * inv_l, __uint_as_float(r[i + 1]) * inv_l);
            uint8_t* dst = o_row + (c >> 1) * CHUNK_BYTES;
            const int u0 = (c & 1) * 4;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                *reinterpret_cast<uint4*>(dst + (((u0 + u) ^ sw) << 4)) =
                    make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
        }
        if (lse != nullptr)
            lse[((long)blockIdx.y) * S + q0 + row] = (m_run + log2f(l_run)) * 0.6931471805599453f;
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (warp == 4 && lane == 0) {
            tma_store_4d(&tmO, sQ, 0, h, q0, b);
            tma_store_4d(&tmO, sQ + CHUNK_BYTES, 64, h, q0, b);
            tma_store_commit();
            tma_store_wait<0>();
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 2) tmem_dealloc<F_TMEM_COLS>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ backward
// delta[b, h, s] = sum_d dO[b, s, h, d] * O[b, s, h, d] (fp32); one warp per row of 128 elements.
template <typename T>
__global__ void __launch_bounds__(256) flash_delta_kernel(const T* __restrict__ dO, const T* __restrict__ O,
                                                          float* __restrict__ delta, int B, int S, int H, long do_sb,
                                                          long do_ss, long do_sh, long o_sb, long o_ss, long o_sh) {
    const long rows = (long)B * S * H;
    const int lane = threadIdx.x & 31;
    for (long r = (long)blockIdx.x * 8 + (threadIdx.x >> 5); r < rows; r += (long)gridDim.x * 8) {
        const int h = (int)(r % H);
        const long t = r / H;
        const int s = (int)(t % S);
        const int b = (int)(t / S);
        const uint2 a = *reinterpret_cast<const uint2*>(dO + b * do_sb + s * do_ss + h * do_sh + lane * 4);
        const uint2 c = *reinterpret_cast<const uint2*>(O + b * o_sb + s * o_ss + h * o_sh + lane * 4);
        const float2 a0 = Cvt<T>::unpack2(a.x), a1 = Cvt<T>::unpack2(a.y);
        const float2 c0 = Cvt<T>::unpack2(c.x), c1 = Cvt<T>::unpack2(c.y);
        float v = a0.x * c0.x + a0.y * c0.y + a1.x * c1.x + a1.y * c1.y;
        v = warp_sum(v);
        if (lane == 0) delta[((long)b * H + h) * S + s] = v;
    }
}

constexpr int B_OFF_K = 0;
constexpr int B_OFF_V = B_OFF_K + TILE_BYTES;
constexpr int B_OFF_Q = B_OFF_V + TILE_BYTES;
constexpr int B_OFF_DO = B_OFF_Q + TILE_BYTES;
constexpr int B_OFF_P = B_OFF_DO + TILE_BYTES;
constexpr int B_OFF_DS = B_OFF_P + TILE_BYTES;
constexpr int B_OFF_Q1 = B_OFF_DS + TILE_BYTES;   // second Q stage (the next query block is prefetched)
constexpr int B_OFF_BAR = B_OFF_Q1 + TILE_BYTES;
constexpr int B_SMEM_BYTES = B_OFF_BAR + 256 + 1024;
constexpr int TM_DP = 128, TM_DV = 256, TM_DK = 384;

DSB_DEVICE void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Bulk tensor reduction shared -> global (fp32 add), executed by the TMA unit / L2; replaces per-thread red.global.
DSB_DEVICE void tma_reduce_add_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                     reinterpret_cast<uint64_t>(m)),
                 "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}

// One CTA per (key block j, batch, head): K_j / V_j stay in shared memory, the CTA walks the query blocks i >= j
// (causal) and per block issues five UMMAs: S = Q K^T and dP = dO V^T (TMEM), then - once the softmax warps have
// turned them into P and dS in shared memory - dV += P^T dO, dK += dS^T Q (accumulated in TMEM over the whole walk;
// P and dS are read MN-major, i.e. transposed for free) and dQ_i = dS K into the TMEM columns S has vacated. dQ_i is
// added to an fp32 accumulator in global memory with vector reductions (other key blocks add to the same rows).
template <typename T>
__global__ void __launch_bounds__(384, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                 const __grid_constant__ CUtensorMap tmdK, const __grid_constant__ CUtensorMap tmdV,
                 const __grid_constant__ CUtensorMap tmDQ,
                 const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_acc, int S,
                 int H, float scale, float scale_log2, int causal, long long* dbg) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sK = smem + B_OFF_K;
    uint8_t* sV = smem + B_OFF_V;
    uint8_t* sQ0 = smem + B_OFF_Q;
    uint8_t* sQ1 = smem + B_OFF_Q1;
    uint8_t* sdO = smem + B_OFF_DO;
    uint8_t* sP = smem + B_OFF_P;
    uint8_t* sdS = smem + B_OFF_DS;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + B_OFF_BAR);
    uint64_t* kv_full = bars;          // K_j, V_j landed
    uint64_t* do_full = bars + 1;      // dO_i landed
    uint64_t* sdp_full = bars + 2;     // S and dP are in TMEM
    uint64_t* pds_ready = bars + 3;    // P and dS are in shared memory (8 warps)
    uint64_t* mma2_done = bars + 4;    // dV, dK, dQ UMMAs of this query block completed
    uint64_t* s_free = bars + 5;       // dQ drained from TMEM (8 warps)
    uint64_t* do_free = bars + 6;      // dV UMMAs (the last readers of dO_i) completed
    uint64_t* q_full = bars + 9;       // [2] Q ring: Q_{i+1} travels while block i is being worked on
    uint64_t* q_free = bars + 11;      // [2] dK UMMAs (the last readers of that Q stage) completed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int jb = blockIdx.x;
    const int b = blockIdx.y / H, h = blockIdx.y - b * H;
    const int nq = S / BQ;
    const int i0 = causal ? jb : 0;
    const int nit = nq - i0;
    const bool tm = dbg != nullptr && blockIdx.x == 0 && blockIdx.y == 0;
    long long t0 = 0, acc[6] = {0, 0, 0, 0, 0, 0};

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmdO);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(kv_full, 1);
        mbar_init(do_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&q_full[i], 1);
            mbar_init(&q_free[i], 1);
        }
        mbar_init(sdp_full, 1);
        mbar_init(pds_ready, 8);
        mbar_init(mma2_done, 1);
        mbar_init(s_free, 8);
        mbar_init(do_free, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (elect_one()) {
            mbar_arrive_expect_tx(kv_full, 2 * TILE_BYTES);
            tma_load_4d(sK, &tmK, kv_full, 0, h, jb * BKV, b);
            tma_load_4d(sK + CHUNK_BYTES, &tmK, kv_full, 64, h, jb * BKV, b);
            tma_load_4d(sV, &tmV, kv_full, 0, h, jb * BKV, b);
            tma_load_4d(sV + CHUNK_BYTES, &tmV, kv_full, 64, h, jb * BKV, b);
            for (int it = 0; it < nit; ++it) {   // Q ring, two stages
                const int st = it & 1;
                if (it >= 2) mbar_wait(&q_free[st], ((it >> 1) - 1) & 1);
                const int q0 = (i0 + it) * BQ;
                uint8_t* dst = st ? sQ1 : sQ0;
                mbar_arrive_expect_tx(&q_full[st], TILE_BYTES);
                tma_load_4d(dst, &tmQ, &q_full[st], 0, h, q0, b);
                tma_load_4d(dst + CHUNK_BYTES, &tmQ, &q_full[st], 64, h, q0, b);
            }
        }
    } else if (warp == 3) {
        if (elect_one()) {
            for (int it = 0; it < nit; ++it) {   // dO: one stage, re-requested as soon as dV of the block has completed
                if (it > 0) mbar_wait(do_free, (it - 1) & 1);
                const int q0 = (i0 + it) * BQ;
                mbar_arrive_expect_tx(do_full, TILE_BYTES);
                tma_load_4d(sdO, &tmdO, do_full, 0, h, q0, b);
                tma_load_4d(sdO + CHUNK_BYTES, &tmdO, do_full, 64, h, q0, b);
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            constexpr int is_bf16 = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
            const uint32_t id_kk = make_idesc_f16(128, 128, is_bf16, 0, 0);   // both operands K-major
            const uint32_t id_mm = make_idesc_f16(128, 128, is_bf16, 1, 1);   // both MN-major (transposed reads)
            const uint32_t id_km = make_idesc_f16(128, 128, is_bf16, 0, 1);
            constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
            constexpr uint32_t LBO_MN = (uint32_t)(CHUNK_BYTES >> 4) << 16;
            auto kmaj = [](const uint8_t* p) { return ((smem_u32(p) & 0x3FFFF) >> 4) | (1u << 16); };
            auto mnmaj = [](const uint8_t* p) { return ((smem_u32(p) & 0x3FFFF) >> 4) | LBO_MN; };
            auto koff = [](int k) { return (uint32_t)((k >> 2) * (CHUNK_BYTES >> 4) + (k & 3) * 2); };
            mbar_wait(kv_full, 0);
            t0 = clock64();
            for (int it = 0; it < nit; ++it) {
                const uint8_t* sQ = (it & 1) ? sQ1 : sQ0;
                mbar_wait(&q_full[it & 1], (it >> 1) & 1);
                TICK(0)
                if (it > 0) mbar_wait(s_free, (it - 1) & 1);
                tc_fence_after();
                TICK(1)
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    umma_f16_lohi(tmem_base + TM_S, kmaj(sQ) + koff(k), kmaj(sK) + koff(k), DESC_HI, id_kk, k != 0);
                mbar_wait(do_full, it & 1);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    umma_f16_lohi(tmem_base + TM_DP, kmaj(sdO) + koff(k), kmaj(sV) + koff(k), DESC_HI, id_kk, k != 0);
                umma_commit(sdp_full);
                TICK(2)
                mbar_wait(pds_ready, it & 1);
                tc_fence_after();
                TICK(3)
#pragma unroll
                for (int k = 0; k < 8; ++k)   // dV[kv, d] += sum_q P[q, kv] dO[q, d]
                    umma_f16_lohi(tmem_base + TM_DV, mnmaj(sP) + k * (2048u >> 4), mnmaj(sdO) + k * (2048u >> 4),
                                  DESC_HI, id_mm, (it | k) != 0);
                umma_commit(do_free);
#pragma unroll
                for (int k = 0; k < 8; ++k)   // dK[kv, d] += sum_q dS[q, kv] Q[q, d]
                    umma_f16_lohi(tmem_base + TM_DK, mnmaj(sdS) + k * (2048u >> 4), mnmaj(sQ) + k * (2048u >> 4),
                                  DESC_HI, id_mm, (it | k) != 0);
                umma_commit(&q_free[it & 1]);
#pragma unroll
                for (int k = 0; k < 8; ++k)   // dQ[q, d] = sum_kv dS[q, kv] K[kv, d]
                    umma_f16_lohi(tmem_base + TM_S, kmaj(sdS) + koff(k), mnmaj(sK) + k * (2048u >> 4), DESC_HI, id_km,
                                  k != 0);
                umma_commit(mma2_done);
                TICK(4)
            }
            if (tm) { for (int i = 0; i < 5; ++i) dbg[8 + i] = acc[i]; dbg[15] = nit; }
        }
    } else if (warp >= 4) {
        // eight warps: warp w works on TMEM lane quarter w % 4 (= 32 query rows) and on one half of the 128 columns,
        // so every row of S / dP / dQ / dV / dK is split between two threads
        const int q = warp & 3;
        const int c_lo = ((warp - 4) >> 2) * 2;
        const int row = q * 32 + lane;
        const uint32_t lane_field = static_cast<uint32_t>(q * 32) << 16;
        const int sw = row & 7;
        const long bh = blockIdx.y;
        t0 = clock64();
        for (int it = 0; it < nit; ++it) {
            const int q0 = (i0 + it) * BQ;
            const bool diag = causal && it == 0;
            const float lse2 = lse[bh * S + q0 + row] * 1.4426950408889634f;
            const float dl = delta[bh * S + q0 + row];
            if (it > 0) {   // the previous block's dQ reduction has finished reading the P / dS buffers
                if (warp == 4 && lane == 0) tma_store_wait_read<0>();
                named_bar_sync(2, 256);
            }
            mbar_wait(sdp_full, it & 1);
            tc_fence_after();
            TICK(0)
#pragma unroll 1
            for (int c = c_lo; c < c_lo + 2; ++c) {
                uint32_t rs[32], rp[32];
                tmem_ld_32x32b_x32(tmem_base + lane_field + TM_S + c * 32, rs);
                tmem_ld_32x32b_x32(tmem_base + lane_field + TM_DP + c * 32, rp);
                tmem_ld_wait();
                uint32_t pk[16], dk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float p0 = ex2(__uint_as_float(rs[i]) * scale_log2 - lse2);
                    float p1 = ex2(__uint_as_float(rs[i + 1]) * scale_log2 - lse2);
                    if (diag && c * 32 + i > row) p0 = 0.f;
                    if (diag && c * 32 + i + 1 > row) p1 = 0.f;
                    const float d0 = p0 * (__uint_as_float(rp[i]) - dl) * scale;
                    const float d1 = p1 * (__uint_as_float(rp[i + 1]) - dl) * scale;
                    pk[i >> 1] = Cvt<T>::pack2(p0, p1);
                    dk[i >> 1] = Cvt<T>::pack2(d0, d1);
                }
                const int off = (c >> 1) * CHUNK_BYTES + row * 128;
                const int u0 = (c & 1) * 4;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int o = off + (((u0 + u) ^ sw) << 4);
                    *reinterpret_cast<uint4*>(sP + o) = make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
                    *reinterpret_cast<uint4*>(sdS + o) = make_uint4(dk[u * 4], dk[u * 4 + 1], dk[u * 4 + 2], dk[u * 4 + 3]);
                }
            }
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(pds_ready);
            TICK(1)
            // dQ_i of this key block -> fp32 accumulator
            mbar_wait(mma2_done, it & 1);
            tc_fence_after();
            TICK(2)
            // dQ_i: TMEM -> fp32 tile in the P / dS buffers (free now) -> ONE bulk tensor reduction per 32 columns into
            // the global accumulator. (Per-thread red.global.add.v4.f32 took 3800 of the 9700 cycles of a query block.)
#pragma unroll 1
            for (int c = c_lo; c < c_lo + 2; ++c) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(tmem_base + lane_field + TM_S + c * 32, r);
                tmem_ld_wait();
                uint8_t* dst = (c < 2 ? sP : sdS) + (c & 1) * CHUNK_BYTES + row * 128;
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    *reinterpret_cast<uint4*>(dst + ((u ^ sw) << 4)) = make_uint4(r[u * 4], r[u * 4 + 1], r[u * 4 + 2], r[u * 4 + 3]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_free);
            fence_proxy_async_smem();
            named_bar_sync(2, 256);
            if (warp == 4 && lane == 0) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    tma_reduce_add_4d(&tmDQ, (c < 2 ? sP : sdS) + (c & 1) * CHUNK_BYTES, c * 32, h, q0, b);
                tma_store_commit();
            }
            TICK(3)
        }
        if (warp == 4 && lane == 0) tma_store_wait_read<0>();
        named_bar_sync(2, 256);
        if (tm && warp == 4 && lane == 0) for (int i = 0; i < 4; ++i) dbg[i] = acc[i];
        // dV, dK (TMEM lane = key row) -> shared (P / dS buffers are free) -> TMA store
#pragma unroll 1
        for (int which = 0; which < 2; ++which) {
            uint8_t* stage = which == 0 ? sP : sdS;
            const uint32_t t_acc = tmem_base + lane_field + (which == 0 ? TM_DV : TM_DK);
#pragma unroll 1
            for (int c = c_lo; c < c_lo + 2; ++c) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(t_acc + c * 32, r);
                tmem_ld_wait();
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) pk[i >> 1] = Cvt<T>::pack2(__uint_as_float(r[i]), __uint_as_float(r[i + 1]));
                uint8_t* dst = stage + (c >> 1) * CHUNK_BYTES + row * 128;
                const int u0 = (c & 1) * 4;
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    *reinterpret_cast<uint4*>(dst + (((u0 + u) ^ sw) << 4)) =
                        make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
            }
        }
        fence_proxy_async_smem();
        named_bar_sync(1, 256);
        if (warp == 4 && lane == 0) {
            tma_store_4d(&tmdV, sP, 0, h, jb * BKV, b);
            tma_store_4d(&tmdV, sP + CHUNK_BYTES, 64, h, jb * BKV, b);
            tma_store_4d(&tmdK, sdS, 0, h, jb * BKV, b);
            tma_store_4d(&tmdK, sdS + CHUNK_BYTES, 64, h, jb * BKV, b);
            tma_store_commit();
            tma_store_wait<0>();
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
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
// [B, S, H, D] tensor with element strides (sb, ss, sh, 1) as a 4-D map ordered (D, H, S, B); box = 64 x 1 x 128 x 1.
static int make_map_bshd(CUtensorMap* m, int is_bf16, const void* base, int B, int S, int H, int D, long sb, long ss,
                         long sh, int box_rows = 128) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return -1;
    cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)H, (cuuint64_t)S, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)sh * 2, (cuuint64_t)ss * 2, (cuuint64_t)sb * 2};
    cuuint32_t box[4] = {64, 1, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                     const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -(int)r - 1000;
}

}  // namespace fa
}  // namespace dsb

// Forward variant: 2 = two query tiles per CTA with P kept in tensor memory (flash_attn_v2.cu, default);
// 64 = 64-key blocks, two CTAs per SM; 128 = 128-key blocks, one CTA per SM (first generation, kept for comparison).
static int g_fa_fwd_variant = 2;
// Backward variant: 2 = transposed formulation with P^T / dS^T in tensor memory (default); 1 = first generation.
static int g_fa_bwd_variant = 2;
extern "C" void dsb_flash_attn_debug_mode(int m) { dsb::fa2::set_debug_mode(m); }
extern "C" int dsb_flash_attn_bwd_variant(int v) {
    const int prev = g_fa_bwd_variant;
    if (v == 1 || v == 2) g_fa_bwd_variant = v;
    return prev;
}
static long long* g_fa_dbg = nullptr;
extern "C" void dsb_flash_attn_debug_timers(long long* p) {
    g_fa_dbg = p;
    dsb::fa2::set_debug_timers(p);
}
extern "C" int dsb_flash_attn_fwd_variant(int v) {
    const int prev = g_fa_fwd_variant;
    if (v == 2 || v == 64 || v == 128) g_fa_fwd_variant = v;
    return prev;
}

// out[b, s, h, :] = softmax(scale * q[b, s, h, :] . k[b, :, h, :]^T (+ causal mask)) v[b, :, h, :]; lse (optional,
// [B, H, S] fp32) receives the natural-log sum of exponentials of the scaled scores of every row.
// q / k / v / out: [B, S, H, 128] with element strides (s*b, s*s, s*h, 1); S a multiple of 128.
// dropout probability -> the kernels' 16-bit threshold form (fa2::Dropout); p is quantised to 1 / 65536
static dsb::fa2::Dropout make_dropout(float p, unsigned seed) {
    dsb::fa2::Dropout d = {0u, 1.f, seed};
    if (p > 0.f) {
        unsigned thr = (unsigned)(p * 65536.f + 0.5f);
        if (thr > 65535u) thr = 65535u;
        d.thr16 = thr;
        d.inv_keep = 65536.f / (float)(65536u - thr);
    }
    return d;
}

static int flash_fwd_impl(const void* q, const void* k, const void* v, void* out, float* lse, int B, int S, int H, int D,
                          long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh, long v_sb, long v_ss,
                          long v_sh, long o_sb, long o_ss, long o_sh, float scale, int causal, int dtype,
                          const int* sched, int sched_hstride, int sched_pstride, const float* kpm, void* stream_,
                          float dropout_p = 0.f, unsigned dropout_seed = 0) {
    using namespace dsb;
    using namespace dsb::fa;
    ensure_context_on_this_thread();
    if (B <= 0 || S <= 0 || H <= 0) return 0;
    if (D < 8 || D > HD || (D % 8) || (S % BQ)) return DSB_ERR_SIZE;      // D < 128: see flash_v2.cuh
    if (dtype != DSB_DTYPE_F16 && dtype != DSB_DTYPE_BF16) return DSB_ERR_DTYPE;
    if (dropout_p < 0.f || dropout_p >= 1.f) return DSB_ERR_ARG;
    const long strides[12] = {q_sb, q_ss, q_sh, k_sb, k_ss, k_sh, v_sb, v_ss, v_sh, o_sb, o_ss, o_sh};
    for (int i = 0; i < 12; ++i)
        if (strides[i] % 8) return DSB_ERR_ALIGN;
    if ((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v) |
         reinterpret_cast<uintptr_t>(out)) & 15)
        return DSB_ERR_ALIGN;
    const int bf = dtype == DSB_DTYPE_BF16;
    CUtensorMap tq, tk, tv, to;
    int rc = make_map_bshd(&tq, bf, q, B, S, H, D, q_sb, q_ss, q_sh);
    if (rc) return rc;
    rc = make_map_bshd(&tk, bf, k, B, S, H, D, k_sb, k_ss, k_sh);
    if (rc) return rc;
    rc = make_map_bshd(&tv, bf, v, B, S, H, D, v_sb, v_ss, v_sh);
    if (rc) return rc;
    rc = make_map_bshd(&to, bf, out, B, S, H, D, o_sb, o_ss, o_sh);
    if (rc) return rc;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const float scale_log2 = scale * 1.4426950408889634f;
    dim3 grid(S / BQ, B * H);
    cudaError_t e;
    const bool extras = sched != nullptr || kpm != nullptr || dropout_p > 0.f || D != HD;
    if (sched != nullptr && S % 256) return DSB_ERR_SIZE;   // schedules are per 256-row query block
    if (g_fa_fwd_variant == 2 || extras) {
        rc = make_map_bshd(&tk, bf, k, B, S, H, D, k_sb, k_ss, k_sh, K64);
        if (rc) return rc;
        rc = make_map_bshd(&tv, bf, v, B, S, H, D, v_sb, v_ss, v_sh, K64);
        if (rc) return rc;
        return dsb::fa2::launch_fwd(tq, tk, tv, to, lse, B, S, H, scale_log2, causal, bf, sched, sched_hstride,
                                    sched_pstride, kpm, make_dropout(dropout_p, dropout_seed), stream, D);
    }
    if (g_fa_fwd_variant != 128) {
        rc = make_map_bshd(&tk, bf, k, B, S, H, D, k_sb, k_ss, k_sh, K64);
        if (rc) return rc;
        rc = make_map_bshd(&tv, bf, v, B, S, H, D, v_sb, v_ss, v_sh, K64);
        if (rc) return rc;
        if (bf) {
            auto kern = flash_fwd64_kernel<__nv_bfloat16>;
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_BYTES);
            if (e != cudaSuccess) return -(int)e;
            kern<<<grid, 256, F_SMEM_BYTES, stream>>>(tq, tk, tv, to, lse, S, H, scale_log2, causal, g_fa_dbg);
        } else {
            auto kern = flash_fwd64_kernel<__half>;
            e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM_BYTES);
            if (e != cudaSuccess) return -(int)e;
            kern<<<grid, 256, F_SMEM_BYTES, stream>>>(tq, tk, tv, to, lse, S, H, scale_log2, causal, g_fa_dbg);
        }
    } else if (bf) {
        auto kern = flash_fwd_kernel<__nv_bfloat16>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 384, SMEM_BYTES, stream>>>(tq, tk, tv, to, lse, S, H, scale_log2, causal);
    } else {
        auto kern = flash_fwd_kernel<__half>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 384, SMEM_BYTES, stream>>>(tq, tk, tv, to, lse, S, H, scale_log2, causal);
    }
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

extern "C" int dsb_flash_attn_fwd(const void* q, const void* k, const void* v, void* out, float* lse, int B, int S,
                                  int H, int D, long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh,
                                  long v_sb, long v_ss, long v_sh, long o_sb, long o_ss, long o_sh, float scale,
                                  int causal, int dtype, void* stream_) {
    return flash_fwd_impl(q, k, v, out, lse, B, S, H, D, q_sb, q_ss, q_sh, k_sb, k_ss, k_sh, v_sb, v_ss, v_sh, o_sb, o_ss,
                          o_sh, scale, causal, dtype, nullptr, 0, 0, nullptr, stream_);
}

// Same with a block-sparse schedule and / or an additive key mask (S % 256 == 0):
//   sched : device int32, per (head, 256-row query block) a record [n, (64-key block index, visibility word of the first
//           128 rows, of the second 128 rows) * n] at sched + head * sched_hstride + block * sched_pstride
//           (sched_hstride = 0: one layout for all heads). Bit qgroup * 4 + kgroup of a word = 16-query group qgroup of the
//           tile sees 16-key group kgroup. `causal` is ignored when a schedule is given (the layout says what is visible).
//   kpm   : fp32 [B, S], added to the scaled scores of every query; ALREADY multiplied by log2(e).
// Reference: deepspeed/ops/sparse_attention (SDD matmul + block-sparse softmax + DSD matmul as three Triton kernels).
extern "C" int dsb_flash_attn_fwd_ex(const void* q, const void* k, const void* v, void* out, float* lse, int B, int S,
                                     int H, int D, const long* strides /* 4 x (sb, ss, sh): q k v out */, float scale,
                                     int causal, int dtype, const int* sched, int sched_hstride, int sched_pstride,
                                     const float* kpm, void* stream_) {
    return flash_fwd_impl(q, k, v, out, lse, B, S, H, D, strides[0], strides[1], strides[2], strides[3], strides[4],
                          strides[5], strides[6], strides[7], strides[8], strides[9], strides[10], strides[11], scale,
                          causal, dtype, sched, sched_hstride, sched_pstride, kpm, stream_);
}

// dsb_flash_attn_fwd_ex with attention dropout inside the kernel: every probability is dropped with probability
// `dropout_p` (quantised to 1 / 65536) and the kept ones are scaled by 1 / (1 - p), as torch.nn.functional.dropout does
// on the softmax output. The keep bits are a hash of (seed, batch, head, query, key): pass the same `dropout_seed` to
// dsb_flash_attn_bwd2_drop. Reference: the attention-probability dropout of csrc/transformer/dropout_kernels.cu as used
// by ds_transformer_cuda.cpp:225-236.
extern "C" int dsb_flash_attn_fwd_drop(const void* q, const void* k, const void* v, void* out, float* lse, int B, int S,
                                       int H, int D, const long* strides, float scale, int causal, int dtype,
                                       const int* sched, int sched_hstride, int sched_pstride, const float* kpm,
                                       float dropout_p, unsigned dropout_seed, void* stream_) {
    return flash_fwd_impl(q, k, v, out, lse, B, S, H, D, strides[0], strides[1], strides[2], strides[3], strides[4],
                          strides[5], strides[6], strides[7], strides[8], strides[9], strides[10], strides[11], scale,
                          causal, dtype, sched, sched_hstride, sched_pstride, kpm, stream_, dropout_p, dropout_seed);
}

// Backward of dsb_flash_attn_fwd. dO / q / k / v / out / dk / dv: [B, S, H, 128] with their own element strides (unit
// last stride, multiples of 8). lse: [B, H, S] from the forward. delta: [B, H, S] fp32 scratch. dq_acc: [B, S, H, 128]
// fp32, contiguous, ZEROED by the caller; it receives dQ (key blocks add their contributions with reductions).
extern "C" int dsb_flash_attn_bwd(const void* dO, const void* q, const void* k, const void* v, const void* out,
                                  const float* lse, float* delta, float* dq_acc, void* dk, void* dv, int B, int S,
                                  int H, int D, const long* strides /* 7 x (sb, ss, sh): dO q k v out dk dv */,
                                  float scale, int causal, int dtype, void* stream_) {
    using namespace dsb;
    using namespace dsb::fa;
    ensure_context_on_this_thread();
    if (B <= 0 || S <= 0 || H <= 0) return 0;
    if (D != HD || (S % BQ)) return DSB_ERR_SIZE;
    if (dtype != DSB_DTYPE_F16 && dtype != DSB_DTYPE_BF16) return DSB_ERR_DTYPE;
    for (int i = 0; i < 21; ++i)
        if (strides[i] % 8) return DSB_ERR_ALIGN;
    const int bf = dtype == DSB_DTYPE_BF16;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const long* s_do = strides, *s_q = strides + 3, *s_k = strides + 6, *s_v = strides + 9, *s_o = strides + 12,
               *s_dk = strides + 15, *s_dv = strides + 18;
    const long rows = (long)B * S * H;
    int grid_d = (int)((rows + 7) / 8 < 148L * 16 ? (rows + 7) / 8 : 148L * 16);
    if (bf)
        flash_delta_kernel<__nv_bfloat16><<<grid_d, 256, 0, stream>>>((const __nv_bfloat16*)dO, (const __nv_bfloat16*)out, delta, B, S, H, s_do[0], s_do[1], s_do[2], s_o[0], s_o[1], s_o[2]);
    else
        flash_delta_kernel<__half><<<grid_d, 256, 0, stream>>>((const __half*)dO, (const __half*)out, delta, B, S, H, s_do[0], s_do[1], s_do[2], s_o[0], s_o[1], s_o[2]);
    CUtensorMap tq, tk, tv, tdo, tdk, tdv;
    int rc = make_map_bshd(&tq, bf, q, B, S, H, D, s_q[0], s_q[1], s_q[2]);
    if (rc) return rc;
    rc = make_map_bshd(&tk, bf, k, B, S, H, D, s_k[0], s_k[1], s_k[2]);
    if (rc) return rc;
    rc = make_map_bshd(&tv, bf, v, B, S, H, D, s_v[0], s_v[1], s_v[2]);
    if (rc) return rc;
    rc = make_map_bshd(&tdo, bf, dO, B, S, H, D, s_do[0], s_do[1], s_do[2]);
    if (rc) return rc;
    rc = make_map_bshd(&tdk, bf, dk, B, S, H, D, s_dk[0], s_dk[1], s_dk[2]);
    if (rc) return rc;
    rc = make_map_bshd(&tdv, bf, dv, B, S, H, D, s_dv[0], s_dv[1], s_dv[2]);
    if (rc) return rc;
    CUtensorMap tdq;
    {
        EncodeTiledFn enc = get_encode();
        if (!enc) return -1;
        cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)H, (cuuint64_t)S, (cuuint64_t)B};
        cuuint64_t st[3] = {(cuuint64_t)D * 4, (cuuint64_t)H * D * 4, (cuuint64_t)S * H * D * 4};
        cuuint32_t box[4] = {32, 1, 128, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        CUresult r = enc(&tdq, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dq_acc, dims, st, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return -(int)r - 1000;
    }
    const float scale_log2 = scale * 1.4426950408889634f;
    dim3 grid(S / BKV, B * H);
    cudaError_t e;
    if (bf) {
        auto kern = flash_bwd_kernel<__nv_bfloat16>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, B_SMEM_BYTES);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 384, B_SMEM_BYTES, stream>>>(tq, tk, tv, tdo, tdk, tdv, tdq, lse, delta, dq_acc, S, H, scale, scale_log2, causal, g_fa_dbg);
    } else {
        auto kern = flash_bwd_kernel<__half>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, B_SMEM_BYTES);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 384, B_SMEM_BYTES, stream>>>(tq, tk, tv, tdo, tdk, tdv, tdq, lse, delta, dq_acc, S, H, scale, scale_log2, causal, g_fa_dbg);
    }
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

// Second-generation backward (flash_attn_v2.cu). Same tensors as dsb_flash_attn_bwd plus the query gradient itself:
// dq / dk / dv are written in the input dtype. `strides`: 8 x (sb, ss, sh) for dO, q, k, v, out, dq, dk, dv.
// `scratch`: fp32, (2 + 128) * B * H * S elements: -lse * log2 e, -scale * rowsum(dO * O) and the TRANSPOSED fp32 dQ
// accumulator [B, H, 128, S]; it is cleared here (stream-ordered memset), nothing is allocated.
// rot_dim > 0 (multiple of 16, tables [S, rot_dim / 2] fp32 from dsb_rotary_table): q and k were rotated before the
// forward pass, and dq / dk come back as gradients of the UNROTATED tensors (the inverse rotation is applied in the dK
// epilogue and in the dQ finishing pass, so no separate rotary launch is needed in backward).
static int flash_bwd2_impl(const void* dO, const void* q, const void* k, const void* v, const void* out,
                           const float* lse, float* scratch, void* dq, void* dk, void* dv, int B, int S, int H, int D,
                           const long* strides, float scale, int causal, int dtype, const float* rot_cos,
                           const float* rot_sin, int rot_dim, const int* sched, int sched_hstride, int sched_pstride,
                           const float* kpm, void* stream_, float dropout_p = 0.f, unsigned dropout_seed = 0) {
    using namespace dsb;
    using namespace dsb::fa;
    ensure_context_on_this_thread();
    if (B <= 0 || S <= 0 || H <= 0) return 0;
    if (D < 8 || D > HD || (D % 8) || (S % BQ)) return DSB_ERR_SIZE;      // D < 128: see flash_v2.cuh
    if (dtype != DSB_DTYPE_F16 && dtype != DSB_DTYPE_BF16) return DSB_ERR_DTYPE;
    if (dropout_p < 0.f || dropout_p >= 1.f) return DSB_ERR_ARG;
    if (rot_dim > 0 && (rot_cos == nullptr || rot_sin == nullptr)) return DSB_ERR_ARG;
    for (int i = 0; i < 24; ++i)
        if (strides[i] % 8) return DSB_ERR_ALIGN;
    const int bf = dtype == DSB_DTYPE_BF16;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const long *s_do = strides, *s_q = strides + 3, *s_k = strides + 6, *s_v = strides + 9, *s_o = strides + 12,
               *s_dq = strides + 15, *s_dk = strides + 18, *s_dv = strides + 21;
    const long bhs = (long)B * H * S;
    float* nlse2 = scratch;
    float* ndelta = scratch + bhs;
    float* dq_acc = scratch + 2 * bhs;
    cudaError_t ce = cudaMemsetAsync(dq_acc, 0, sizeof(float) * bhs * HD, stream);
    if (ce != cudaSuccess) return -(int)ce;
    int rc = fa2::launch_prep(dO, out, lse, nlse2, ndelta, B, S, H, s_do, s_o, scale, bf, stream, D);
    if (rc) return rc;
    CUtensorMap tq, tk, tv, tdo, tdk, tdv, tdq;
    if ((rc = make_map_bshd(&tq, bf, q, B, S, H, D, s_q[0], s_q[1], s_q[2], 64))) return rc;
    if ((rc = make_map_bshd(&tdo, bf, dO, B, S, H, D, s_do[0], s_do[1], s_do[2], 64))) return rc;
    if ((rc = make_map_bshd(&tk, bf, k, B, S, H, D, s_k[0], s_k[1], s_k[2]))) return rc;
    if ((rc = make_map_bshd(&tv, bf, v, B, S, H, D, s_v[0], s_v[1], s_v[2]))) return rc;
    if ((rc = make_map_bshd(&tdk, bf, dk, B, S, H, D, s_dk[0], s_dk[1], s_dk[2]))) return rc;
    if ((rc = make_map_bshd(&tdv, bf, dv, B, S, H, D, s_dv[0], s_dv[1], s_dv[2]))) return rc;
    EncodeTiledFn enc = get_encode();
    if (!enc) return -1;
    // the accumulator always has 128 rows per head (see the strides), whatever the head dim; the map exposes only the
    // first D of them, so the reduce-adds of the rows beyond (all zeros) are clipped instead of travelling to L2
    cuuint64_t dims[4] = {(cuuint64_t)S, (cuuint64_t)D, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t st[3] = {(cuuint64_t)S * 4, (cuuint64_t)HD * S * 4, (cuuint64_t)H * HD * S * 4};
    cuuint32_t box[4] = {32, 128, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tdq, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dq_acc, dims, st, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return -(int)r - 1000;
    rc = fa2::launch_bwd(tq, tk, tv, tdo, tdk, tdv, tdq, nlse2, ndelta, B, S, H, scale, scale * 1.4426950408889634f, causal,
                         bf, rot_cos, rot_sin, rot_dim, sched, sched_hstride, sched_pstride, kpm,
                         make_dropout(dropout_p, dropout_seed), stream, D);
    if (rc) return rc;
    return fa2::launch_dq_finish(dq_acc, dq, B, S, H, s_dq, bf, rot_cos, rot_sin, rot_dim, stream, D);
}

extern "C" int dsb_flash_attn_bwd2(const void* dO, const void* q, const void* k, const void* v, const void* out,
                                   const float* lse, float* scratch, void* dq, void* dk, void* dv, int B, int S, int H,
                                   int D, const long* strides, float scale, int causal, int dtype,
                                   const float* rot_cos, const float* rot_sin, int rot_dim, void* stream_) {
    return flash_bwd2_impl(dO, q, k, v, out, lse, scratch, dq, dk, dv, B, S, H, D, strides, scale, causal, dtype, rot_cos,
                           rot_sin, rot_dim, nullptr, 0, 0, nullptr, stream_);
}

// Backward of dsb_flash_attn_fwd_ex. `sched`: per (head, 128-key block) a record [n, (64-row query block index,
// visibility word) * n]; bit kgroup * 4 + qgroup of a word = 16-key group kgroup of the block is seen by 16-query group
// qgroup of the tile. `kpm` as in the forward.
extern "C" int dsb_flash_attn_bwd2_ex(const void* dO, const void* q, const void* k, const void* v, const void* out,
                                      const float* lse, float* scratch, void* dq, void* dk, void* dv, int B, int S,
                                      int H, int D, const long* strides, float scale, int causal, int dtype,
                                      const int* sched, int sched_hstride, int sched_pstride, const float* kpm,
                                      void* stream_) {
    return flash_bwd2_impl(dO, q, k, v, out, lse, scratch, dq, dk, dv, B, S, H, D, strides, scale, causal, dtype, nullptr,
                           nullptr, 0, sched, sched_hstride, sched_pstride, kpm, stream_);
}

// Backward of dsb_flash_attn_fwd_drop (same dropout_p / dropout_seed).
extern "C" int dsb_flash_attn_bwd2_drop(const void* dO, const void* q, const void* k, const void* v, const void* out,
                                        const float* lse, float* scratch, void* dq, void* dk, void* dv, int B, int S,
                                        int H, int D, const long* strides, float scale, int causal, int dtype,
                                        const int* sched, int sched_hstride, int sched_pstride, const float* kpm,
                                        float dropout_p, unsigned dropout_seed, void* stream_) {
    return flash_bwd2_impl(dO, q, k, v, out, lse, scratch, dq, dk, dv, B, S, H, D, strides, scale, causal, dtype, nullptr,
                           nullptr, 0, sched, sched_hstride, sched_pstride, kpm, stream_, dropout_p, dropout_seed);
}
