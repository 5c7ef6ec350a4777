// Second-generation flash attention for sm_100a (head dim 128, fp16 / bf16): the probabilities never leave the tensor
// core's own memory.
//
// Forward (flash_fwd2_kernel): one CTA per 256 query rows of one (batch, head) = two 128-row tiles that share every
//   K / V tile (one TMA load feeds both), 64 keys per step.
//     warp 0 / warp 3 : TMA producers (Q tiles + K ring / V ring, 4 stages each); warp 3 also owns the TMEM allocation
//     warps 1, 2      : UMMA issue for tile 0 / tile 1. S_t(j) = Q_t K_j^T goes to one of two TMEM score buffers per tile;
//                       O_t += P_t(j) V_j takes its A operand (P, 16 bit) straight from TENSOR MEMORY - the softmax
//                       warps overwrite the scores they have just read with the packed probabilities, so there is no
//                       shared-memory round trip, no proxy fence and no bank conflict on that path.
//     warps 4-7 / 8-11: online softmax of tile 0 / tile 1, one query row per thread (TMEM lane). The running maximum
//                       is only raised when it grows by more than 2^8 (the O accumulator is then rescaled by the same
//                       warps); otherwise a block costs one TMEM load, 32 three-input maxima, 32 packed FFMA2, 64
//                       exponentials, 32 packed adds, 32 converts and one TMEM store, and never waits for the P V
//                       product of the previous block.
//   Buffer reuse needs no barrier: the tensor pipe executes the UMMAs of the issuing thread in order, so S_t(j+2) may
//   overwrite the columns P_t(j) was read from as soon as it has been issued after O_t += P_t(j) V_j.
//
// Backward (flash_bwd2_kernel): one CTA per (128-key block, batch, head), K / V resident, walks 64-row query blocks.
//   Everything is computed TRANSPOSED (lane = key row): S^T = K Q^T and dP^T = V dO^T land in TMEM, the softmax warps
//   write P^T and dS^T (16 bit) back over the scores, and dV += P^T dO, dK += dS^T Q read that A operand from TMEM.
//   dQ^T = K^T dS^T needs dS^T as the B operand, so it is also staged in shared memory (one 128-byte row per thread,
//   conflict-free). TMEM: 2 x S^T (64) + dP^T (64) + dQ^T (64) + dV (128) + dK (128) = 512 columns, i.e. dQ has its own
//   region and S^T is double-buffered, which removes the serial chain of the first-generation kernel.
//     warp 0: TMA producer (K, V once; Q ring of 3, dO ring of 2, -lse*log2e / -scale*delta rows with the Q tile)
//     warp 1: issues S^T / dP^T    warp 2: issues dV / dK / dQ^T    warp 3: TMEM allocation
//     warps 4-11 : P^T / dS^T (lane = key row, each warp half of the 64 query columns)
//     warps 12-15: dQ^T drain: TMEM -> fp32 tiles in shared memory -> TMA reduce-add into a TRANSPOSED accumulator
//                  ([B, H, D, S]; flash_dq_finish_kernel transposes and casts it afterwards)
// Arguments of both directions beyond plain causal / full attention:
//   * `sched`: block-sparse schedules (lists of live key tiles / query blocks with 16 x 16 visibility words);
//   * `kpm`: additive key mask in log2 units; `drop`: attention dropout from a hash of (seed, b, h, query, key), the
//     same bits in forward and backward, no mask tensor;
//   * `D`: head dims below 128 - the tensor maps carry the true D (TMA zero-fills / clips the rest) and the products
//     skip the zero columns (D / 16 k-steps, D-wide accumulators, dQ reduce-adds clipped to D rows).
// Reference for the math: csrc/transformer/ds_transformer_cuda.cpp:186-214 and :340-420 (score, softmax, context and
// their gradients as separate strided-batch GEMMs and softmax kernels).
#include "../common/ptx.cuh"
#include "../../include/dsb200.h"
#include "flash_v2.cuh"
#include <type_traits>

namespace dsb {
namespace fa2 {

// ------------------------------------------------------------------------------------------------ device helpers
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
DSB_DEVICE void tma_reduce_add_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                     reinterpret_cast<uint64_t>(m)),
                 "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
// plain (non-tensor) bulk copy global -> shared, completion on an mbarrier; 16-byte aligned, size % 16 == 0
DSB_DEVICE void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]
DSB_DEVICE void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t hi, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %4, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(hi), "r"(idesc), "r"(acc)
        : "memory");
}
// Warp-uniform issue: every lane of the issuing warp runs the role loop (waits, descriptor arithmetic on the uniform
// datapath), only the tcgen05 instruction itself is predicated on the elected lane. With the whole loop inside an
// `if (elected)` branch the descriptors live in per-thread registers and every UMMA pays register -> uniform-register
// moves: measured 38 cycles per issue, which made the issuing thread - not the tensor pipe - the limit of both kernels.
DSB_DEVICE void umma_ss_p(uint32_t lead, uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                          uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "setp.ne.b32 q, %6, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(acc), "r"(lead)
        : "memory");
}
DSB_DEVICE void umma_ts_p(uint32_t lead, uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                          uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .b64 db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "setp.ne.b32 q, %6, 0;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %4, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(hi), "r"(idesc), "r"(acc), "r"(lead)
        : "memory");
}
DSB_DEVICE void umma_commit_p(uint32_t lead, uint64_t* bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "setp.ne.b32 q, %1, 0;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
        ::"r"(smem_u32(bar)), "r"(lead)
        : "memory");
}
DSB_DEVICE void mbar_arrive_p(uint32_t lead, uint64_t* bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "setp.ne.b32 q, %1, 0;\n\t"
        "@q mbarrier.arrive.shared::cta.b64 _, [%0];\n\t}"
        ::"r"(smem_u32(bar)), "r"(lead)
        : "memory");
}
DSB_DEVICE void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
DSB_DEVICE void tmem_st_x32(uint32_t taddr, const uint32_t* r) {
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
DSB_DEVICE void tmem_st_x16(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
DSB_DEVICE void tmem_ld_x32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
DSB_DEVICE float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
DSB_DEVICE float max3(float a, float b, float c) {
    float y;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(y) : "f"(a), "f"(b), "f"(c));
    return y;
}
// packed fp32 pairs (one 64-bit register): FFMA2 / FADD2 / FMUL2 halve the issue slots of the softmax arithmetic
DSB_DEVICE uint64_t pk2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
DSB_DEVICE uint64_t pk2u(uint32_t lo, uint32_t hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
    return r;
}
DSB_DEVICE void upk2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
DSB_DEVICE uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
DSB_DEVICE uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
DSB_DEVICE uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

DSB_DEVICE uint64_t add2_rm(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("add.rm.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
// 2^x for a packed pair on the FMA / ALU pipes (no MUFU): x = floor(x) + f, 2^f by a cubic fitted on [0, 1) (max relative
// error 8.7e-5, below the rounding of a 16-bit probability), 2^floor(x) by an add into the exponent field. x in
// [-125, 126]; callers clamp from below. The MUFU unit does 16 exponentials per clock and SM, which is exactly the
// tensor pipe's appetite at head dim 128; moving half of them here takes the special-function unit off the critical path.
DSB_DEVICE void exp2_poly2(uint64_t x, float& lo, float& hi) {
    const uint64_t magic = pk2(12582912.f, 12582912.f);          // 1.5 * 2^23: the low mantissa bits of x + magic are floor(x)
    const uint64_t t = add2_rm(x, magic);
    const uint64_t fl = add2(t, pk2(-12582912.f, -12582912.f));   // floor(x) as a float
    const uint64_t f = fma2(fl, pk2(-1.f, -1.f), x);             // x - floor(x) in [0, 1)
    uint64_t pz = fma2(pk2(0.07705804705619812f, 0.07705804705619812f), f, pk2(0.22764557600021362f, 0.22764557600021362f));
    pz = fma2(pz, f, pk2(0.6951228976249695f, 0.6951228976249695f));
    pz = fma2(pz, f, pk2(1.f, 1.f));
    float t0, t1, p0, p1;
    upk2(t, t0, t1);
    upk2(pz, p0, p1);
    lo = __uint_as_float(__float_as_uint(p0) + (__float_as_uint(t0) << 23));
    hi = __uint_as_float(__float_as_uint(p1) + (__float_as_uint(t1) << 23));
}

constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);   // SBO 1024 B, descriptor version, 128 B swizzle
DSB_DEVICE uint32_t desc_lo(const void* p, uint32_t lbo_bytes) {
    return ((smem_u32(p) & 0x3FFFF) >> 4) | ((lbo_bytes >> 4) << 16);
}
constexpr uint32_t NEG_INF_BITS = 0xff800000u;

// ---- dropout bits (see fa2::Dropout): one 32-bit hash serves the two keys of an even / odd pair
constexpr uint32_t DROP_C1 = 0x9E3779B1u, DROP_C2 = 0x85EBCA77u;
__device__ __forceinline__ uint32_t lowbias32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t drop_seed_bh(uint32_t seed, uint32_t bh) { return lowbias32(seed ^ (bh * 0x27d4eb2fu)); }

// ================================================================================================ forward
constexpr int F_BQ = 128;                  // rows per query tile
constexpr int F_BK = 64;                   // keys per step
constexpr int F_NS = 4;                    // K / V ring stages
constexpr int F_QTILE = 128 * 128 * 2;     // 32 KiB = two chunks of [128 rows x 128 B]
constexpr int F_QCHUNK = 128 * 128;        // 16 KiB
constexpr int F_KVTILE = 64 * 128 * 2;     // 16 KiB = two chunks of [64 rows x 128 B]
constexpr int F_KVCHUNK = 64 * 128;        // 8 KiB
constexpr int F_OFF_Q = 0;
constexpr int F_OFF_K = F_OFF_Q + 2 * F_QTILE;
constexpr int F_OFF_V = F_OFF_K + F_NS * F_KVTILE;
constexpr int F_OFF_BAR = F_OFF_V + F_NS * F_KVTILE;
constexpr int F_SMEM = F_OFF_BAR + 512 + 1024;
constexpr bool F_POLY = true;              // half of the exponentials as a polynomial on the FMA pipe
constexpr float F_TAU = 8.f;               // the running maximum may lag the true one by 2^8

template <typename T>
__global__ void __launch_bounds__(384, 1)
flash_fwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                  float* __restrict__ lse, int S, int H, float scale_log2, int causal, int dbg, long long* prof,
                  const int* __restrict__ sched, int sched_hstride, int sched_pstride,
                  const float* __restrict__ kpm, const Dropout drop, const int D) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem + F_OFF_Q;
    uint8_t* sK = smem + F_OFF_K;
    uint8_t* sV = smem + F_OFF_V;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + F_OFF_BAR);
    uint64_t* q_full = bars;                    // [2]
    uint64_t* k_full = bars + 2;                // [NS]
    uint64_t* k_empty = k_full + F_NS;          // [NS]
    uint64_t* v_full = k_empty + F_NS;          // [NS]
    uint64_t* v_empty = v_full + F_NS;          // [NS]
    uint64_t* s_full = v_empty + F_NS;          // [tile][buf]
    uint64_t* p_ready = s_full + 4;             // [tile][buf]
    uint64_t* pv_done = p_ready + 4;            // [tile] every P V product (parity = block index; rescale path only)
    uint64_t* o_final = pv_done + 2;            // [tile] the LAST P V product of the tile (single use)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_final + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int pr = gridDim.y - 1 - blockIdx.y;   // the long (late) row blocks are dispatched first
    const int q0 = pr * 2 * F_BQ;
    // Block-sparse mode: `sc` = [n, (key block, visibility word of tile 0, of tile 1) * n] for this (head, row pair). A
    // visibility word has one bit per (16-row query group, 16-key group) of a 128 x 64 tile: bit = qgroup * 4 + kgroup.
    const int* sc = sched != nullptr ? sched + (long)h * sched_hstride + (long)pr * sched_pstride : nullptr;
    const int nA = sc ? sc[0] : (causal ? (q0 + F_BQ) / F_BK : S / F_BK);
    const bool tileB = q0 + F_BQ < S;             // S % 256 == 128: the last CTA has only its first tile
    const int nB = !tileB ? 0 : (sc ? sc[0] : (causal ? (q0 + 2 * F_BQ) / F_BK : S / F_BK));
    const int nmax = nA > nB ? nA : nB;
    auto key_block = [&](int j) { return sc ? sc[1 + 3 * j] : j; };
    // phase timers of CTA (0, 0) (the longest one), only when a buffer is registered with dsb_flash_attn_debug_timers
    const bool tm = prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0;
    long long t0 = 0, acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define FA2_TICK(n) if (tm) { const long long t1 = clock64(); acc[n] += t1 - t0; t0 = t1; }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmO);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(&q_full[0], 1);
        mbar_init(&q_full[1], 1);
        for (int i = 0; i < F_NS; ++i) {
            mbar_init(&k_full[i], 1);
            mbar_init(&k_empty[i], 2);   // released by the issuing warps of both tiles
            mbar_init(&v_full[i], 1);
            mbar_init(&v_empty[i], 2);
        }
        for (int i = 0; i < 4; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&p_ready[i], 4);
        }
        mbar_init(&pv_done[0], 1);
        mbar_init(&pv_done[1], 1);
        mbar_init(&o_final[0], 1);
        mbar_init(&o_final[1], 1);
        fence_barrier_init();
    }
    if (warp == 3) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // TMEM columns: scores of tile t, buffer u at t * 128 + u * 64 (P overwrites the first 32 of them); O_t at 256 + t * 128

    if (warp == 0) {
        if (elect_one()) {
            for (int t = 0; t < (tileB ? 2 : 1); ++t) {
                mbar_arrive_expect_tx(&q_full[t], F_QTILE);
                tma_load_4d(sQ + t * F_QTILE, &tmQ, &q_full[t], 0, h, q0 + t * F_BQ, b);
                tma_load_4d(sQ + t * F_QTILE + F_QCHUNK, &tmQ, &q_full[t], 64, h, q0 + t * F_BQ, b);
            }
            for (int j = 0; j < nmax; ++j) {
                const int st = j % F_NS;
                if (j >= F_NS) mbar_wait(&k_empty[st], ((j / F_NS) - 1) & 1);
                if ((dbg & 32) && j >= F_NS) { mbar_arrive(&k_full[st]); continue; }
                mbar_arrive_expect_tx(&k_full[st], F_KVTILE);
                uint8_t* dst = sK + st * F_KVTILE;
                tma_load_4d(dst, &tmK, &k_full[st], 0, h, key_block(j) * F_BK, b);
                tma_load_4d(dst + F_KVCHUNK, &tmK, &k_full[st], 64, h, key_block(j) * F_BK, b);
            }
        }
    } else if (warp == 3) {
        if (elect_one()) {
            for (int j = 0; j < nmax; ++j) {
                const int st = j % F_NS;
                if (j >= F_NS) mbar_wait(&v_empty[st], ((j / F_NS) - 1) & 1);
                if ((dbg & 32) && j >= F_NS) { mbar_arrive(&v_full[st]); continue; }
                mbar_arrive_expect_tx(&v_full[st], F_KVTILE);
                uint8_t* dst = sV + st * F_KVTILE;
                tma_load_4d(dst, &tmV, &v_full[st], 0, h, key_block(j) * F_BK, b);
                tma_load_4d(dst + F_KVCHUNK, &tmV, &v_full[st], 64, h, key_block(j) * F_BK, b);
            }
        }
    } else if (warp == 1 || warp == 2) {
        // One issuing warp per query tile: S_t(j + 1) = Q_t K_{j+1}^T, then O_t += P_t(j) V_j. Both warps release every
        // K / V stage (the empty barriers count two arrivals); the shorter tile keeps arriving for the blocks it skips.
        const int t = warp - 1;
        const uint32_t lead = elect_one() ? 1u : 0u;
        const int n = t ? nB : nA;
        constexpr int is_bf16 = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
        const uint32_t idesc_s = make_idesc_f16(F_BQ, (dbg & 16) ? 128 : F_BK, is_bf16, 0, 0);
        // head dims below 128: TMA zero-filled the columns beyond D, so the products skip them - S = Q K^T runs D / 16
        // k-steps and O += P V is D columns wide (the accumulator columns beyond are never written nor stored)
        const int ksteps = (D + 15) >> 4;
        const uint32_t idesc_pv = make_idesc_f16(F_BQ, ksteps * 16, is_bf16, 0, 1);
        const uint32_t q_lo = desc_lo(sQ + t * F_QTILE, 16);
        const uint32_t t_sc = tmem_base + t * 128;
        const uint32_t t_acc = tmem_base + 256 + t * 128;
        auto issue_s = [&](int j) {
            const uint32_t k_lo = desc_lo(sK + (j % F_NS) * F_KVTILE, 16);
            const uint32_t d = t_sc + (j & 1) * 64;
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (k < ksteps)
                    umma_ss_p(lead, d, q_lo + (k >> 2) * (F_QCHUNK >> 4) + (k & 3) * 2,
                              k_lo + (k >> 2) * (F_KVCHUNK >> 4) + (k & 3) * 2, DESC_HI, idesc_s, k != 0);
            umma_commit_p(lead, &s_full[t * 2 + (j & 1)]);
            umma_commit_p(lead, &k_empty[j % F_NS]);
        };
        auto issue_pv = [&](int j) {
            // A = P_t(j) in TMEM (16 keys = 8 columns per k-step); B = V_j read MN-major exactly as TMA delivered it
            const uint32_t v_lo = desc_lo(sV + (j % F_NS) * F_KVTILE, F_KVCHUNK);
            const uint32_t a = t_sc + (j & 1) * 64;
            if (!(dbg & 4)) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_ts_p(lead, t_acc, a + k * 8, v_lo + k * (2048u >> 4), DESC_HI, idesc_pv, (j | k) != 0);
            }
            umma_commit_p(lead, &pv_done[t]);
            umma_commit_p(lead, &v_empty[j % F_NS]);
            if (j == n - 1) umma_commit_p(lead, &o_final[t]);
        };
        const bool tmw = tm && lead;
        if (n > 0) {
            mbar_wait(&q_full[t], 0);
            mbar_wait(&k_full[0], 0);
            tc_fence_after();
            issue_s(0);
        }
        t0 = clock64();
        const long long t_begin = t0;
        for (int j = 0; j < n; ++j) {
            if (j + 1 < n) {
                const int nx = j + 1;
                mbar_wait(&k_full[nx % F_NS], (nx / F_NS) & 1);
                tc_fence_after();
                FA2_TICK(0)
                issue_s(nx);
                FA2_TICK(1)
            }
            mbar_wait(&v_full[j % F_NS], (j / F_NS) & 1);
            FA2_TICK(2)
            mbar_wait(&p_ready[t * 2 + (j & 1)], (j >> 1) & 1);
            tc_fence_after();
            FA2_TICK(3)
            issue_pv(j);
            FA2_TICK(4)
        }
        if (tmw) {
            for (int i = 0; i < 5; ++i) prof[16 + t * 6 + i] = acc[i];
            prof[30 + t] = clock64() - t_begin;
            prof[28 + t] = n;
        }
        for (int j = n; j < nmax; ++j) {   // blocks only the other tile works on: keep the two-arrival stage protocol going
            mbar_wait(&k_full[j % F_NS], (j / F_NS) & 1);
            mbar_arrive_p(lead, &k_empty[j % F_NS]);
            mbar_wait(&v_full[j % F_NS], (j / F_NS) & 1);
            mbar_arrive_p(lead, &v_empty[j % F_NS]);
        }
    } else if (warp >= 4) {
        const int t = (warp - 4) >> 2;
        const int qd = warp & 3;
        const int row = qd * 32 + lane;
        const uint32_t lane_field = static_cast<uint32_t>(qd * 32) << 16;
        const int n = t ? nB : nA;
        const int qrow = q0 + t * F_BQ + row;
        const uint32_t t_o = tmem_base + lane_field + 256 + t * 128;
        // with an additive key mask the scores are scaled and biased first, and the rest works on them with scale 1
        const float sc_eff = kpm != nullptr ? 1.f : scale_log2;
        const uint64_t sc2 = pk2(sc_eff, sc_eff);
        const float* kpm_b = kpm != nullptr ? kpm + (long)b * S : nullptr;
        float m_run = -INFINITY, l_run = 0.f;
        // dropout: this row's hash base; the row sum l keeps ALL probabilities, only the P handed to P V is thinned,
        // and the 1 / keep factor is applied once, with 1 / l, in the epilogue
        const uint32_t drop_row = drop_seed_bh(drop.seed, blockIdx.x) + static_cast<uint32_t>(qrow) * DROP_C1;
        t0 = clock64();
        for (int j = 0; j < n; ++j) {
            const int buf = j & 1;
            mbar_wait(&s_full[t * 2 + buf], (j >> 1) & 1);
            tc_fence_after();
            FA2_TICK(0)
            const uint32_t t_s = tmem_base + lane_field + t * 128 + buf * 64;
            uint32_t r[64];
            if (!(dbg & 2)) {
                tmem_ld_x32(t_s, r);
                tmem_ld_x32(t_s + 32, r + 32);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 64; ++i) r[i] = 0x3f800000u + i + j;
            }
            FA2_TICK(1)
            if (dbg & 1) {
                tmem_st_x32(t_s, r);
                tmem_st_wait();
                FA2_TICK(4)
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_ready[t * 2 + buf]);
                FA2_TICK(5)
                l_run = 1.f;
                m_run = 0.f;
                continue;
            }
            if (kpm_b != nullptr) {
                const float4* km = reinterpret_cast<const float4*>(kpm_b + key_block(j) * F_BK);
#pragma unroll
                for (int i = 0; i < 64; i += 4) {
                    const float4 k4 = __ldg(km + (i >> 2));
                    float a0, a1, b0, b1;
                    upk2(fma2(pk2u(r[i], r[i + 1]), pk2(scale_log2, scale_log2), pk2(k4.x, k4.y)), a0, a1);
                    upk2(fma2(pk2u(r[i + 2], r[i + 3]), pk2(scale_log2, scale_log2), pk2(k4.z, k4.w)), b0, b1);
                    r[i] = __float_as_uint(a0);
                    r[i + 1] = __float_as_uint(a1);
                    r[i + 2] = __float_as_uint(b0);
                    r[i + 3] = __float_as_uint(b1);
                }
            }
            if (sc != nullptr) {
                const uint32_t vis = (static_cast<uint32_t>(sc[2 + 3 * j + t]) >> ((row >> 4) * 4)) & 15u;
                if (__any_sync(0xffffffffu, vis != 15u)) {
#pragma unroll
                    for (int i = 0; i < 64; ++i)
                        if (!((vis >> (i >> 4)) & 1u)) r[i] = NEG_INF_BITS;
                }
            } else if (causal) {
                const int lim = qrow - j * F_BK;   // columns <= lim of this block are visible to this row
                if (__any_sync(0xffffffffu, lim < F_BK - 1)) {
#pragma unroll
                    for (int i = 0; i < 64; ++i)
                        if (i > lim) r[i] = NEG_INF_BITS;
                }
            }
            float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
            for (int i = 0; i < 64; i += 8) {
                m0 = max3(m0, __uint_as_float(r[i]), __uint_as_float(r[i + 1]));
                m1 = max3(m1, __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
                m2 = max3(m2, __uint_as_float(r[i + 4]), __uint_as_float(r[i + 5]));
                m3 = max3(m3, __uint_as_float(r[i + 6]), __uint_as_float(r[i + 7]));
            }
            const float m_cand = max3(m0, m1, fmaxf(m2, m3)) * sc_eff;
            const bool grow = m_cand > m_run + F_TAU;   // always true for the first block (m_run = -inf)
            if (__any_sync(0xffffffffu, grow)) {
                const float m_new = grow ? m_cand : m_run;
                // rows that do not grow keep their state (also rows that have seen nothing yet: -inf - -inf is not 0)
                const float alpha = grow ? ex2(m_run - m_new) : 1.f;
                if (j > 0) {
                    mbar_wait(&pv_done[t], (j - 1) & 1);   // O_t is quiescent: P V of the previous block has completed
                    tc_fence_after();
#pragma unroll 1
                    for (int c = 0; c < 4; ++c) {
                        uint32_t o[32];
                        tmem_ld_x32(t_o + c * 32, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st_x32(t_o + c * 32, o);
                    }
                }
                l_run *= alpha;
                m_run = m_new;
            }
            FA2_TICK(2)
            const float m_use = m_run == -INFINITY ? 0.f : m_run;   // a row with nothing visible yet: exp2(-inf - 0) = 0
            const uint64_t nm2 = pk2(-m_use, -m_use);
            uint64_t sa = 0ull, sb = 0ull;
            uint32_t pk[32];
#pragma unroll
            for (int i = 0; i < 64; i += 4) {
                float a0, a1, b0, b1;
                upk2(fma2(pk2u(r[i], r[i + 1]), sc2, nm2), a0, a1);
                upk2(fma2(pk2u(r[i + 2], r[i + 3]), sc2, nm2), b0, b1);
                a0 = ex2(a0);
                a1 = ex2(a1);
                if (F_POLY && (i & 4)) {   // every fourth pair on the FMA pipe (balances MUFU time against issue slots)
                    exp2_poly2(pk2(fmaxf(b0, -125.f), fmaxf(b1, -125.f)), b0, b1);
                } else {
                    b0 = ex2(b0);
                    b1 = ex2(b1);
                }
                sa = add2(sa, pk2(a0, a1));
                sb = add2(sb, pk2(b0, b1));
                if (drop.thr16) {
                    const uint32_t kp = static_cast<uint32_t>(key_block(j) * F_BK + i) >> 1;      // key pair index
                    const uint32_t h0 = lowbias32(drop_row + kp * DROP_C2);
                    const uint32_t h1 = lowbias32(drop_row + (kp + 1) * DROP_C2);
                    if ((h0 & 0xffffu) < drop.thr16) a0 = 0.f;
                    if ((h0 >> 16) < drop.thr16) a1 = 0.f;
                    if ((h1 & 0xffffu) < drop.thr16) b0 = 0.f;
                    if ((h1 >> 16) < drop.thr16) b1 = 0.f;
                }
                pk[i >> 1] = Cvt<T>::pack2(a0, a1);
                pk[(i >> 1) + 1] = Cvt<T>::pack2(b0, b1);
            }
            FA2_TICK(3)
            tmem_st_x32(t_s, pk);
            float s0, s1;
            upk2(add2(sa, sb), s0, s1);
            l_run += s0 + s1;
            tmem_st_wait();
            FA2_TICK(4)
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_ready[t * 2 + buf]);
            FA2_TICK(5)
        }
        if (tm && qd == 0 && lane == 0)
            for (int i = 0; i < 6; ++i) prof[t * 8 + i] = acc[i];
        // epilogue: O / l -> shared (the tile's Q buffer, same swizzled layout) -> TMA store; log-sum-exp for backward
        // (a parity wait on pv_done would be ambiguous here: this thread may be two products behind)
        if (n > 0) {
        mbar_wait(&o_final[t], 0);
        tc_fence_after();
        const float inv_l = l_run > 0.f ? drop.inv_keep / l_run : 0.f;
        uint8_t* o_row = sQ + t * F_QTILE + row * 128;
        const int sw = row & 7;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            tmem_ld_x32(t_o + c * 32, o);
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2)
                pk[i >> 1] = Cvt<T>::pack2(__uint_as_float(o[i]) * inv_l, __uint_as_float(o[i + 1]) * inv_l);
            uint8_t* dst = o_row + (c >> 1) * F_QCHUNK;
            const int u0 = (c & 1) * 4;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                *reinterpret_cast<uint4*>(dst + (((u0 + u) ^ sw) << 4)) =
                    make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
        }
        if (lse != nullptr) lse[((long)blockIdx.x) * S + qrow] = (m_run + log2f(l_run)) * 0.6931471805599453f;
        fence_proxy_async_smem();
        named_bar_sync(1 + t, 128);
        if ((warp & 3) == 0 && lane == 0) {
            tma_store_4d(&tmO, sQ + t * F_QTILE, 0, h, q0 + t * F_BQ, b);
            tma_store_4d(&tmO, sQ + t * F_QTILE + F_QCHUNK, 64, h, q0 + t * F_BQ, b);
            tma_store_commit();
            tma_store_wait<0>();
        }
        }   // n > 0
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 3) tmem_dealloc<512>(tmem_base);
}

// ================================================================================================ backward
// nlse2[bh, s] = -lse * log2(e); ndelta[bh, s] = -scale * sum_d dO[b, s, h, d] * O[b, s, h, d]; one warp per row.
template <typename T>
__global__ void __launch_bounds__(256) flash_prep_kernel(const T* __restrict__ dO, const T* __restrict__ O,
                                                         const float* __restrict__ lse, float* __restrict__ nlse2,
                                                         float* __restrict__ ndelta, int B, int S, int H, long do_sb,
                                                         long do_ss, long do_sh, long o_sb, long o_ss, long o_sh,
                                                         float scale, int D) {
    const long rows = (long)B * S * H;
    const int lane = threadIdx.x & 31;
    for (long r = (long)blockIdx.x * 8 + (threadIdx.x >> 5); r < rows; r += (long)gridDim.x * 8) {
        const int h = (int)(r % H);
        const long t = r / H;
        const int s = (int)(t % S);
        const int b = (int)(t / S);
        float v = 0.f;
        if (lane * 4 < D) {      // head dims below 128: the rows are only D wide
            const uint2 a = *reinterpret_cast<const uint2*>(dO + b * do_sb + s * do_ss + h * do_sh + lane * 4);
            const uint2 c = *reinterpret_cast<const uint2*>(O + b * o_sb + s * o_ss + h * o_sh + lane * 4);
            const float2 a0 = Cvt<T>::unpack2(a.x), a1 = Cvt<T>::unpack2(a.y);
            const float2 c0 = Cvt<T>::unpack2(c.x), c1 = Cvt<T>::unpack2(c.y);
            v = a0.x * c0.x + a0.y * c0.y + a1.x * c1.x + a1.y * c1.y;
        }
        v = warp_sum(v);
        if (lane == 0) {
            const long o = ((long)b * H + h) * S + s;
            ndelta[o] = -scale * v;
            nlse2[o] = -1.4426950408889634f * lse[o];
        }
    }
}

// dq[b, s, h, :] (16 bit, strided) = dq_acc[b, h, :, s] (fp32, transposed accumulator of the backward kernel).
// One block per (32 queries, batch * head): coalesced 128-byte reads along s, padded shared tile, 256-byte row writes.
template <typename T>
__global__ void __launch_bounds__(256) flash_dq_finish_kernel(const float* __restrict__ acc, T* __restrict__ dq, int S,
                                                              int H, long sb, long ss, long sh,
                                                              const float* __restrict__ rot_cs,
                                                              const float* __restrict__ rot_sn, int rot, int D) {
    __shared__ float tile[128][33];
    const int s0 = blockIdx.x * 32;
    const int bh = blockIdx.y;
    const int b = bh / H, h = bh - b * H;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* src = acc + (long)bh * 128 * S + s0 + lane;
#pragma unroll
    for (int i = 0; i < 16; ++i) tile[w * 16 + i][lane] = src[(long)(w * 16 + i) * S];
    __syncthreads();
    T* dst = dq + b * sb + h * sh;
    const int hr = rot >> 1;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int s = w * 4 + i;
        float v[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int d = lane * 4 + t;
            float x = tile[d][s];
            if (d < rot) {   // inverse rotary: the accumulator holds the gradient of the rotated queries
                const int j = d < hr ? d : d - hr;
                const float c = rot_cs[(long)(s0 + s) * hr + j], sg = rot_sn[(long)(s0 + s) * hr + j];
                x = d < hr ? x * c + tile[d + hr][s] * sg : x * c - tile[d - hr][s] * sg;
            }
            v[t] = x;
        }
        uint2 o;
        o.x = Cvt<T>::pack2(v[0], v[1]);
        o.y = Cvt<T>::pack2(v[2], v[3]);
        if (lane * 4 < D) *reinterpret_cast<uint2*>(dst + (long)(s0 + s) * ss + lane * 4) = o;
    }
}

constexpr int B_OFF_K = 0;                          // [128 kv x 128 d] 32 KiB (two chunks of 16 KiB)
constexpr int B_OFF_V = B_OFF_K + 32768;
constexpr int B_OFF_Q = B_OFF_V + 32768;            // 3 x [64 q x 128 d] 16 KiB (two chunks of 8 KiB)
constexpr int B_OFF_DO = B_OFF_Q + 3 * 16384;       // 3 x 16 KiB
constexpr int B_OFF_DS = B_OFF_DO + 3 * 16384;      // [128 kv x 64 q] 16 KiB
constexpr int B_OFF_DQ = B_OFF_DS + 16384;          // fp32 dQ^T staging: 2 x [128 d x 32 q] 16 KiB, swizzled
constexpr int B_OFF_LD = B_OFF_DQ + 32768;          // 3 x (64 -lse2 | 64 -delta) fp32
constexpr int B_OFF_BAR = B_OFF_LD + 3 * 512;
constexpr int B_SMEM = B_OFF_BAR + 512 + 1024;
constexpr bool B_POLY = true;   // half of the exponentials as a polynomial on the FMA pipe
constexpr int TM_ST = 0, TM_DPT = 128, TM_DQT = 192, TM_DV = 256, TM_DK = 384;

template <typename T>
__global__ void __launch_bounds__(512, 1)
flash_bwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                  const __grid_constant__ CUtensorMap tmdK, const __grid_constant__ CUtensorMap tmdV,
                  const __grid_constant__ CUtensorMap tmDQ, const float* __restrict__ nlse2,
                  const float* __restrict__ ndelta, int S, int H, float scale, float scale_log2, int causal, int dbg,
                  long long* prof, const float* __restrict__ rot_cs, const float* __restrict__ rot_sn, int rot,
                  const int* __restrict__ sched, int sched_hstride, int sched_pstride,
                  const float* __restrict__ kpm, const Dropout drop, const int D) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sK = smem + B_OFF_K;
    uint8_t* sV = smem + B_OFF_V;
    uint8_t* sQ = smem + B_OFF_Q;
    uint8_t* sdO = smem + B_OFF_DO;
    uint8_t* sdS = smem + B_OFF_DS;
    float* sdQ = reinterpret_cast<float*>(smem + B_OFF_DQ);
    float* sLD = reinterpret_cast<float*>(smem + B_OFF_LD);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + B_OFF_BAR);
    uint64_t* kv_full = bars;            // [1]
    uint64_t* q_full = bars + 1;         // [3] Q tile + its -lse2 / -delta rows
    uint64_t* q_free = bars + 4;         // [3] dK UMMAs (last readers of the stage) completed
    uint64_t* do_full = bars + 7;        // [3]
    uint64_t* do_free = bars + 10;       // [3] dV UMMAs completed
    uint64_t* s_full = bars + 13;        // [2] S^T in TMEM
    uint64_t* dp_full = bars + 15;       // [1] dP^T in TMEM
    uint64_t* dp_free = bars + 16;       // [1] dP^T read into registers (8 warps)
    uint64_t* pds_ready = bars + 17;     // [2] P^T, dS^T stored in TMEM and dS^T in shared memory (8 warps)
    uint64_t* ds_free = bars + 19;       // [1] dQ^T UMMAs (readers of the shared dS^T tile) completed
    uint64_t* dq_full = bars + 20;       // [1] dQ^T in TMEM
    uint64_t* dq_free = bars + 21;       // [1] dQ^T read into registers (4 warps)
    uint64_t* fin = bars + 22;           // [1] everything issued has completed
    uint64_t* st_free = bars + 23;       // [2] dV / dK UMMAs (readers of P^T / dS^T in the score buffer) completed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 26);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x / H, h = blockIdx.x - b * H;
    const int jb = blockIdx.y;                       // key block; block 0 walks the most query blocks and starts first
    // Block-sparse mode: `sc` = [n, (64-row query block, visibility word) * n] for this (head, key block); bit
    // kgroup * 4 + qgroup of a word says whether 16-key group kgroup sees 16-query group qgroup of the 128 x 64 tile.
    const int* sc = sched != nullptr ? sched + (long)h * sched_hstride + (long)jb * sched_pstride : nullptr;
    const int i0 = causal ? jb * 2 : 0;              // first 64-row query block that sees these keys
    const int nit = sc ? sc[0] : S / 64 - i0;
    auto q_block = [&](int it) { return sc ? sc[1 + 2 * it] : i0 + it; };
    const bool tm = prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0;
    long long t0 = 0, acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmdO);
        tma_prefetch_desc(&tmDQ);
    }
    if (warp == 1 && lane == 0) {
        mbar_init(kv_full, 1);
        for (int i = 0; i < 3; ++i) {
            mbar_init(&q_full[i], 1);
            mbar_init(&q_free[i], 1);
            mbar_init(&do_full[i], 1);
            mbar_init(&do_free[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1);
            mbar_init(&pds_ready[i], 8);
            mbar_init(&st_free[i], 1);
        }
        mbar_init(ds_free, 1);
        mbar_init(dp_full, 1);
        mbar_init(dp_free, 8);
        mbar_init(dq_full, 1);
        mbar_init(dq_free, 4);
        mbar_init(fin, 1);
        fence_barrier_init();
    }
    if (warp == 3) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (elect_one()) {
            mbar_arrive_expect_tx(kv_full, 65536);
            tma_load_4d(sK, &tmK, kv_full, 0, h, jb * 128, b);
            tma_load_4d(sK + 16384, &tmK, kv_full, 64, h, jb * 128, b);
            tma_load_4d(sV, &tmV, kv_full, 0, h, jb * 128, b);
            tma_load_4d(sV + 16384, &tmV, kv_full, 64, h, jb * 128, b);
            const long bh_off = (long)blockIdx.x * S;
            for (int it = 0; it < nit; ++it) {
                const int q0 = q_block(it) * 64;
                const int s3 = it % 3;
                if (it >= 3) mbar_wait(&q_free[s3], ((it / 3) - 1) & 1);
                mbar_arrive_expect_tx(&q_full[s3], 16384 + 512);
                tma_load_4d(sQ + s3 * 16384, &tmQ, &q_full[s3], 0, h, q0, b);
                tma_load_4d(sQ + s3 * 16384 + 8192, &tmQ, &q_full[s3], 64, h, q0, b);
                bulk_load(sLD + s3 * 128, nlse2 + bh_off + q0, 256, &q_full[s3]);
                bulk_load(sLD + s3 * 128 + 64, ndelta + bh_off + q0, 256, &q_full[s3]);
                if (it >= 3) mbar_wait(&do_free[s3], ((it / 3) - 1) & 1);
                mbar_arrive_expect_tx(&do_full[s3], 16384);
                tma_load_4d(sdO + s3 * 16384, &tmdO, &do_full[s3], 0, h, q0, b);
                tma_load_4d(sdO + s3 * 16384 + 8192, &tmdO, &do_full[s3], 64, h, q0, b);
            }
        }
    } else if (warp == 1) {
        // Issuing warp A: the two score products of every block, as far ahead as their buffers allow.
        const uint32_t lead = elect_one() ? 1u : 0u;
        constexpr int is_bf16 = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
        const uint32_t id_st = make_idesc_f16(128, 64, is_bf16, 0, 0);    // K-major x K-major, N = 64 queries
        const int ksteps = (D + 15) >> 4;    // head dims below 128: the zero-filled columns are skipped
        const uint32_t k_km = desc_lo(sK, 16), v_km = desc_lo(sV, 16);
        auto koff_a = [](int k) { return (uint32_t)((k >> 2) * (16384 >> 4) + (k & 3) * 2); };
        auto koff_b = [](int k) { return (uint32_t)((k >> 2) * (8192 >> 4) + (k & 3) * 2); };
        mbar_wait(kv_full, 0);
        t0 = clock64();
        for (int n = 0; n < nit; ++n) {
            // S^T[kv, q] = K[kv, :] . Q[q, :] into score buffer n & 1 (free once dV / dK of block n - 2 have completed)
            mbar_wait(&q_full[n % 3], (n / 3) & 1);
            FA2_TICK(0)
            if (n >= 2) mbar_wait(&st_free[n & 1], ((n >> 1) - 1) & 1);
            tc_fence_after();
            FA2_TICK(1)
            {
                const uint32_t q_km = desc_lo(sQ + (n % 3) * 16384, 16);
                const uint32_t d = tmem_base + TM_ST + (n & 1) * 64;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < ksteps) umma_ss_p(lead, d, k_km + koff_a(k), q_km + koff_b(k), DESC_HI, id_st, k != 0);
                umma_commit_p(lead, &s_full[n & 1]);
            }
            FA2_TICK(2)
            // dP^T[kv, q] = V[kv, :] . dO[q, :]: one buffer, refilled once block n - 1 has been pulled into registers
            if (n >= 1) mbar_wait(dp_free, (n - 1) & 1);
            FA2_TICK(3)
            mbar_wait(&do_full[n % 3], (n / 3) & 1);
            tc_fence_after();
            FA2_TICK(4)
            {
                const uint32_t do_km = desc_lo(sdO + (n % 3) * 16384, 16);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < ksteps)
                        umma_ss_p(lead, tmem_base + TM_DPT, v_km + koff_a(k), do_km + koff_b(k), DESC_HI, id_st, k != 0);
                umma_commit_p(lead, dp_full);
            }
            FA2_TICK(5)
        }
        if (tm && lead) for (int i = 0; i < 6; ++i) prof[i] = acc[i];
    } else if (warp == 2) {
        // Issuing warp B: the three products that consume P^T / dS^T.
        // P^T / dS^T of query columns [32 u, 32 u + 32) live in columns 32 u + [0, 16) / 32 u + [16, 32) of the score buffer
        // (every softmax warp writes only over columns it has read itself); 16 queries = 8 columns per k-step.
        const uint32_t lead = elect_one() ? 1u : 0u;
        constexpr int is_bf16 = std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
        // A from TMEM, B MN-major, N = head dim (rounded up to 16; the accumulator columns beyond are never stored)
        const uint32_t id_acc = make_idesc_f16(128, ((D + 15) >> 4) * 16, is_bf16, 0, 1);
        const uint32_t id_dq = make_idesc_f16(128, 64, is_bf16, 1, 1);    // both MN-major, N = 64 queries
        const uint32_t k_mn = desc_lo(sK, 16384);
        const uint32_t ds_mn = desc_lo(sdS, 8192);
        t0 = clock64();
        const long long t_begin = t0;
        for (int it = 0; it < nit; ++it) {
            mbar_wait(&pds_ready[it & 1], (it >> 1) & 1);
            tc_fence_after();
            FA2_TICK(0)
            const uint32_t a = tmem_base + TM_ST + (it & 1) * 64;
            const uint32_t do_mn = desc_lo(sdO + (it % 3) * 16384, 8192);
            const uint32_t q_mn = desc_lo(sQ + (it % 3) * 16384, 8192);
            if (!(dbg & 32))
#pragma unroll
            for (int k = 0; k < 4; ++k)   // dV[kv, d] += P^T[kv, q] dO[q, d]
                umma_ts_p(lead, tmem_base + TM_DV, a + (k >> 1) * 32 + (k & 1) * 8, do_mn + k * (2048u >> 4), DESC_HI,
                          id_acc, (it | k) != 0);
            umma_commit_p(lead, &do_free[it % 3]);
            if (!(dbg & 32))
#pragma unroll
            for (int k = 0; k < 4; ++k)   // dK[kv, d] += dS^T[kv, q] Q[q, d]
                umma_ts_p(lead, tmem_base + TM_DK, a + 16 + (k >> 1) * 32 + (k & 1) * 8, q_mn + k * (2048u >> 4), DESC_HI,
                          id_acc, (it | k) != 0);
            umma_commit_p(lead, &q_free[it % 3]);
            umma_commit_p(lead, &st_free[it & 1]);
            FA2_TICK(1)
            if (it > 0) {
                mbar_wait(dq_free, (it - 1) & 1);
                tc_fence_after();
            }
            FA2_TICK(2)
            if (!(dbg & 16))
#pragma unroll
            for (int k = 0; k < 8; ++k)   // dQ^T[d, q] = K^T[d, kv] dS^T[kv, q]
                umma_ss_p(lead, tmem_base + TM_DQT, k_mn + k * (2048u >> 4), ds_mn + k * (2048u >> 4), DESC_HI, id_dq,
                          k != 0);
            umma_commit_p(lead, dq_full);
            umma_commit_p(lead, ds_free);
            FA2_TICK(3)
        }
        umma_commit_p(lead, fin);
        if (tm && lead) {
            for (int i = 0; i < 4; ++i) prof[8 + i] = acc[i];
            prof[30] = clock64() - t_begin;
            prof[31] = nit;
        }
    } else if (warp >= 4 && warp < 12) {
        const int qd = warp & 3;
        const int half = (warp - 4) >> 2;
        const int row = qd * 32 + lane;                       // key row of the block = TMEM lane
        const int kvg = jb * 128 + row;
        const uint32_t lane_field = static_cast<uint32_t>(qd * 32) << 16;
        const int sw = row & 7;
        const uint64_t sc2 = pk2(scale_log2, scale_log2);
        const uint64_t s2 = pk2(scale, scale);
        const float kp_row = kpm != nullptr ? kpm[(long)b * S + kvg] : 0.f;   // additive key mask (already in log2 units)
        // dropout: the forward's bits again. This thread's key selects the half of the pair hash; the query varies
        const uint32_t drop_key = drop_seed_bh(drop.seed, blockIdx.x) + (static_cast<uint32_t>(kvg) >> 1) * DROP_C2;
        const int drop_sh = (kvg & 1) * 16;
        const uint64_t ik2 = pk2(drop.inv_keep, drop.inv_keep);
        for (int it = 0; it < nit; ++it) {
            const int buf = it & 1;
            const int q0 = q_block(it) * 64 + half * 32;      // global query index of this thread's first column
            if (it == 0) t0 = clock64();
            mbar_wait(&q_full[it % 3], (it / 3) & 1);         // the -lse2 / -delta rows travel with the Q tile
            const float4* nl = reinterpret_cast<const float4*>(sLD + (it % 3) * 128 + half * 32);
            const float4* nd = reinterpret_cast<const float4*>(sLD + (it % 3) * 128 + 64 + half * 32);
            mbar_wait(&s_full[buf], (it >> 1) & 1);
            FA2_TICK(0)
            mbar_wait(dp_full, it & 1);
            tc_fence_after();
            FA2_TICK(1)
            const uint32_t t_s = tmem_base + lane_field + TM_ST + buf * 64 + half * 32;
            uint32_t rs[32], rp[32];
            if (!(dbg & 8)) {
                tmem_ld_x32(t_s, rs);
                tmem_ld_x32(tmem_base + lane_field + TM_DPT + half * 32, rp);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) rs[i] = rp[i] = 0x3c000000u + i + it;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dp_free);
            FA2_TICK(2)
            const bool diag = causal && it < 2 && sc == nullptr;
            // this thread's two 16-query groups of the tile: visible or not (block-sparse mode)
            const uint32_t vis = sc ? (static_cast<uint32_t>(sc[2 + 2 * it]) >> ((row >> 4) * 4 + half * 2)) & 3u : 3u;
            uint32_t pk[32];   // [0, 16) P^T, [16, 32) dS^T: the 32 columns this warp has just read
            if (dbg & 1) {
#pragma unroll
                for (int i = 0; i < 32; ++i) pk[i] = rs[i] ^ rp[i];
            } else
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                const float4 l4 = nl[i >> 2];
                const float4 d4 = nd[i >> 2];
                float p0, p1, p2, p3;
                upk2(fma2(pk2u(rs[i], rs[i + 1]), sc2, pk2(l4.x + kp_row, l4.y + kp_row)), p0, p1);
                const uint64_t x23 = fma2(pk2u(rs[i + 2], rs[i + 3]), sc2, pk2(l4.z + kp_row, l4.w + kp_row));
                p0 = ex2(p0);
                p1 = ex2(p1);
                if (B_POLY) {
                    float c2, c3;
                    upk2(x23, c2, c3);
                    exp2_poly2(pk2(fmaxf(c2, -125.f), fmaxf(c3, -125.f)), p2, p3);
                } else {
                    upk2(x23, p2, p3);
                    p2 = ex2(p2);
                    p3 = ex2(p3);
                }
                if (!((vis >> (i >> 4)) & 1u)) p0 = p1 = p2 = p3 = 0.f;
                if (diag) {
                    if (q0 + i < kvg) p0 = 0.f;
                    if (q0 + i + 1 < kvg) p1 = 0.f;
                    if (q0 + i + 2 < kvg) p2 = 0.f;
                    if (q0 + i + 3 < kvg) p3 = 0.f;
                }
                float a0, a1, b0, b1;
                if (drop.thr16) {
                    // dropped elements: no contribution to dV (P^T operand) and no gradient through dP; kept ones carry
                    // the 1 / keep factor on both
                    float g0, g1, g2, g3;
                    upk2(mul2(pk2u(rp[i], rp[i + 1]), ik2), g0, g1);
                    upk2(mul2(pk2u(rp[i + 2], rp[i + 3]), ik2), g2, g3);
                    const uint32_t qb = static_cast<uint32_t>(q0 + i) * DROP_C1 + drop_key;
                    const bool k0 = ((lowbias32(qb) >> drop_sh) & 0xffffu) >= drop.thr16;
                    const bool k1 = ((lowbias32(qb + DROP_C1) >> drop_sh) & 0xffffu) >= drop.thr16;
                    const bool k2 = ((lowbias32(qb + 2 * DROP_C1) >> drop_sh) & 0xffffu) >= drop.thr16;
                    const bool k3 = ((lowbias32(qb + 3 * DROP_C1) >> drop_sh) & 0xffffu) >= drop.thr16;
                    upk2(mul2(fma2(pk2(k0 ? g0 : 0.f, k1 ? g1 : 0.f), s2, pk2(d4.x, d4.y)), pk2(p0, p1)), a0, a1);
                    upk2(mul2(fma2(pk2(k2 ? g2 : 0.f, k3 ? g3 : 0.f), s2, pk2(d4.z, d4.w)), pk2(p2, p3)), b0, b1);
                    p0 = k0 ? p0 * drop.inv_keep : 0.f;
                    p1 = k1 ? p1 * drop.inv_keep : 0.f;
                    p2 = k2 ? p2 * drop.inv_keep : 0.f;
                    p3 = k3 ? p3 * drop.inv_keep : 0.f;
                } else {
                upk2(mul2(fma2(pk2u(rp[i], rp[i + 1]), s2, pk2(d4.x, d4.y)), pk2(p0, p1)), a0, a1);
                upk2(mul2(fma2(pk2u(rp[i + 2], rp[i + 3]), s2, pk2(d4.z, d4.w)), pk2(p2, p3)), b0, b1);
                }
                pk[i >> 1] = Cvt<T>::pack2(p0, p1);
                pk[(i >> 1) + 1] = Cvt<T>::pack2(p2, p3);
                pk[16 + (i >> 1)] = Cvt<T>::pack2(a0, a1);
                pk[17 + (i >> 1)] = Cvt<T>::pack2(b0, b1);
            }
            FA2_TICK(3)
            tmem_st_x32(t_s, pk);
            if (it >= 1) mbar_wait(ds_free, (it - 1) & 1);   // dQ^T of the previous block has finished reading the tile
            FA2_TICK(4)
            uint8_t* dst = sdS + row * 128;
#pragma unroll
            for (int u = 0; u < 4; ++u)
                *reinterpret_cast<uint4*>(dst + (((half * 4 + u) ^ sw) << 4)) =
                    make_uint4(pk[16 + u * 4], pk[17 + u * 4], pk[18 + u * 4], pk[19 + u * 4]);
            tmem_st_wait();
            tc_fence_before();
            if (!(dbg & 64)) fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&pds_ready[buf]);
            FA2_TICK(5)
        }
        if (tm && warp == 4 && lane == 0) for (int i = 0; i < 6; ++i) prof[16 + i] = acc[i];
        // dV / dK: TMEM (lane = key row) -> 16 bit -> shared (the K / V tiles are dead) -> TMA store
        mbar_wait(fin, 0);
        tc_fence_after();
#pragma unroll 1
        for (int which = 0; which < 2; ++which) {
            uint8_t* stage = (which == 0 ? sK : sV) + half * 16384 + row * 128;
            const uint32_t t_acc = tmem_base + lane_field + (which == 0 ? TM_DV : TM_DK) + half * 64;
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t r[32];
                tmem_ld_x32(t_acc + c * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    *reinterpret_cast<uint4*>(stage + (((c * 4 + u) ^ sw) << 4)) = make_uint4(
                        Cvt<T>::pack2(__uint_as_float(r[u * 8]), __uint_as_float(r[u * 8 + 1])),
                        Cvt<T>::pack2(__uint_as_float(r[u * 8 + 2]), __uint_as_float(r[u * 8 + 3])),
                        Cvt<T>::pack2(__uint_as_float(r[u * 8 + 4]), __uint_as_float(r[u * 8 + 5])),
                        Cvt<T>::pack2(__uint_as_float(r[u * 8 + 6]), __uint_as_float(r[u * 8 + 7])));
            }
        }
        if (rot > 0) {
            // dK is the gradient of the ROTATED keys: turn it back (inverse rotary, transpose of the forward rotation)
            // in the staged tile. Pairs (j, j + rot / 2) of a row may sit in different 64-column chunks, hence the
            // barrier; every thread then rotates whole 8-element units of its own row.
            named_bar_sync(1, 256);
            const int hr = rot >> 1;
            const float* cs = rot_cs + (long)(jb * 128 + row) * hr;
            const float* sn = rot_sn + (long)(jb * 128 + row) * hr;
            uint8_t* rowp = sV + row * 128;
            for (int j = half * 8; j < hr; j += 16) {
                const int d1 = j + hr;
                uint4* p0 = reinterpret_cast<uint4*>(rowp + (j >> 6) * 16384 + ((((j & 63) >> 3) ^ sw) << 4));
                uint4* p1 = reinterpret_cast<uint4*>(rowp + (d1 >> 6) * 16384 + ((((d1 & 63) >> 3) ^ sw) << 4));
                float x0[8], x1[8], y0[8], y1[8];
                unpack8<T>(*p0, x0);
                unpack8<T>(*p1, x1);
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const float c = cs[j + t], sg = sn[j + t];
                    y0[t] = x0[t] * c + x1[t] * sg;
                    y1[t] = x1[t] * c - x0[t] * sg;
                }
                *p0 = pack8<T>(y0);
                *p1 = pack8<T>(y1);
            }
        }
        fence_proxy_async_smem();
        named_bar_sync(1, 256);
        if (warp == 4 && lane == 0) {
            tma_store_4d(&tmdV, sK, 0, h, jb * 128, b);
            tma_store_4d(&tmdV, sK + 16384, 64, h, jb * 128, b);
            tma_store_4d(&tmdK, sV, 0, h, jb * 128, b);
            tma_store_4d(&tmdK, sV + 16384, 64, h, jb * 128, b);
            tma_store_commit();
            tma_store_wait<0>();
        }
    } else if (warp >= 12) {
        const int qd = warp & 3;
        const int d = qd * 32 + lane;                          // head-dim index = TMEM lane of dQ^T
        const uint32_t lane_field = static_cast<uint32_t>(qd * 32) << 16;
        for (int it = 0; it < nit; ++it) {
            const int q0 = q_block(it) * 64;
            if (it == 0) t0 = clock64();
            mbar_wait(dq_full, it & 1);
            tc_fence_after();
            FA2_TICK(0)
            uint32_t r[64];
            tmem_ld_x32(tmem_base + lane_field + TM_DQT, r);
            tmem_ld_x32(tmem_base + lane_field + TM_DQT + 32, r + 32);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(dq_free);
            FA2_TICK(1)
            if (it > 0) {   // the previous reduction must have finished reading the staging tile
                if (warp == 12 && lane == 0) tma_store_wait_read<0>();
                named_bar_sync(3, 128);
            }
            FA2_TICK(2)
            if (dbg & 2) continue;
            // The accumulator in global memory is kept TRANSPOSED ([B, H, D, S]) so that this thread's 64 values are
            // contiguous: 16 vector stores into two 128-byte-swizzled [128 d x 32 q] tiles instead of 64 scalar stores
            // (the proxy fence that follows costs time in proportion to the stores in flight: ~1000 cycles with 64).
            {
                uint8_t* base = reinterpret_cast<uint8_t*>(sdQ) + d * 128;
                const int sw = d & 7;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh)
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        *reinterpret_cast<uint4*>(base + hh * 16384 + ((u ^ sw) << 4)) =
                            make_uint4(r[hh * 32 + u * 4], r[hh * 32 + u * 4 + 1], r[hh * 32 + u * 4 + 2],
                                       r[hh * 32 + u * 4 + 3]);
            }
            if (!(dbg & 64)) fence_proxy_async_smem();
            named_bar_sync(3, 128);
            if (warp == 12 && lane == 0 && !(dbg & 4)) {
                tma_reduce_add_4d(&tmDQ, sdQ, q0, 0, h, b);
                tma_reduce_add_4d(&tmDQ, reinterpret_cast<uint8_t*>(sdQ) + 16384, q0 + 32, 0, h, b);
                tma_store_commit();
            }
            FA2_TICK(3)
        }
        if (tm && warp == 12 && lane == 0) for (int i = 0; i < 4; ++i) prof[24 + i] = acc[i];
        if (warp == 12 && lane == 0) tma_store_wait<0>();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 3) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ host launchers
// Experiment switches (0 in production): the kernels skip parts of their work so that the remaining parts can be timed
// alone (results are then garbage). Forward: 1 = no softmax arithmetic, 2 = no TMEM score load, 4 = no P V products.
// (forward, timing only: 16 = score UMMAs issued with N = 128, 32 = K / V tiles not reloaded after the first ring fill)
// Backward: 1 = no P / dS arithmetic, 2 = dQ drain stops after its TMEM load, 4 = no dQ reduce-add, 8 = no TMEM loads
// in the softmax warps, 16 = no dQ^T UMMAs, 32 = no dV / dK UMMAs, 64 = no proxy fences.
static int g_dbg = 0;
static long long* g_prof = nullptr;
void set_debug_mode(int m) { g_dbg = m; }
void set_debug_timers(long long* p) { g_prof = p; }
int launch_fwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& to, float* lse,
               int B, int S, int H, float scale_log2, int causal, int bf16, const int* sched, int sched_hstride,
               int sched_pstride, const float* kpm, Dropout drop, cudaStream_t stream, int D) {
    if (S % 128) return DSB_ERR_SIZE;
    dim3 grid(B * H, (S + 255) / 256);
    cudaError_t e;
    if (bf16) {
        auto kern = flash_fwd2_kernel<__nv_bfloat16>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 384, F_SMEM, stream>>>(tq, tk, tv, to, lse, S, H, scale_log2, causal, g_dbg, g_prof, sched, sched_hstride,
                                            sched_pstride, kpm, drop, D);
    } else {
        auto kern = flash_fwd2_kernel<__half>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, F_SMEM);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 384, F_SMEM, stream>>>(tq, tk, tv, to, lse, S, H, scale_log2, causal, g_dbg, g_prof, sched, sched_hstride,
                                            sched_pstride, kpm, drop, D);
    }
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

int launch_prep(const void* dO, const void* O, const float* lse, float* nlse2, float* ndelta, int B, int S, int H,
                const long* s_do, const long* s_o, float scale, int bf16, cudaStream_t stream, int D) {
    const long rows = (long)B * S * H;
    const int grid = (int)((rows + 7) / 8 < 148L * 16 ? (rows + 7) / 8 : 148L * 16);
    if (bf16)
        flash_prep_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)dO, (const __nv_bfloat16*)O, lse,
                                                                   nlse2, ndelta, B, S, H, s_do[0], s_do[1], s_do[2],
                                                                   s_o[0], s_o[1], s_o[2], scale, D);
    else
        flash_prep_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)dO, (const __half*)O, lse, nlse2, ndelta, B, S,
                                                            H, s_do[0], s_do[1], s_do[2], s_o[0], s_o[1], s_o[2], scale, D);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

int launch_dq_finish(const float* acc, void* dq, int B, int S, int H, const long* s_dq, int bf16, const float* rot_cs,
                     const float* rot_sn, int rot, cudaStream_t stream, int D) {
    dim3 grid(S / 32, B * H);
    if (bf16)
        flash_dq_finish_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(acc, (__nv_bfloat16*)dq, S, H, s_dq[0], s_dq[1],
                                                                        s_dq[2], rot_cs, rot_sn, rot, D);
    else
        flash_dq_finish_kernel<__half><<<grid, 256, 0, stream>>>(acc, (__half*)dq, S, H, s_dq[0], s_dq[1], s_dq[2],
                                                                 rot_cs, rot_sn, rot, D);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

int launch_bwd(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& tdo,
               const CUtensorMap& tdk, const CUtensorMap& tdv, const CUtensorMap& tdq, const float* nlse2,
               const float* ndelta, int B, int S, int H, float scale, float scale_log2, int causal, int bf16,
               const float* rot_cs, const float* rot_sn, int rot, const int* sched, int sched_hstride, int sched_pstride,
               const float* kpm, Dropout drop, cudaStream_t stream, int D) {
    if (S % 128) return DSB_ERR_SIZE;
    if (rot < 0 || rot > 128 || (rot % 16)) return DSB_ERR_ARG;
    dim3 grid(B * H, S / 128);
    cudaError_t e;
    if (bf16) {
        auto kern = flash_bwd2_kernel<__nv_bfloat16>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, B_SMEM);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 512, B_SMEM, stream>>>(tq, tk, tv, tdo, tdk, tdv, tdq, nlse2, ndelta, S, H, scale, scale_log2,
                                            causal, g_dbg, g_prof, rot_cs, rot_sn, rot, sched, sched_hstride, sched_pstride, kpm, drop, D);
    } else {
        auto kern = flash_bwd2_kernel<__half>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, B_SMEM);
        if (e != cudaSuccess) return -(int)e;
        kern<<<grid, 512, B_SMEM, stream>>>(tq, tk, tv, tdo, tdk, tdv, tdq, nlse2, ndelta, S, H, scale, scale_log2,
                                            causal, g_dbg, g_prof, rot_cs, rot_sn, rot, sched, sched_hstride, sched_pstride, kpm, drop, D);
    }
    e = cudaGetLastError();
    return e == cudaSuccess ? 0 : -(int)e;
}

}  // namespace fa2
}  // namespace dsb
