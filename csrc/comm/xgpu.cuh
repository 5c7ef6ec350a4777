// Cross-GPU building blocks shared by the peer-memory kernels (fused ZeRO step, 1-bit all-reduce):
//   * peer pointer tables (one unicast address per rank of the group, this rank included),
//   * epoch-flag barrier and scalar mailboxes inside each rank's symmetric-heap flag block,
//   * NVLS: multimem.ld_reduce / multimem.st on the MULTICAST address of the heap (NVSwitch reduces 16-byte vectors of
//     every rank's copy in the switch and returns one result; a multimem store lands in every rank's copy). These are
//     plain load / store instructions on a multicast mapping - not TMA.
//
// Flag block layout (4 KiB = 1024 u32 at the start of every rank's heap):
//   [slot * MAX_RANKS + rank]              slot 0,1: barrier arrivals (epoch values)
//   [MAILBOX_BASE + rank * 4 + {0,1,2}]    scalar mailbox written by `rank`: two payload words, then the epoch
#pragma once
#include "../common/ptx.cuh"

namespace dsb {

constexpr int MAX_RANKS = 16;
constexpr int MAILBOX_BASE = 64;

struct PeerPtrs {
    void* p[MAX_RANKS];
};
struct FlagPtrs {
    uint32_t* p[MAX_RANKS];
};

// All ranks arrive at `slot` with value `epoch`; returns when every rank has arrived. Called by one CTA.
DSB_DEVICE void xgpu_barrier(const FlagPtrs& flags, int R, int me, int slot, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x < R) {
        __threadfence_system();
        st_release_sys(flags.p[threadIdx.x] + slot * MAX_RANKS + me, epoch);
        const uint32_t* mine = flags.p[me] + slot * MAX_RANKS + threadIdx.x;
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {}
    }
    __syncthreads();
}

// Every rank contributes two floats; every rank gets the sums, added in rank order (bitwise identical everywhere).
// Called by one CTA (threads 0..R-1 do the work); also a barrier: returning means every rank has reached this point.
DSB_DEVICE void xgpu_sum2(const FlagPtrs& flags, int R, int me, uint32_t epoch, float a, float b, float* out2) {
    __shared__ float box[2 * MAX_RANKS];
    __syncthreads();
    if (threadIdx.x < R) {
        uint32_t* dst = flags.p[threadIdx.x] + MAILBOX_BASE + me * 4;
        dst[0] = __float_as_uint(a);
        dst[1] = __float_as_uint(b);
        __threadfence_system();
        st_release_sys(dst + 2, epoch);
        const uint32_t* src = flags.p[me] + MAILBOX_BASE + threadIdx.x * 4;
        while ((int32_t)(ld_acquire_sys(src + 2) - epoch) < 0) {}
        box[2 * threadIdx.x] = __uint_as_float(ld_relaxed_sys(src));
        box[2 * threadIdx.x + 1] = __uint_as_float(ld_relaxed_sys(src + 1));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float sa = 0.f, sb = 0.f;
        for (int r = 0; r < R; ++r) {
            sa += box[2 * r];
            sb += box[2 * r + 1];
        }
        out2[0] = sa;
        out2[1] = sb;
    }
    __syncthreads();
}

// ---- NVLS (multimem) -------------------------------------------------------------------------------------------
// Sum of the 8 half / bf16 values at `mc` over every rank's copy, accumulated in fp32 inside the switch.
template <typename T>
DSB_DEVICE uint4 multimem_ld_reduce_add8(const void* mc);
template <>
DSB_DEVICE uint4 multimem_ld_reduce_add8<__half>(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
template <>
DSB_DEVICE uint4 multimem_ld_reduce_add8<__nv_bfloat16>(const void* mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(mc)
                 : "memory");
    return r;
}
// 16 bytes stored into every rank's copy.
DSB_DEVICE void multimem_st_v4(void* mc, const uint4& v) {
    // the element type of a multimem store only names the register width; the 16 bytes are stored as they are
    asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

}  // namespace dsb
