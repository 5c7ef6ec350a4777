// GEMM tile-policy autotuner + dispatching entry point.
//
// The reference times cuBLAS algorithm ids 99..115 for the forward / dW / dX GEMMs of its transformer layer and keeps
// the best triple (csrc/includes/gemm_test.h:27-293). Our GEMMs are our own kernels, so the tunable is the kernel
// variant: single-CTA tiles, CTA-pair tiles, and the tail-wave K split of the pair kernel. `dsb_gemm_tune` times every
// variant for one problem (shape, layouts, dtype) on scratch operands with CUDA events and records the winner;
// `dsb_gemm_auto` is the entry point the framework calls: tuned problems use their recorded variant, everything else
// a static rule.
#include <cuda_runtime.h>
#include <atomic>
#include <mutex>
#include <unordered_map>
#include <cstdint>
#include "../../include/dsb200.h"

extern "C" int dsb_gemm(const void*, const void*, void*, int, int, int, int, long, long, long, long, long, long, int,
                        int, int, int, float, const void*, const void*, long, long, void*, long, long, void*);
extern "C" int dsb_gemm_pair_ex(const void*, const void*, void*, int, int, int, int, long, long, long, long, long,
                                long, int, int, int, int, float, const void*, const void*, long, long, void*, long,
                                long, int, void*);

namespace {

// variant id: 0 = single-CTA kernel; 1..4 = pair kernel with that K split of the tail wave (1 = whole tiles);
// 5 = pair kernel choosing the split itself.
constexpr int N_VARIANTS = 6;

struct Key {
    int M, N, K, batch, layout, dtype;
    bool operator==(const Key& o) const {
        return M == o.M && N == o.N && K == o.K && batch == o.batch && layout == o.layout && dtype == o.dtype;
    }
};
struct KeyHash {
    size_t operator()(const Key& k) const {
        uint64_t h = 1469598103934665603ull;
        const int v[6] = {k.M, k.N, k.K, k.batch, k.layout, k.dtype};
        for (int x : v) { h ^= (uint64_t)(uint32_t)x; h *= 1099511628211ull; }
        return (size_t)h;
    }
};
std::unordered_map<Key, int, KeyHash> g_table;
std::mutex g_mu;
std::atomic<bool> g_nonempty{false};      // read without the lock on the dispatch path

int run_variant(int variant, const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb,
                long ldd, long sA, long sB, long sD, int a_mn, int b_mn, int dtype, int epi, float alpha,
                const void* bias, const void* aux, long ldaux, long s_aux, void* aux_out, long ldaux_out,
                long s_aux_out, void* stream) {
    if (variant == 0)
        return dsb_gemm(A, B, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, a_mn, b_mn, dtype, epi, alpha, bias, aux,
                        ldaux, s_aux, aux_out, ldaux_out, s_aux_out, stream);
    return dsb_gemm_pair_ex(A, B, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, a_mn, b_mn, dtype, epi, alpha, bias,
                            aux, ldaux, s_aux, aux_out, ldaux_out, s_aux_out, variant == 5 ? -1 : variant, stream);
}

int static_rule(int M, int N, int batch) {
    const long units = (long)((M + 255) / 256) * ((N + 127) / 128) * batch;
    return (M >= 256 && N >= 128 && units >= 64) ? 5 : 0;
}

}  // namespace

// Time every kernel variant for one problem and remember the fastest. ms_out (optional, N_VARIANTS floats) receives the
// per-variant median time in ms (-1 where a variant does not apply). Returns the winning variant id or a negative error.
extern "C" int dsb_gemm_tune(int M, int N, int K, int batch, int a_mn, int b_mn, int dtype, int iters, float* ms_out,
                             void* stream_) {
    if (M <= 0 || N <= 0 || K <= 0 || batch <= 0 || iters <= 0 || iters > 64) return DSB_ERR_ARG;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const size_t ea = (size_t)M * K * batch * 2, eb = (size_t)N * K * batch * 2, ed = (size_t)M * N * batch * 2;
    void *A = nullptr, *B = nullptr, *D = nullptr;
    if (cudaMalloc(&A, ea) != cudaSuccess || cudaMalloc(&B, eb) != cudaSuccess || cudaMalloc(&D, ed) != cudaSuccess) {
        cudaFree(A); cudaFree(B); cudaFree(D);
        cudaGetLastError();
        return DSB_ERR_ARG;
    }
    cudaMemsetAsync(A, 0x3c, ea, stream);   // 0x3c3c = 1.06 (fp16) / 0.0115 (bf16): finite, non-zero operands
    cudaMemsetAsync(B, 0x3c, eb, stream);
    const long lda = a_mn ? M : K, ldb = b_mn ? N : K, ldd = N;
    const long sA = (long)M * K, sB = (long)N * K, sD = (long)M * N;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    int best = -1;
    float best_ms = 1e30f;
    for (int v = 0; v < N_VARIANTS; ++v) {
        float med = -1.f;
        if (v >= 1 && v <= 4 && (M < 256 || N < 128)) { if (ms_out) ms_out[v] = med; continue; }
        if (v == 5 && (M < 256 || N < 128)) { if (ms_out) ms_out[v] = med; continue; }
        int rc = 0;
        for (int w = 0; w < 2 && rc == 0; ++w)
            rc = run_variant(v, A, B, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, a_mn, b_mn, dtype, DSB_EPI_NONE,
                             1.f, nullptr, nullptr, 0, 0, nullptr, 0, 0, stream_);
        if (rc == 0) {
            float t[64];
            for (int i = 0; i < iters && rc == 0; ++i) {
                cudaEventRecord(e0, stream);
                rc = run_variant(v, A, B, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, a_mn, b_mn, dtype,
                                 DSB_EPI_NONE, 1.f, nullptr, nullptr, 0, 0, nullptr, 0, 0, stream_);
                cudaEventRecord(e1, stream);
                cudaEventSynchronize(e1);
                cudaEventElapsedTime(&t[i], e0, e1);
            }
            if (rc == 0) {
                for (int i = 1; i < iters; ++i)   // insertion sort, iters <= 64
                    for (int j = i; j > 0 && t[j] < t[j - 1]; --j) { float x = t[j]; t[j] = t[j - 1]; t[j - 1] = x; }
                med = t[iters / 2];
                if (med < best_ms) { best_ms = med; best = v; }
            }
        }
        if (ms_out) ms_out[v] = med;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaStreamSynchronize(stream);
    cudaFree(A); cudaFree(B); cudaFree(D);
    if (best < 0) return DSB_ERR_ARG;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        g_table[Key{M, N, K, batch, a_mn * 2 + b_mn, dtype}] = best;
        g_nonempty.store(true, std::memory_order_release);
    }
    return best;
}

// Variant recorded for a problem (-1 if it was never tuned).
extern "C" int dsb_gemm_tuned_variant(int M, int N, int K, int batch, int a_mn, int b_mn, int dtype) {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_table.find(Key{M, N, K, batch, a_mn * 2 + b_mn, dtype});
    return it == g_table.end() ? -1 : it->second;
}

extern "C" void dsb_gemm_tune_clear(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_table.clear();
    g_nonempty.store(false, std::memory_order_release);
}

// Same contract as dsb_gemm; picks the kernel variant (tuned table first, static rule otherwise).
extern "C" int dsb_gemm_auto(const void* A, const void* B, void* D, int M, int N, int K, int batch, long lda, long ldb,
                             long ldd, long strideA, long strideB, long strideD, int a_mn_major, int b_mn_major,
                             int dtype, int epilogue, float alpha, const void* bias, const void* aux, long ldaux,
                             long stride_aux, void* aux_out, long ldaux_out, long stride_aux_out, void* stream) {
    int variant = -1;
    if (g_nonempty.load(std::memory_order_acquire)) {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_table.find(Key{M, N, K, batch, a_mn_major * 2 + b_mn_major, dtype});
        if (it != g_table.end()) variant = it->second;
    }
    if (variant < 0) variant = static_rule(M, N, batch);
    return run_variant(variant, A, B, D, M, N, K, batch, lda, ldb, ldd, strideA, strideB, strideD, a_mn_major,
                       b_mn_major, dtype, epilogue, alpha, bias, aux, ldaux, stride_aux, aux_out, ldaux_out,
                       stride_aux_out, stream);
}
