// Host-side Adam/AdamW for ZeRO-Offload: AVX-512 / AVX2 / scalar kernels selected at run time (works on AMD and
// Intel alike), OpenMP over cache-sized chunks, and an optional fused low-precision copy-back: each updated tile
// is converted to fp16/bf16 into one of two pinned staging buffers and shipped with cudaMemcpyAsync on alternating
// streams, so the H2D copy of tile t overlaps the CPU update of tile t+1.
// Reference behaviour: csrc/adam/cpu_adam.cpp:21-690, csrc/includes/cpu_adam.h:23-163 (math in SURVEY C.6).
// The CUDA runtime is resolved lazily with dlopen so the library also loads on CPU-only hosts.
#include <dlfcn.h>
#include <immintrin.h>
#include <math.h>
#include <omp.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <map>
#include <mutex>

namespace {

struct AdamState {
    float lr, beta1, beta2, eps, wd;
    int adamw;
    float beta1_t = 1.f, beta2_t = 1.f;
    long step = 0;
};
std::map<int, AdamState> g_opts;
std::mutex g_mu;

// ---- lazy CUDA runtime ---------------------------------------------------------------------------------------
typedef int (*fn_malloc_host)(void**, size_t);
typedef int (*fn_free_host)(void*);
typedef int (*fn_memcpy_async)(void*, const void*, size_t, int, void*);
typedef int (*fn_stream_create)(void**);
typedef int (*fn_stream_sync)(void*);
struct Cuda {
    bool ok = false;
    fn_malloc_host mallocHost = nullptr;
    fn_free_host freeHost = nullptr;
    fn_memcpy_async memcpyAsync = nullptr;
    fn_stream_create streamCreate = nullptr;
    fn_stream_sync streamSync = nullptr;
    void* streams[2] = {nullptr, nullptr};
    void* staging[2] = {nullptr, nullptr};
    size_t staging_bytes = 0;
};
Cuda g_cuda;
const long TILE = 1L << 24;  // elements per staged tile (32 MiB of fp16)

bool cuda_init() {
    if (g_cuda.ok) return true;
    void* h = dlopen("libcudart.so.12", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libcudart.so.12", RTLD_NOW);
    if (!h) h = dlopen("libcudart.so", RTLD_NOW);
    if (!h) return false;
    g_cuda.mallocHost = (fn_malloc_host)dlsym(h, "cudaMallocHost");
    g_cuda.freeHost = (fn_free_host)dlsym(h, "cudaFreeHost");
    g_cuda.memcpyAsync = (fn_memcpy_async)dlsym(h, "cudaMemcpyAsync");
    g_cuda.streamCreate = (fn_stream_create)dlsym(h, "cudaStreamCreate");
    g_cuda.streamSync = (fn_stream_sync)dlsym(h, "cudaStreamSynchronize");
    if (!g_cuda.mallocHost || !g_cuda.memcpyAsync || !g_cuda.streamCreate || !g_cuda.streamSync) return false;
    for (int i = 0; i < 2; ++i) {
        if (g_cuda.streamCreate(&g_cuda.streams[i]) != 0) return false;
        if (g_cuda.mallocHost(&g_cuda.staging[i], TILE * 2) != 0) return false;
    }
    g_cuda.staging_bytes = TILE * 2;
    g_cuda.ok = true;
    return true;
}

// ---- conversions -------------------------------------------------------------------------------------------
inline uint16_t f32_to_bf16(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffff) > 0x7f800000) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fff + ((u >> 16) & 1);
    return (uint16_t)(u >> 16);
}
inline uint16_t f32_to_f16(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000;
    int32_t e = (int32_t)((x >> 23) & 0xff) - 127 + 15;
    uint32_t m = x & 0x7fffff;
    if (((x >> 23) & 0xff) == 0xff) return (uint16_t)(sign | 0x7c00 | (m ? 0x200 : 0));
    if (e >= 31) return (uint16_t)(sign | 0x7c00);
    if (e <= 0) {
        if (e < -10) return (uint16_t)sign;
        m |= 0x800000;
        uint32_t shift = (uint32_t)(14 - e);
        uint32_t r = m >> shift;
        uint32_t rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (r & 1))) r++;
        return (uint16_t)(sign | r);
    }
    uint32_t r = (uint32_t)(e << 10) | (m >> 13);
    uint32_t rem = m & 0x1fff;
    if (rem > 0x1000 || (rem == 0x1000 && (r & 1))) r++;
    return (uint16_t)(sign | r);
}

struct Coef {
    float beta1, beta2, omb1, omb2, eps, wd, step_size, bc2_rsqrt, lr_wd;
    int adamw;
};

// ---- scalar kernel -------------------------------------------------------------------------------------------
void adam_scalar(float* p, const float* g, float* m, float* v, long n, const Coef& c, uint16_t* lp, int lp_bf16) {
    for (long i = 0; i < n; ++i) {
        float grad = g[i], param = p[i];
        if (!c.adamw && c.wd > 0.f) grad += c.wd * param;
        float mm = c.beta1 * m[i] + c.omb1 * grad;
        float vv = c.beta2 * v[i] + c.omb2 * grad * grad;
        float denom = sqrtf(vv) * c.bc2_rsqrt + c.eps;
        if (c.adamw && c.wd > 0.f) param += c.lr_wd * param;
        param += c.step_size * (mm / denom);
        p[i] = param; m[i] = mm; v[i] = vv;
        if (lp) lp[i] = lp_bf16 ? f32_to_bf16(param) : f32_to_f16(param);
    }
}

// ---- AVX2 kernel ----------------------------------------------------------------------------------------------
__attribute__((target("avx2,fma,f16c")))
void adam_avx2(float* p, const float* g, float* m, float* v, long n, const Coef& c, uint16_t* lp, int lp_bf16) {
    const __m256 b1 = _mm256_set1_ps(c.beta1), b2 = _mm256_set1_ps(c.beta2), ob1 = _mm256_set1_ps(c.omb1),
                 ob2 = _mm256_set1_ps(c.omb2), eps = _mm256_set1_ps(c.eps), wd = _mm256_set1_ps(c.wd),
                 ss = _mm256_set1_ps(c.step_size), bc2 = _mm256_set1_ps(c.bc2_rsqrt), lrwd = _mm256_set1_ps(c.lr_wd);
    long i = 0;
    for (; i + 8 <= n; i += 8) {
        __m256 grad = _mm256_loadu_ps(g + i), param = _mm256_loadu_ps(p + i);
        __m256 mm = _mm256_loadu_ps(m + i), vv = _mm256_loadu_ps(v + i);
        if (!c.adamw && c.wd > 0.f) grad = _mm256_fmadd_ps(wd, param, grad);
        mm = _mm256_fmadd_ps(ob1, grad, _mm256_mul_ps(b1, mm));
        vv = _mm256_fmadd_ps(ob2, _mm256_mul_ps(grad, grad), _mm256_mul_ps(b2, vv));
        __m256 denom = _mm256_fmadd_ps(_mm256_sqrt_ps(vv), bc2, eps);
        if (c.adamw && c.wd > 0.f) param = _mm256_fmadd_ps(lrwd, param, param);
        param = _mm256_fmadd_ps(ss, _mm256_div_ps(mm, denom), param);
        _mm256_storeu_ps(p + i, param);
        _mm256_storeu_ps(m + i, mm);
        _mm256_storeu_ps(v + i, vv);
        if (lp) {
            if (!lp_bf16) {
                _mm_storeu_si128((__m128i*)(lp + i), _mm256_cvtps_ph(param, _MM_FROUND_TO_NEAREST_INT));
            } else {
                float tmp[8];
                _mm256_storeu_ps(tmp, param);
                for (int k = 0; k < 8; ++k) lp[i + k] = f32_to_bf16(tmp[k]);
            }
        }
    }
    if (i < n) adam_scalar(p + i, g + i, m + i, v + i, n - i, c, lp ? lp + i : nullptr, lp_bf16);
}

// ---- AVX-512 kernel -------------------------------------------------------------------------------------------
__attribute__((target("avx512f,avx512bw,avx512vl,fma,f16c")))
void adam_avx512(float* p, const float* g, float* m, float* v, long n, const Coef& c, uint16_t* lp, int lp_bf16) {
    const __m512 b1 = _mm512_set1_ps(c.beta1), b2 = _mm512_set1_ps(c.beta2), ob1 = _mm512_set1_ps(c.omb1),
                 ob2 = _mm512_set1_ps(c.omb2), eps = _mm512_set1_ps(c.eps), wd = _mm512_set1_ps(c.wd),
                 ss = _mm512_set1_ps(c.step_size), bc2 = _mm512_set1_ps(c.bc2_rsqrt), lrwd = _mm512_set1_ps(c.lr_wd);
    long i = 0;
    for (; i + 16 <= n; i += 16) {
        __m512 grad = _mm512_loadu_ps(g + i), param = _mm512_loadu_ps(p + i);
        __m512 mm = _mm512_loadu_ps(m + i), vv = _mm512_loadu_ps(v + i);
        if (!c.adamw && c.wd > 0.f) grad = _mm512_fmadd_ps(wd, param, grad);
        mm = _mm512_fmadd_ps(ob1, grad, _mm512_mul_ps(b1, mm));
        vv = _mm512_fmadd_ps(ob2, _mm512_mul_ps(grad, grad), _mm512_mul_ps(b2, vv));
        __m512 denom = _mm512_fmadd_ps(_mm512_sqrt_ps(vv), bc2, eps);
        if (c.adamw && c.wd > 0.f) param = _mm512_fmadd_ps(lrwd, param, param);
        param = _mm512_fmadd_ps(ss, _mm512_div_ps(mm, denom), param);
        _mm512_storeu_ps(p + i, param);
        _mm512_storeu_ps(m + i, mm);
        _mm512_storeu_ps(v + i, vv);
        if (lp) {
            if (!lp_bf16) {
                _mm256_storeu_si256((__m256i*)(lp + i), _mm512_cvtps_ph(param, _MM_FROUND_TO_NEAREST_INT));
            } else {
                // round-to-nearest-even bf16: add 0x7fff + lsb, take the high half
                __m512i u = _mm512_castps_si512(param);
                __m512i lsb = _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1));
                u = _mm512_add_epi32(u, _mm512_add_epi32(lsb, _mm512_set1_epi32(0x7fff)));
                _mm256_storeu_si256((__m256i*)(lp + i), _mm512_cvtepi32_epi16(_mm512_srli_epi32(u, 16)));
            }
        }
    }
    if (i < n) adam_scalar(p + i, g + i, m + i, v + i, n - i, c, lp ? lp + i : nullptr, lp_bf16);
}

typedef void (*kernel_fn)(float*, const float*, float*, float*, long, const Coef&, uint16_t*, int);
int g_simd_level = -1;  // 0 scalar, 1 avx2, 2 avx512
kernel_fn pick_kernel() {
    if (g_simd_level < 0) {
        __builtin_cpu_init();
        if (__builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512vl"))
            g_simd_level = 2;
        else if (__builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma") && __builtin_cpu_supports("f16c"))
            g_simd_level = 1;
        else
            g_simd_level = 0;
    }
    return g_simd_level == 2 ? adam_avx512 : (g_simd_level == 1 ? adam_avx2 : adam_scalar);
}

void run_range(float* p, const float* g, float* m, float* v, long n, const Coef& c, uint16_t* lp, int lp_bf16) {
    kernel_fn k = pick_kernel();
    const long CHUNK = 1L << 16;
    const long nchunks = (n + CHUNK - 1) / CHUNK;
#pragma omp parallel for schedule(static)
    for (long ci = 0; ci < nchunks; ++ci) {
        long s = ci * CHUNK, len = (s + CHUNK <= n) ? CHUNK : n - s;
        k(p + s, g + s, m + s, v + s, len, c, lp ? lp + s : nullptr, lp_bf16);
    }
}

}  // namespace

extern "C" {

int dsb_cpu_adam_simd_level() {
    pick_kernel();
    return g_simd_level;
}

int dsb_cpu_adam_create(int id, float lr, float beta1, float beta2, float eps, float weight_decay, int adamw_mode) {
    std::lock_guard<std::mutex> lk(g_mu);
    AdamState s;
    s.lr = lr; s.beta1 = beta1; s.beta2 = beta2; s.eps = eps; s.wd = weight_decay; s.adamw = adamw_mode;
    g_opts[id] = s;
    return 0;
}

int dsb_cpu_adam_destroy(int id) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_opts.erase(id);
    return 0;
}

// One Adam step over n contiguous fp32 elements. dev_lp (optional) is a DEVICE pointer that receives the updated
// parameters in fp16 (lp_bf16 = 0) or bf16 (1) through the pinned double-buffered pipeline; host_lp (optional) is a
// HOST pointer receiving the same conversion without any CUDA call.
int dsb_cpu_adam_step(int id, long step, float lr, float beta1, float beta2, float eps, float weight_decay,
                      int bias_correction, float* p, const float* g, float* m, float* v, long n, void* dev_lp,
                      void* host_lp, int lp_bf16) {
    AdamState* st;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_opts.find(id);
        if (it == g_opts.end()) return -1;
        st = &it->second;
    }
    // beta powers in double, straight from the step counter: 1 - beta2^t cancels badly in float, and an incrementally
    // multiplied power would make a run resumed from a checkpoint differ from the uninterrupted one
    st->beta1_t = (float)pow((double)beta1, (double)step);
    st->beta2_t = (float)pow((double)beta2, (double)step);
    st->step = step; st->lr = lr; st->beta1 = beta1; st->beta2 = beta2; st->eps = eps; st->wd = weight_decay;
    Coef c;
    c.beta1 = beta1; c.beta2 = beta2; c.omb1 = 1.f - beta1; c.omb2 = 1.f - beta2; c.eps = eps; c.wd = weight_decay;
    c.adamw = st->adamw;
    float bc1 = 1.f, bc2 = 1.f;
    if (bias_correction) {
        bc1 = (float)(1.0 - pow((double)beta1, (double)step));
        bc2 = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)step)));
    }
    c.step_size = -lr / bc1;
    c.bc2_rsqrt = bc2;
    c.lr_wd = -lr * weight_decay;

    if (dev_lp == nullptr) {
        run_range(p, g, m, v, n, c, reinterpret_cast<uint16_t*>(host_lp), lp_bf16);
        return 0;
    }
    if (!cuda_init()) return -2;
    int buf = 0;
    for (long off = 0; off < n; off += TILE, buf ^= 1) {
        const long len = (off + TILE <= n) ? TILE : n - off;
        // the staging buffer about to be overwritten must have finished its previous copy
        g_cuda.streamSync(g_cuda.streams[buf]);
        uint16_t* stage = reinterpret_cast<uint16_t*>(g_cuda.staging[buf]);
        run_range(p + off, g + off, m + off, v + off, len, c, stage, lp_bf16);
        if (host_lp) memcpy(reinterpret_cast<uint16_t*>(host_lp) + off, stage, len * 2);
        if (g_cuda.memcpyAsync(reinterpret_cast<uint16_t*>(dev_lp) + off, stage, len * 2, 1 /*H2D*/,
                               g_cuda.streams[buf]) != 0)
            return -3;
    }
    g_cuda.streamSync(g_cuda.streams[0]);
    g_cuda.streamSync(g_cuda.streams[1]);
    return 0;
}

// fp32 -> fp16/bf16 host conversion (used by the offload path when the copy is issued from Python)
int dsb_cpu_cast_lp(const float* src, void* dst, long n, int lp_bf16) {
    uint16_t* d = reinterpret_cast<uint16_t*>(dst);
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i) d[i] = lp_bf16 ? f32_to_bf16(src[i]) : f32_to_f16(src[i]);
    return 0;
}

}  // extern "C"
