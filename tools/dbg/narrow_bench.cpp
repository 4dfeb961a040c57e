// host micro-benchmark: double -> float narrowing rate of the bench host's cores (what bounds pcb_decode_batch with double input)
#include <immintrin.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>
__attribute__((target("avx"))) static void k_avx(const double* s, float* d, size_t n) {
    for (size_t i = 0; i + 8 <= n; i += 8) {
        const __m128 lo = _mm256_cvtpd_ps(_mm256_loadu_pd(s + i)), hi = _mm256_cvtpd_ps(_mm256_loadu_pd(s + i + 4));
        _mm256_stream_ps(d + i, _mm256_insertf128_ps(_mm256_castps128_ps256(lo), hi, 1));
    }
    _mm_sfence();
}
__attribute__((target("avx"))) static void k_avx_pf(const double* s, float* d, size_t n) {
    for (size_t i = 0; i + 32 <= n; i += 32) {
        _mm_prefetch((const char*)(s + i + 256), _MM_HINT_NTA); _mm_prefetch((const char*)(s + i + 264), _MM_HINT_NTA);
        _mm_prefetch((const char*)(s + i + 272), _MM_HINT_NTA); _mm_prefetch((const char*)(s + i + 280), _MM_HINT_NTA);
        for (int j = 0; j < 32; j += 8) {
            const __m128 lo = _mm256_cvtpd_ps(_mm256_loadu_pd(s + i + j)), hi = _mm256_cvtpd_ps(_mm256_loadu_pd(s + i + j + 4));
            _mm256_stream_ps(d + i + j, _mm256_insertf128_ps(_mm256_castps128_ps256(lo), hi, 1));
        }
    }
    _mm_sfence();
}
__attribute__((target("avx512f,avx512dq"))) static void k_avx512(const double* s, float* d, size_t n) {
    for (size_t i = 0; i + 16 <= n; i += 16) {
        const __m256 lo = _mm512_cvtpd_ps(_mm512_loadu_pd(s + i)), hi = _mm512_cvtpd_ps(_mm512_loadu_pd(s + i + 8));
        _mm512_stream_ps(d + i, _mm512_insertf32x8(_mm512_castps256_ps512(lo), hi, 1));
    }
    _mm_sfence();
}
static void k_copy(const double* s, float* d, size_t n) { memcpy(d, s, n * 4); }
typedef void (*fn)(const double*, float*, size_t);
int main() {
    const size_t N = (size_t)1 << 28;   // 2 GiB of doubles
    double* src = (double*)aligned_alloc(64, N * 8); float* dst = (float*)aligned_alloc(64, N * 4);
    for (size_t i = 0; i < N; i++) src[i] = (double)(i & 1023) * 0.37 - 100.;
    memset(dst, 0, N * 4);
    printf("hardware threads %u, avx512f %d\n", std::thread::hardware_concurrency(), __builtin_cpu_supports("avx512f"));
    struct { const char* name; fn f; bool ok; } ks[] = {{"avx + stream", k_avx, true}, {"avx + stream + prefetchnta", k_avx_pf, true},
                                                        {"avx512 + stream", k_avx512, (bool)__builtin_cpu_supports("avx512f")}, {"memcpy of n*4 bytes", k_copy, true}};
    for (auto& k : ks) {
        if (!k.ok) continue;
        for (int T : {1, 2, 4, 8, 12, 14, 16, 24, 32}) {
            if (T > (int)std::thread::hardware_concurrency()) break;
            double best = 0;
            for (int rep = 0; rep < 3; rep++) {
                auto t0 = std::chrono::steady_clock::now();
                std::vector<std::thread> th;
                const size_t per = (N / T) & ~(size_t)63;
                for (int t = 0; t < T; t++) th.emplace_back([&, t] { k.f(src + t * per, dst + t * per, per); });
                for (auto& x : th) x.join();
                const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                best = std::max(best, (double)per * T / dt / 1e9);
            }
            printf("%-28s T=%2d  %6.2f Gvalues/s\n", k.name, T, best);
        }
    }
    return 0;
}
