// gather.cpp -- host micro-benchmark of the pack step: 400 000 records of 1.5-3.9 KB scattered over the heap are
// gathered into one buffer by T threads.  mode 0: the library's stream_copy (SSE loads, non-temporal stores); 3: the same
// with software prefetch inside the record and of the next record; 4: memcpy; 5: AVX2 loads and non-temporal stores.
//     g++ -O2 -march=native -pthread gather.cpp -o gather && ./gather THREADS MODE
#include <immintrin.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <thread>
#include <chrono>
#include <algorithm>
#include <random>
static inline void copy_a(uint8_t* dst, const uint8_t* src, size_t n) {   // current
    if (n < 256) { memcpy(dst, src, n); return; }
    const size_t head = (16 - ((uintptr_t)dst & 15)) & 15;
    if (head) { memcpy(dst, src, head); dst += head; src += head; n -= head; }
    size_t i = 0;
    for (; i + 64 <= n; i += 64) {
        const __m128i a = _mm_loadu_si128((const __m128i*)(src + i));
        const __m128i b = _mm_loadu_si128((const __m128i*)(src + i + 16));
        const __m128i c = _mm_loadu_si128((const __m128i*)(src + i + 32));
        const __m128i d = _mm_loadu_si128((const __m128i*)(src + i + 48));
        _mm_stream_si128((__m128i*)(dst + i), a);
        _mm_stream_si128((__m128i*)(dst + i + 16), b);
        _mm_stream_si128((__m128i*)(dst + i + 32), c);
        _mm_stream_si128((__m128i*)(dst + i + 48), d);
    }
    if (i < n) memcpy(dst + i, src + i, n - i);
}
template <int PF>
static inline void copy_b(uint8_t* dst, const uint8_t* src, size_t n) {   // + prefetch inside the record
    if (n < 256) { memcpy(dst, src, n); return; }
    const size_t head = (16 - ((uintptr_t)dst & 15)) & 15;
    if (head) { memcpy(dst, src, head); dst += head; src += head; n -= head; }
    size_t i = 0;
    for (; i + 64 <= n; i += 64) {
        if (PF) _mm_prefetch((const char*)(src + i + PF), _MM_HINT_NTA);
        const __m128i a = _mm_loadu_si128((const __m128i*)(src + i));
        const __m128i b = _mm_loadu_si128((const __m128i*)(src + i + 16));
        const __m128i c = _mm_loadu_si128((const __m128i*)(src + i + 32));
        const __m128i d = _mm_loadu_si128((const __m128i*)(src + i + 48));
        _mm_stream_si128((__m128i*)(dst + i), a);
        _mm_stream_si128((__m128i*)(dst + i + 16), b);
        _mm_stream_si128((__m128i*)(dst + i + 32), c);
        _mm_stream_si128((__m128i*)(dst + i + 48), d);
    }
    if (i < n) memcpy(dst + i, src + i, n - i);
}
static inline void copy_avx2(uint8_t* dst, const uint8_t* src, size_t n) {
    if (n < 256) { memcpy(dst, src, n); return; }
    const size_t head = (32 - ((uintptr_t)dst & 31)) & 31;
    if (head) { memcpy(dst, src, head); dst += head; src += head; n -= head; }
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        _mm_prefetch((const char*)(src + i + 512), _MM_HINT_NTA);
        _mm_prefetch((const char*)(src + i + 576), _MM_HINT_NTA);
        const __m256i a = _mm256_loadu_si256((const __m256i*)(src + i));
        const __m256i b = _mm256_loadu_si256((const __m256i*)(src + i + 32));
        const __m256i c = _mm256_loadu_si256((const __m256i*)(src + i + 64));
        const __m256i d = _mm256_loadu_si256((const __m256i*)(src + i + 96));
        _mm256_stream_si256((__m256i*)(dst + i), a);
        _mm256_stream_si256((__m256i*)(dst + i + 32), b);
        _mm256_stream_si256((__m256i*)(dst + i + 64), c);
        _mm256_stream_si256((__m256i*)(dst + i + 96), d);
    }
    if (i < n) memcpy(dst + i, src + i, n - i);
}
int main(int argc, char** argv) {
    const int T = argc > 1 ? atoi(argv[1]) : 4;
    const int mode = argc > 2 ? atoi(argv[2]) : 0;
    const size_t N = 400000;                       // records
    std::mt19937 rng(1);
    std::vector<uint8_t*> src(N); std::vector<uint32_t> len(N);
    size_t total = 0;
    std::vector<uint8_t*> pool;
    for (size_t i = 0; i < N; ++i) { len[i] = 1500 + rng() % 2400; src[i] = (uint8_t*)malloc(len[i] + 33 + rng() % 64); memset(src[i], i, len[i]); total += len[i]; }
    std::shuffle(src.begin(), src.end(), rng);     // consumption order != allocation order
    uint8_t* dst = (uint8_t*)aligned_alloc(4096, total + 4096);
    memset(dst, 0, total);
    std::vector<size_t> start(N + 1); start[0] = 0; for (size_t i = 0; i < N; ++i) start[i + 1] = start[i] + len[i];
    for (int rep = 0; rep < 3; ++rep) {
        auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th;
        for (int t = 0; t < T; ++t) th.emplace_back([&, t]() {
            const size_t a = N * t / T, b = N * (t + 1) / T;
            for (size_t i = a; i < b; ++i) {
                if (mode >= 2 && i + 1 < b) { const uint8_t* nx = src[i + 1]; for (int k = 0; k < 512; k += 64) _mm_prefetch((const char*)(nx + k), _MM_HINT_NTA); }
                if (mode == 0) copy_a(dst + start[i], src[i], len[i]);
                else if (mode == 1) copy_b<512>(dst + start[i], src[i], len[i]);
                else if (mode == 2) copy_a(dst + start[i], src[i], len[i]);
                else if (mode == 3) copy_b<512>(dst + start[i], src[i], len[i]);
                else if (mode == 4) memcpy(dst + start[i], src[i], len[i]);
                else copy_avx2(dst + start[i], src[i], len[i]);
            }
            _mm_sfence();
        });
        for (auto& x : th) x.join();
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("T=%d mode=%d  %.2f GB/s total, %.2f per thread\n", T, mode, total / dt / 1e9, total / dt / 1e9 / T);
    }
}
