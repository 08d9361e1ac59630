// host-side micro-benchmark (tools, not product): how fast can ONE thread answer "does this RGBA8 view contain 0x00FF00FF"?
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <vector>
#include <immintrin.h>
static bool plain(const uint32_t* p, size_t n) { uint32_t any = 0; for (size_t i = 0; i < n; ++i) any |= (uint32_t)(p[i] == 0x00FF00FFu); return any != 0; }
template <int S>
static bool streams(const uint32_t* p, size_t n) {   // S interleaved linear streams: more hardware-prefetcher streams / misses in flight
    const size_t c = n / S; uint32_t any = 0;
    for (size_t i = 0; i < c; ++i) {
#pragma GCC unroll 8
        for (int s = 0; s < S; ++s) any |= (uint32_t)(p[s * c + i] == 0x00FF00FFu);
    }
    for (size_t i = c * S; i < n; ++i) any |= (uint32_t)(p[i] == 0x00FF00FFu);
    return any != 0;
}
template <int S>
static bool streams_blk(const uint32_t* p, size_t n) {   // S streams, 64-byte blocks (one cache line per stream per iteration), SSE2
    const size_t c = (n / S) & ~(size_t)15; __m128i acc = _mm_setzero_si128(); const __m128i k = _mm_set1_epi32(0x00FF00FF);
    for (size_t i = 0; i < c; i += 16) {
#pragma GCC unroll 8
        for (int s = 0; s < S; ++s) {
            const __m128i* q = reinterpret_cast<const __m128i*>(p + s * c + i);
            acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q), k)); acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 1), k));
            acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 2), k)); acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 3), k));
        }
    }
    uint32_t any = _mm_movemask_epi8(acc) != 0;
    for (size_t i = c * S; i < n; ++i) any |= (uint32_t)(p[i] == 0x00FF00FFu);
    return any != 0;
}
static bool prefetching(const uint32_t* p, size_t n) {
    __m128i acc = _mm_setzero_si128(); const __m128i k = _mm_set1_epi32(0x00FF00FF); const size_t c = n & ~(size_t)15;
    for (size_t i = 0; i < c; i += 16) {
        _mm_prefetch(reinterpret_cast<const char*>(p + i + 1024), _MM_HINT_NTA);
        const __m128i* q = reinterpret_cast<const __m128i*>(p + i);
        acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q), k)); acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 1), k));
        acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 2), k)); acc = _mm_or_si128(acc, _mm_cmpeq_epi32(_mm_loadu_si128(q + 3), k));
    }
    uint32_t any = _mm_movemask_epi8(acc) != 0;
    for (size_t i = c; i < n; ++i) any |= (uint32_t)(p[i] == 0x00FF00FFu);
    return any != 0;
}
int main() {
    const size_t n = 800 * 800; const int V = 100;
    std::vector<std::vector<uint32_t>> views(V, std::vector<uint32_t>(n));
    for (auto& v : views) for (size_t i = 0; i < n; ++i) v[i] = (uint32_t)rand() | 0xFF000000u;
    struct { const char* name; bool (*f)(const uint32_t*, size_t); } fs[] = {{"plain", plain}, {"2 streams", streams<2>}, {"4 streams", streams<4>}, {"8 streams", streams<8>},
        {"4 streams, line blocks", streams_blk<4>}, {"8 streams, line blocks", streams_blk<8>}, {"16 streams, line blocks", streams_blk<16>}, {"prefetch 4 KB ahead", prefetching}};
    for (auto& f : fs) {
        bool any = false; double best = 1e30;
        for (int r = 0; r < 4; ++r) {
            auto t0 = std::chrono::steady_clock::now();
            for (int v = 0; v < V; ++v) any |= f.f(views[v].data(), n);
            best = std::min(best, std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / V);
        }
        printf("%-28s %.1f us per view (%d)\n", f.name, best, (int)any);
    }
}
