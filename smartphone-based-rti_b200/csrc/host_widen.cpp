// Host half of rti_ptm_compute_host's narrow transfer (host_pipeline.cu): the int32 table tile
// (prepare_images_data's dtype, analysis.py:400-409) crosses PCIe as int16 when every value of
// the tile fits, and is widened back here into the caller's int32 table by a few host threads.
// Plain C++ (compiled by the host compiler, not cudafe) so the AVX2 path can use intrinsics;
// non-temporal stores: the destination (tens of GB) is written once and not read back here.
#include <immintrin.h>
#include "host_widen.h"

namespace rti {

__attribute__((target("avx2"))) static void widen_avx2(const int16_t *src, int32_t *dst, size_t n) {
    size_t i = 0;
    while (i < n && (reinterpret_cast<uintptr_t>(dst + i) & 31u)) {          // until dst is 32-byte aligned
        dst[i] = src[i];
        ++i;
    }
    for (; i + 32 <= n; i += 32) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 8));
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 16));
        const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i + 24));
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i), _mm256_cvtepi16_epi32(a));
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 8), _mm256_cvtepi16_epi32(b));
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 16), _mm256_cvtepi16_epi32(c));
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 24), _mm256_cvtepi16_epi32(d));
    }
    for (; i + 8 <= n; i += 8)
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i),
                            _mm256_cvtepi16_epi32(_mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i))));
    for (; i < n; ++i) dst[i] = src[i];
}

// dst[i] = src[i] for i < n (sign-extending); any alignment of either pointer.
void widen_i16_to_i32(const int16_t *src, int32_t *dst, size_t n) {
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) {
        widen_avx2(src, dst, n);
        return;
    }
    for (size_t i = 0; i < n; ++i) dst[i] = src[i];
}

// Orders the non-temporal stores of this thread before whatever it publishes next.
void widen_fence() { _mm_sfence(); }

}  // namespace rti
