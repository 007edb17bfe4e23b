#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <thread>
#include <chrono>
#include <cstring>
namespace rti { void widen_i16_to_i32(const int16_t*, int32_t*, size_t); void widen_fence(); }
int main(int argc, char** argv) {
    int T = argc > 1 ? atoi(argv[1]) : 8;
    // correctness at odd alignments and lengths
    for (int off = 0; off < 9; ++off) for (size_t n : {0, 1, 7, 8, 31, 32, 33, 100, 1000, 4099}) {
        std::vector<int16_t> s(n + 16); std::vector<int32_t> d(n + 32, 0x7fffffff);
        for (size_t i = 0; i < s.size(); ++i) s[i] = (int16_t)(i * 2654435761u >> 7);
        rti::widen_i16_to_i32(s.data() + off % 3, d.data() + off, n);
        for (size_t i = 0; i < n; ++i) if (d[off + i] != s[off % 3 + i]) { printf("MISMATCH off %d n %zu i %zu\n", off, n, i); return 1; }
        if (d[off + n] != 0x7fffffff || (off && d[off - 1] != 0x7fffffff)) { printf("OVERRUN\n"); return 1; }
    }
    printf("correct\n");
    size_t total = (size_t)1 << 30;     // elements: 2 GiB in, 4 GiB out
    int16_t* src = (int16_t*)aligned_alloc(64, total * 2); int32_t* dst = (int32_t*)aligned_alloc(64, total * 4);
    memset(src, 1, total * 2); memset(dst, 0, total * 4);
    for (int rep = 0; rep < 3; ++rep) {
        auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th;
        for (int j = 0; j < T; ++j) th.emplace_back([=] { size_t a = total * j / T, b = total * (j + 1) / T; rti::widen_i16_to_i32(src + a, dst + a, b - a); rti::widen_fence(); });
        for (auto& x : th) x.join();
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("T=%d: %.3f s, %.1f GB/s written\n", T, dt, total * 4 / dt / 1e9);
    }
}
