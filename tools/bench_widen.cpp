// Host-only driver for the widening half of the narrow transfer (csrc/host_widen.h, host_widen.cpp):
//   widen_bench --check     correctness: widen_i16_to_i32 at every alignment / ragged length, and the Wideners
//                           thread protocol (tiles in order, skipped tiles untouched, early destruction) against
//                           a plain loop; prints "check ok" (tests/test_host_widen.py builds and runs this)
//   widen_bench T           throughput of T threads widening 2 GiB of int16 into 4 GiB of int32
//   g++ -O3 -std=c++17 -pthread tools/bench_widen.cpp smartphone-based-rti_b200/csrc/host_widen.cpp -o scratch/widen_bench
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <random>
#include <unistd.h>
#include "../smartphone-based-rti_b200/csrc/host_widen.h"

static int check() {
    // ---- the kernel of it: any alignment of either pointer, lengths around the vector widths
    for (int off = 0; off < 9; ++off)
        for (size_t n : {0, 1, 7, 8, 15, 16, 31, 32, 33, 63, 100, 1000, 4099}) {
            std::vector<int16_t> s(n + 16);
            std::vector<int32_t> d(n + 32, 0x7fffffff);
            for (size_t i = 0; i < s.size(); ++i) s[i] = (int16_t)((i + 1) * 2654435761u >> 7);
            if (n > 2) { s[off % 3] = -32768; s[off % 3 + 1] = 32767; }
            rti::widen_i16_to_i32(s.data() + off % 3, d.data() + off, n);
            for (size_t i = 0; i < n; ++i)
                if (d[off + i] != s[off % 3 + i]) { printf("MISMATCH off %d n %zu i %zu\n", off, n, i); return 1; }
            if (d[off + n] != 0x7fffffff || (off && d[off - 1] != 0x7fffffff)) { printf("OVERRUN off %d n %zu\n", off, n); return 1; }
        }
    // ---- the thread protocol: a slab of `slab` pixels per image, cut into ragged tiles; some tiles are
    // skipped (the caller re-copied them as int32) and must stay untouched; staging buffers are reused
    // three tiles later, which is only safe once wait_done() says so
    std::mt19937 rng(7);
    for (int trial = 0; trial < 24; ++trial) {
        const int threads = 1 + (int)(rng() % 7);
        const int64_t gg = 1 + (int64_t)(rng() % 37);                      // also fewer images than threads
        const int64_t n_tiles = 1 + (int64_t)(rng() % 9);
        std::vector<int64_t> tp((size_t)n_tiles), off((size_t)n_tiles);
        int64_t slab = 0;
        for (int64_t k = 0; k < n_tiles; ++k) { off[(size_t)k] = slab; tp[(size_t)k] = 1 + (int64_t)(rng() % 300); slab += tp[(size_t)k]; }
        std::vector<int32_t> table((size_t)(gg * slab), -7), ref((size_t)(gg * slab), -7);
        int64_t tp_max = 0;
        for (auto v : tp) tp_max = v > tp_max ? v : tp_max;
        std::vector<int32_t> stage((size_t)(3 * gg * tp_max));            // holds int16 or int32 tiles
        {
            rti::Wideners w(threads, n_tiles, slab, table.data());
            for (int64_t k = 0; k < n_tiles; ++k) {
                w.wait_done(k - 2);                                        // staging buffer k % 3 is free again
                int32_t *buf32 = stage.data() + (size_t)((k % 3) * gg * tp_max);
                int16_t *buf16 = reinterpret_cast<int16_t *>(buf32);
                const rti::Wideners::Kind kind = (rti::Wideners::Kind)(rng() % 3);      // skipped, int16 or int32
                // a tile may also cover only a range of the images (the chunks of rti_table_tile_to_host)
                const int64_t img0 = rng() % 2 ? 0 : (int64_t)(rng() % (uint64_t)gg);
                const int64_t n_img = img0 ? gg - img0 : gg;
                for (int64_t i = 0; i < n_img; ++i)
                    for (int64_t x = 0; x < tp[(size_t)k]; ++x) {
                        const int32_t v = kind == rti::Wideners::I32 ? (int32_t)rng() : (int32_t)(int16_t)(rng() & 0xffff);
                        if (kind == rti::Wideners::I32) buf32[i * tp[(size_t)k] + x] = v;
                        else buf16[i * tp[(size_t)k] + x] = (int16_t)v;
                        if (kind != rti::Wideners::SKIP) ref[(size_t)((img0 + i) * slab + off[(size_t)k] + x)] = v;
                    }
                w.publish({buf32, img0, n_img, off[(size_t)k], tp[(size_t)k], kind});
            }
            w.wait_done(n_tiles);
        }
        if (table != ref) { printf("WIDENERS MISMATCH trial %d\n", trial); return 1; }
    }
    // ---- the same protocol into a FILE: image i's payload sits at byte offsets[i] (gaps in between, as in the
    // result pickle); tiles bring int16 or int32 values for ranges of rows, skipped tiles leave holes untouched
    for (int trial = 0; trial < 8; ++trial) {
        char path[] = "/tmp/widen_check_XXXXXX";
        const int fd = mkstemp(path);
        if (fd < 0) { printf("mkstemp failed\n"); return 1; }
        const int threads = 1 + (int)(rng() % 5);
        const int64_t gg = 1 + (int64_t)(rng() % 23), n_tiles = 1 + (int64_t)(rng() % 6);
        std::vector<int64_t> tp((size_t)n_tiles), off((size_t)n_tiles), offsets((size_t)gg);
        int64_t slab = 0;
        for (int64_t k = 0; k < n_tiles; ++k) { off[(size_t)k] = slab; tp[(size_t)k] = 1 + (int64_t)(rng() % 200); slab += tp[(size_t)k]; }
        const int64_t gap = 3 + (int64_t)(rng() % 40), per = slab * 4 + gap, size = 11 + gg * per;
        for (int64_t i = 0; i < gg; ++i) offsets[(size_t)i] = 11 + i * per;
        std::vector<unsigned char> ref((size_t)size, 0xEE);
        if (pwrite(fd, ref.data(), ref.size(), 0) != (ssize_t)ref.size()) { printf("pwrite failed\n"); return 1; }
        int64_t tp_max = 0;
        for (auto v : tp) tp_max = v > tp_max ? v : tp_max;
        std::vector<int32_t> stage((size_t)(3 * gg * tp_max));
        {
            rti::Wideners w(threads, n_tiles, 0, nullptr, fd, offsets.data());
            for (int64_t k = 0; k < n_tiles; ++k) {
                w.wait_done(k - 2);
                int32_t *buf32 = stage.data() + (size_t)((k % 3) * gg * tp_max);
                int16_t *buf16 = reinterpret_cast<int16_t *>(buf32);
                const rti::Wideners::Kind kind = (rti::Wideners::Kind)(rng() % 3);
                for (int64_t i = 0; i < gg; ++i)
                    for (int64_t x = 0; x < tp[(size_t)k]; ++x) {
                        const int32_t v = kind == rti::Wideners::I32 ? (int32_t)rng() : (int32_t)(int16_t)(rng() & 0xffff);
                        if (kind == rti::Wideners::I32) buf32[i * tp[(size_t)k] + x] = v;
                        else buf16[i * tp[(size_t)k] + x] = (int16_t)v;
                        if (kind != rti::Wideners::SKIP) memcpy(&ref[(size_t)(offsets[(size_t)i] + (off[(size_t)k] + x) * 4)], &v, 4);
                    }
                w.publish({buf32, 0, gg, off[(size_t)k], tp[(size_t)k], kind});
            }
            w.wait_done(n_tiles);
            if (w.error()) { printf("FILE SINK errno %d\n", w.error()); return 1; }
        }
        std::vector<unsigned char> got((size_t)size);
        const bool ok = pread(fd, got.data(), got.size(), 0) == (ssize_t)got.size() && got == ref;
        close(fd);
        unlink(path);
        if (!ok) { printf("FILE SINK MISMATCH trial %d\n", trial); return 1; }
    }
    {   // a bad descriptor is reported, not ignored
        std::vector<int16_t> stage(64, 1);
        const int64_t offsets[1] = {0};
        rti::Wideners w(2, 1, 0, nullptr, 1 << 20, offsets);
        w.publish({stage.data(), 0, 1, 0, 64, rti::Wideners::I16});
        w.wait_done(1);
        if (w.error() == 0) { printf("FILE SINK: bad descriptor not reported\n"); return 1; }
    }
    // ---- early destruction: fewer tiles published than announced, and none at all; must join, not hang
    for (int published = 0; published < 3; ++published) {
        std::vector<int32_t> table(64 * 10, 0);
        std::vector<int16_t> stage(64 * 10, 5);
        rti::Wideners w(4, 5, 64, table.data());
        for (int k = 0; k < published; ++k) w.publish({stage.data(), 0, 10, 0, 64, rti::Wideners::I16});
        if (published) w.wait_done(published);
    }
    printf("check ok\n");
    return 0;
}

int main(int argc, char **argv) {
    if (argc > 1 && !strcmp(argv[1], "--check")) return check();
    if (argc > 3 && !strcmp(argv[1], "--file")) {
        // widen_bench --file T path [tiles]: T threads widen int16 tiles of 26 624 pixels x 10 000 images and pwrite
        // them into `path` laid out like the result pickle (what rti_table_tile_to_file does per tile)
        const int T = atoi(argv[2]);
        const int64_t gg = 10000, tp = 26624, n_tiles = argc > 4 ? atoi(argv[4]) : 4, per = n_tiles * tp * 4 + 165;
        const int fd = open(argv[3], O_RDWR | O_CREAT | O_TRUNC, 0644);
        if (fd < 0 || ftruncate(fd, 64 + gg * per) != 0) { printf("cannot create %s\n", argv[3]); return 1; }
        std::vector<int64_t> offsets((size_t)gg);
        for (int64_t i = 0; i < gg; ++i) offsets[(size_t)i] = 64 + i * per;
        std::vector<int16_t> stage((size_t)(gg * tp), 7);
        const auto t0 = std::chrono::steady_clock::now();
        {
            rti::Wideners w(T, n_tiles, 0, nullptr, fd, offsets.data());
            for (int64_t k = 0; k < n_tiles; ++k) w.publish({stage.data(), 0, gg, k * tp, tp, rti::Wideners::I16});
            w.wait_done(n_tiles);
            if (w.error()) { printf("errno %d\n", w.error()); return 1; }
        }
        const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("T=%d: %.2f GB written in %.3f s, %.2f GB/s\n", T, gg * n_tiles * tp * 4 / 1e9, dt, gg * n_tiles * tp * 4 / dt / 1e9);
        close(fd);
        return 0;
    }
    const int T = argc > 1 ? atoi(argv[1]) : 8;
    const size_t total = (size_t)1 << 30;     // elements: 2 GiB in, 4 GiB out
    int16_t *src = (int16_t *)aligned_alloc(64, total * 2);
    int32_t *dst = (int32_t *)aligned_alloc(64, total * 4);
    memset(src, 1, total * 2);
    memset(dst, 0, total * 4);
    for (int rep = 0; rep < 3; ++rep) {
        const auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> th;
        for (int j = 0; j < T; ++j)
            th.emplace_back([=] {
                const size_t a = total * j / T, b = total * (j + 1) / T;
                rti::widen_i16_to_i32(src + a, dst + a, b - a);
                rti::widen_fence();
            });
        for (auto &x : th) x.join();
        const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("T=%d: %.3f s, %.1f GB/s written\n", T, dt, total * 4 / dt / 1e9);
    }
    return 0;
}
