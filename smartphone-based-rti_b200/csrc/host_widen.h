// Host half of rti_ptm_compute_host's narrow transfer: plain C++ so that host_pipeline.cu,
// host_widen.cpp and the CPU-only test driver (tools/bench_widen.cpp) share one definition.
#pragma once
#include <atomic>
#include <condition_variable>
#include <errno.h>
#include <mutex>
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include <thread>
#include <unistd.h>
#include <vector>

namespace rti {

// dst[i] = src[i] for i < n (sign-extending), any alignment; AVX2 + non-temporal stores when the CPU has them
void widen_i16_to_i32(const int16_t *src, int32_t *dst, size_t n);
// orders the calling thread's non-temporal stores before whatever it publishes next
void widen_fence();

// Host threads that move staged tiles into the caller's int32 table, tile by tile in the order they are
// published.  A tile is `n_img` images of `tp` values each, packed [n_img][tp] as int16 (widened) or int32
// (copied); image i of the tile goes to h_table + (img0 + i) * slab_pixels + pix_off -- or, with a file
// sink, its int32 values are written with pwrite at byte offsets[img0 + i] + 4 pix_off of the file (the
// result pickle's image payloads, table_io.TableTileWriter).  Thread j owns images
// [n_img j / T, n_img (j + 1) / T) of every tile.  The destructor stops and joins the threads, so an early
// error return of the caller cannot leave any behind.
class Wideners {
public:
    enum Kind { SKIP = 0, I16 = 1, I32 = 2 };
    struct Job { const void *src; int64_t img0, n_img, pix_off, tp; Kind kind; };
    Wideners(int threads, int64_t n_tiles, int64_t slab_pixels, int32_t *h_table, int fd = -1,
             const int64_t *file_offsets = nullptr)
        : jobs_((size_t)n_tiles), left_((size_t)n_tiles, threads), n_tiles_(n_tiles),
          slab_pixels_(slab_pixels), h_table_(h_table), threads_(threads), fd_(fd), file_offsets_(file_offsets) {
        try {
            for (int j = 0; j < threads; ++j) th_.emplace_back([this, j] { run(j); });
        } catch (...) {
            stop();                                      // the threads that did start
            throw;
        }
    }
    ~Wideners() { stop(); }
    Wideners(const Wideners &) = delete;
    Wideners &operator=(const Wideners &) = delete;
    void publish(const Job &job) {                       // the next tile is in its staging buffer (or skipped)
        {
            std::lock_guard<std::mutex> lk(m_);
            jobs_[(size_t)published_] = job;
            ++published_;
        }
        cv_.notify_all();
    }
    void wait_done(int64_t tiles) {                      // blocks until the first `tiles` tiles are in h_table
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return done_ >= tiles; });
    }
    int64_t done() {                                     // tiles finished so far
        std::lock_guard<std::mutex> lk(m_);
        return done_;
    }
    int error() const { return errno_.load(); }          // file sink: errno of the first failed pwrite, else 0

private:
    void stop() {
        {
            std::lock_guard<std::mutex> lk(m_);
            abort_ = true;
        }
        cv_.notify_all();
        for (auto &t : th_) t.join();
        th_.clear();
    }
    bool write_all(const void *data, size_t bytes, int64_t at) {
        const char *p = static_cast<const char *>(data);
        while (bytes > 0) {
            const ssize_t w = pwrite(fd_, p, bytes, (off_t)at);
            if (w < 0) {
                if (errno == EINTR) continue;
                int expected = 0;
                errno_.compare_exchange_strong(expected, errno ? errno : EIO);
                return false;
            }
            p += w;
            at += w;
            bytes -= (size_t)w;
        }
        return true;
    }
    void run(int j) {
        std::vector<int32_t> wide;                       // file sink: one image's rows, widened
        for (int64_t k = 0; k < n_tiles_; ++k) {
            Job job;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return abort_ || published_ > k; });
                if (published_ <= k) return;             // stopped before this tile arrived
                job = jobs_[(size_t)k];
            }
            if (job.kind != SKIP) {
                const int64_t i0 = job.n_img * j / threads_, i1 = job.n_img * (j + 1) / threads_;
                if (fd_ >= 0 && job.kind == I16) wide.resize((size_t)job.tp);
                for (int64_t i = i0; i < i1; ++i) {
                    if (fd_ >= 0) {
                        if (errno_.load() != 0) break;
                        const int64_t at = file_offsets_[job.img0 + i] + job.pix_off * (int64_t)sizeof(int32_t);
                        const void *data = static_cast<const int32_t *>(job.src) + i * job.tp;
                        if (job.kind == I16) {           // through the cache: the page-cache copy reads it right away
                            const int16_t *s16 = static_cast<const int16_t *>(job.src) + i * job.tp;
                            for (int64_t x = 0; x < job.tp; ++x) wide[(size_t)x] = s16[x];
                            data = wide.data();
                        }
                        write_all(data, (size_t)job.tp * sizeof(int32_t), at);
                        continue;
                    }
                    int32_t *dst = h_table_ + (job.img0 + i) * slab_pixels_ + job.pix_off;
                    if (job.kind == I16)
                        widen_i16_to_i32(static_cast<const int16_t *>(job.src) + i * job.tp, dst, (size_t)job.tp);
                    else
                        memcpy(dst, static_cast<const int32_t *>(job.src) + i * job.tp, (size_t)job.tp * sizeof(int32_t));
                }
                widen_fence();
            }
            bool last;
            {
                std::lock_guard<std::mutex> lk(m_);
                last = --left_[(size_t)k] == 0;
                if (last) done_ = k + 1;                 // tiles finish in order: every thread walks them in order
            }
            if (last) cv_.notify_all();
        }
    }
    std::vector<std::thread> th_;
    std::vector<Job> jobs_;
    std::vector<int> left_;
    std::mutex m_;
    std::condition_variable cv_;
    int64_t published_ = 0, done_ = 0, n_tiles_, slab_pixels_;
    int32_t *h_table_;
    int threads_;
    int fd_;
    const int64_t *file_offsets_;
    std::atomic<int> errno_{0};
    bool abort_ = false;
};

}  // namespace rti
