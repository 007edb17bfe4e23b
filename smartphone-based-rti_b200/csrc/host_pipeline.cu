// Host-buffer pipeline for analysis.compute() steps 2-4 with PTM interpolation
// (analysis.py:466-472): frames on the host in, int32 table / coefficients on the
// host out.  The ROI slab is cut into row tiles; the H2D copy of tile t+1, the
// fit + table kernels of tile t and the D2H copy of tile t-1 run on three streams.
//
// The int32 table is 4 B per pixel and light direction and PCIe is the end-to-end bound (42 GB per
// 1024^2 ROI at ~55 GB/s).  Narrow transfer (default, RTI_HOST_TRANSFER=wide switches it off): the
// tile is narrowed to int16 on the device (narrow_table_kernel, with a flag when a value does not
// fit), crosses PCIe at half the bytes into a pinned staging ring, and host threads widen it into
// the caller's table (host_widen.cpp) while the next tile is in flight; a tile whose flag is set is
// copied again as int32 straight from the device tile, so the result is the same table either way.
#include <memory>
#include <mutex>
#include <sched.h>
#include <stdexcept>
#include <stdlib.h>
#include "common.cuh"
#include "host_widen.h"

namespace rti {

constexpr int N_STAGE = 3;            // pinned int16 staging tiles: one filling by DMA, one being widened, one spare

struct HostPipe {
    int device = -1;
    cudaStream_t s_h2d = nullptr, s_comp = nullptr, s_d2h = nullptr;
    cudaEvent_t ev_h2d[2] = {}, ev_fit[2] = {}, ev_comp[2] = {}, ev_d2h[2] = {};
    uint8_t *d_frames[2] = {};
    double *d_coef[2] = {};
    int32_t *d_table[2] = {};
    double *d_cams = nullptr, *d_axis = nullptr;
    int32_t *d_bad = nullptr;           // [0] singular pixels, [1] unrepresentable values
    size_t frames_bytes = 0, coef_bytes = 0, table_bytes = 0, cams_bytes = 0, axis_bytes = 0;
    // narrow transfer
    int16_t *d_narrow[2] = {};
    int32_t *d_flag = nullptr;          // [2]: tile of buffer b holds a value outside int16
    int16_t *h_stage = nullptr;         // pinned, N_STAGE tiles
    int32_t *h_flag = nullptr;          // pinned [N_STAGE]
    cudaEvent_t ev_stage[N_STAGE] = {};
    int16_t *d_narrow_tile = nullptr;   // rti_table_tile_to_host: the whole tile narrowed
    size_t narrow_bytes = 0, narrow_tile_bytes = 0, stage_bytes = 0;
    bool ready = false;
};

// One workspace per device ordinal, each behind its own mutex: calls for different devices (one
// host thread per GPU) run concurrently, calls for the same device serialise on its workspace.
constexpr int MAX_DEVICES = 64;
static HostPipe g_pipes[MAX_DEVICES];
static std::mutex g_pipe_mutex[MAX_DEVICES];

static int grow(void **ptr, size_t *have, size_t need) {
    if (*have >= need && *ptr) return RTI_OK;
    if (*ptr) RTI_CUDA(cudaFree(*ptr));
    *ptr = nullptr;
    *have = 0;
    RTI_CUDA(cudaMalloc(ptr, need));
    *have = need;
    return RTI_OK;
}

static int pipe_release(HostPipe &c) {
    if (!c.ready) return RTI_OK;
    cudaSetDevice(c.device);
    for (int b = 0; b < 2; ++b) {
        cudaFree(c.d_frames[b]); cudaFree(c.d_coef[b]); cudaFree(c.d_table[b]);
        cudaEventDestroy(c.ev_h2d[b]); cudaEventDestroy(c.ev_fit[b]);
        cudaEventDestroy(c.ev_comp[b]); cudaEventDestroy(c.ev_d2h[b]);
    }
    cudaFree(c.d_cams); cudaFree(c.d_axis); cudaFree(c.d_bad);
    cudaFree(c.d_narrow[0]); cudaFree(c.d_narrow[1]); cudaFree(c.d_narrow_tile); cudaFree(c.d_flag);
    if (c.h_stage) cudaFreeHost(c.h_stage);
    if (c.h_flag) cudaFreeHost(c.h_flag);
    for (int s = 0; s < N_STAGE; ++s) cudaEventDestroy(c.ev_stage[s]);
    cudaStreamDestroy(c.s_h2d); cudaStreamDestroy(c.s_comp); cudaStreamDestroy(c.s_d2h);
    c = HostPipe();
    return RTI_OK;
}

static int pipe_init(HostPipe &c, int dev) {
    if (c.ready) return RTI_OK;
    c.device = dev;
    RTI_CUDA(cudaStreamCreateWithFlags(&c.s_h2d, cudaStreamNonBlocking));
    RTI_CUDA(cudaStreamCreateWithFlags(&c.s_comp, cudaStreamNonBlocking));
    RTI_CUDA(cudaStreamCreateWithFlags(&c.s_d2h, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        RTI_CUDA(cudaEventCreateWithFlags(&c.ev_h2d[b], cudaEventDisableTiming));
        RTI_CUDA(cudaEventCreateWithFlags(&c.ev_fit[b], cudaEventDisableTiming));
        RTI_CUDA(cudaEventCreateWithFlags(&c.ev_comp[b], cudaEventDisableTiming));
        RTI_CUDA(cudaEventCreateWithFlags(&c.ev_d2h[b], cudaEventDisableTiming | cudaEventBlockingSync));
    }
    // the host thread sleeps in cudaEventSynchronize on these instead of spinning: the cores belong to the wideners
    for (int s = 0; s < N_STAGE; ++s)
        RTI_CUDA(cudaEventCreateWithFlags(&c.ev_stage[s], cudaEventDisableTiming | cudaEventBlockingSync));
    RTI_CUDA(cudaMalloc(&c.d_bad, 2 * sizeof(int32_t)));
    RTI_CUDA(cudaMalloc(&c.d_flag, 2 * sizeof(int32_t)));
    RTI_CUDA(cudaHostAlloc(&c.h_flag, N_STAGE * sizeof(int32_t), cudaHostAllocDefault));
    c.ready = true;
    return RTI_OK;
}

// int32 table tile [gg][image_stride] -> int16 [gg][tp] (packed: one contiguous D2H copy); *flag |= 1 when a value
// is outside [-limit - 1, limit] (limit = 32767 unless a test lowers it).  Thread = 8 consecutive pixels of one
// image (two 16-byte loads, one 16-byte store) when the shapes allow it.
template <bool VEC>
__global__ void __launch_bounds__(256)
narrow_table_kernel(const int32_t *__restrict__ table, int64_t image_stride, int64_t tp, int64_t gg, int limit,
                    int16_t *__restrict__ out, int32_t *__restrict__ flag) {
    bool over = false;
    const int64_t x0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * (VEC ? 8 : 1);
    if (x0 < tp) {
        for (int64_t i = blockIdx.y; i < gg; i += gridDim.y) {
            const int32_t *src = table + i * image_stride + x0;
            int16_t *dst = out + i * tp + x0;
            if (VEC) {
                const int4 a = __ldcs(reinterpret_cast<const int4 *>(src));
                const int4 b = __ldcs(reinterpret_cast<const int4 *>(src) + 1);
                const int v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
                for (int k = 0; k < 8; ++k) over |= (v[k] > limit) | (v[k] < -limit - 1);
                int4 o;
                o.x = (v[0] & 0xffff) | (v[1] << 16);
                o.y = (v[2] & 0xffff) | (v[3] << 16);
                o.z = (v[4] & 0xffff) | (v[5] << 16);
                o.w = (v[6] & 0xffff) | (v[7] << 16);
                __stcs(reinterpret_cast<int4 *>(dst), o);
            } else {
                const int v = __ldcs(src);
                over |= (v > limit) | (v < -limit - 1);
                *dst = (int16_t)v;
            }
        }
    }
    if (__any_sync(0xffffffffu, over) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

static int launch_narrow(const int32_t *d_table, int64_t image_stride, int64_t tp, int64_t gg, int limit,
                         int16_t *d_out, int32_t *d_flag, cudaStream_t stream) {
    const bool vec = tp % 8 == 0 && image_stride % 4 == 0;
    const int64_t items = vec ? tp / 8 : tp;
    const unsigned bx = (unsigned)((items + 255) / 256);
    const unsigned by = (unsigned)(gg < 32768 ? gg : 32768);
    if (vec) narrow_table_kernel<true><<<dim3(bx, by), 256, 0, stream>>>(d_table, image_stride, tp, gg, limit, d_out, d_flag);
    else narrow_table_kernel<false><<<dim3(bx, by), 256, 0, stream>>>(d_table, image_stride, tp, gg, limit, d_out, d_flag);
    RTI_LAUNCHED("narrow_table_kernel");
    return RTI_OK;
}

// Cores this process may use for widening: RTI_HOST_THREADS, else the affinity mask shared out between the
// ranks of this host (torchrun's LOCAL_WORLD_SIZE), at most 16.
static int local_world() {
    const char *e = getenv("LOCAL_WORLD_SIZE");
    const int w = e ? atoi(e) : 1;
    return w > 0 ? w : 1;
}
static int widen_threads() {
    const char *e = getenv("RTI_HOST_THREADS");
    if (e && atoi(e) > 0) return atoi(e) < 64 ? atoi(e) : 64;
    cpu_set_t set;
    int cores = sched_getaffinity(0, sizeof(set), &set) == 0 ? CPU_COUNT(&set) : (int)std::thread::hardware_concurrency();
    cores /= local_world();
    return cores < 1 ? 1 : (cores > 16 ? 16 : cores);
}
// narrow transfer: RTI_HOST_TRANSFER=narrow|wide, else when this process has the host to itself and >= 8 cores.
// (With several ranks per host the shared limit is the host's memory path, which the int32 stores of the
// widening load as much as the DMA writes they replace: measured 91 GB/s of DMA for 8 ranks, ~105 GB/s of
// widening stores on 16 cores.)
// "mixed" (opt-in) adds one rule to "narrow": when the staging ring is full, i.e. the wideners are the slower side
// right now, the tile does not wait for them but goes as int32 straight into the caller's table (if that is
// pinned).  Measured on a quiet 16-core host it is SLOWER than plain narrow (533-551 ms against 497 ms per step at
// 1024^2 x 500, a quarter of the tiles as int32): while an int32 tile holds the link for twice as long, nothing
// new arrives for the wideners and they run dry.  Kept for hosts with few cores.
enum TransferMode { TRANSFER_WIDE, TRANSFER_NARROW, TRANSFER_MIXED };
static TransferMode transfer_mode() {
    const char *e = getenv("RTI_HOST_TRANSFER");
    if (e && e[0] == 'n') return TRANSFER_NARROW;
    if (e && e[0] == 'w') return TRANSFER_WIDE;
    if (e && e[0] == 'm') return TRANSFER_MIXED;
    return local_world() == 1 && widen_threads() >= 8 ? TRANSFER_NARROW : TRANSFER_WIDE;
}

static thread_local int64_t t_last_d2h_bytes = 0;

}  // namespace rti

using namespace rti;

extern "C" int64_t rti_host_d2h_bytes(void) { return t_last_d2h_bytes; }

extern "C" int rti_ptm_compute_host(const uint8_t *h_frames, const double *h_cams, int n, int rows_total,
                                    int cols, int row0, int rows, const double *h_axis, int g,
                                    int32_t *h_table, double *h_coef, int tile_rows) {
    RTI_REQUIRE(h_frames && h_cams && h_axis, "rti_ptm_compute_host: NULL input");
    RTI_REQUIRE(h_table || h_coef, "rti_ptm_compute_host: no output requested");
    RTI_REQUIRE(rows_total > 0 && cols > 0 && row0 >= 0 && rows > 0 && row0 + rows <= rows_total,
                "rti_ptm_compute_host: bad slab %d+%d of %d x %d", row0, rows, rows_total, cols);
    RTI_REQUIRE(g > 0 && g <= 2048, "rti_ptm_compute_host: bad grid size %d", g);
    if (n < 6) {
        set_error("rti_ptm_compute_host: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    int dev = 0;
    RTI_CUDA(cudaGetDevice(&dev));
    RTI_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "rti_ptm_compute_host: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_pipe_mutex[dev]);
    HostPipe &c = g_pipes[dev];
    int rc = pipe_init(c, dev);
    if (rc != RTI_OK) return rc;

    const int64_t gg = (int64_t)g * g;
    if (tile_rows <= 0) {
        // ~1 GiB of table per tile buffer (two buffers): big enough to amortise launch and
        // copy set-up, small enough that the three-stage overlap has many tiles to work on
        const int64_t target_pixels = h_table ? (int64_t)(1 << 30) / (gg * 4) : (1 << 20);
        tile_rows = (int)((target_pixels + cols - 1) / cols);
        if (tile_rows < 1) tile_rows = 1;
    }
    if (tile_rows > rows) tile_rows = rows;
    // keep tile pixel counts a multiple of 16: 16-byte table stores, and 16-byte aligned frame rows
    // on the device, which the warp-specialised fused kernel's cp.async staging needs
    while (tile_rows < rows && ((int64_t)tile_rows * cols) % 16 != 0) ++tile_rows;
    const int64_t tile_pixels = (int64_t)tile_rows * cols;
    const int64_t slab_pixels = (int64_t)rows * cols;
    const int64_t frame_pitch = (int64_t)rows_total * cols;
    const int64_t n_tiles = (rows + tile_rows - 1) / tile_rows;
    const TransferMode mode = h_table != nullptr ? transfer_mode() : TRANSFER_WIDE;
    const bool narrow = mode != TRANSFER_WIDE;
    bool mixed = false;
    if (mode == TRANSFER_MIXED) {                    // int32 tiles go straight into h_table: only worth it when pinned
        cudaPointerAttributes attr;
        mixed = cudaPointerGetAttributes(&attr, h_table) == cudaSuccess && attr.type == cudaMemoryTypeHost;
        cudaGetLastError();
    }
    const char *env_every = getenv("RTI_HOST_MIX_EVERY");                  // tests: every k-th tile goes as int32
    const int mix_every = mixed && env_every ? atoi(env_every) : 0;
    const char *env_limit = getenv("RTI_HOST_NARROW_LIMIT");              // tests lower it to force the int32 re-copy
    const int limit = env_limit && atoi(env_limit) > 0 && atoi(env_limit) < 32767 ? atoi(env_limit) : 32767;

    for (int b = 0; b < 2; ++b) {
        size_t fb = c.frames_bytes, cb = c.coef_bytes, tb = c.table_bytes, nb = c.narrow_bytes;
        if ((rc = grow((void **)&c.d_frames[b], &fb, (size_t)n * tile_pixels)) != RTI_OK) return rc;
        if ((rc = grow((void **)&c.d_coef[b], &cb, (size_t)6 * tile_pixels * sizeof(double))) != RTI_OK) return rc;
        if (h_table && (rc = grow((void **)&c.d_table[b], &tb, (size_t)gg * tile_pixels * sizeof(int32_t))) != RTI_OK) return rc;
        if (narrow && (rc = grow((void **)&c.d_narrow[b], &nb, (size_t)gg * tile_pixels * sizeof(int16_t))) != RTI_OK) return rc;
        if (b == 1) { c.frames_bytes = fb; c.coef_bytes = cb; if (h_table) c.table_bytes = tb; if (narrow) c.narrow_bytes = nb; }
    }
    if ((rc = grow((void **)&c.d_cams, &c.cams_bytes, (size_t)n * 3 * sizeof(double))) != RTI_OK) return rc;
    if ((rc = grow((void **)&c.d_axis, &c.axis_bytes, (size_t)g * sizeof(double))) != RTI_OK) return rc;
    const size_t stage_elems = (size_t)gg * tile_pixels;                   // int16 per staging tile
    if (narrow && c.stage_bytes < N_STAGE * stage_elems * sizeof(int16_t)) {
        if (c.h_stage) RTI_CUDA(cudaFreeHost(c.h_stage));
        c.h_stage = nullptr;
        c.stage_bytes = 0;
        RTI_CUDA(cudaHostAlloc(&c.h_stage, N_STAGE * stage_elems * sizeof(int16_t), cudaHostAllocDefault));
        c.stage_bytes = N_STAGE * stage_elems * sizeof(int16_t);
    }

    RTI_CUDA(cudaMemsetAsync(c.d_bad, 0, 2 * sizeof(int32_t), c.s_comp));
    RTI_CUDA(cudaMemcpyAsync(c.d_cams, h_cams, (size_t)n * 3 * sizeof(double), cudaMemcpyHostToDevice, c.s_comp));
    RTI_CUDA(cudaMemcpyAsync(c.d_axis, h_axis, (size_t)g * sizeof(double), cudaMemcpyHostToDevice, c.s_comp));

    int64_t d2h_bytes = 0;
    {
        // declared before the loop: its destructor joins the threads on every way out of this block
        std::unique_ptr<Wideners> wide;
        if (narrow) {
            try {
                wide.reset(new Wideners(widen_threads(), n_tiles, slab_pixels, h_table));
            } catch (const std::exception &e) {           // no threads to be had: say so instead of terminating
                set_error("rti_ptm_compute_host: cannot start the widening threads (%s); RTI_HOST_TRANSFER=wide avoids them",
                          e.what());
                return RTI_ERR_ARG;
            }
        }
        // a staged tile is handed to the wideners one iteration after it was queued, so the GPU always has the
        // next tile's copies queued behind the one the host waits for
        std::vector<int> slot_of((size_t)n_tiles, -1);       // staging slot of a tile; -1: it went as int32
        int64_t slot_last[N_STAGE];                          // the tile that used a slot last
        for (int s = 0; s < N_STAGE; ++s) slot_last[s] = -1;
        int next_slot = 0;
        auto hand_over = [&](int64_t k) -> int {
            const int b = (int)(k & 1), sb = slot_of[(size_t)k];
            const int r = (int)(k * tile_rows);
            const int tr = (rows - r < tile_rows) ? rows - r : tile_rows;
            const int64_t tp = (int64_t)tr * cols, pix_off = (int64_t)r * cols;
            if (sb < 0) {                                    // nothing to widen; keeps the tile count in step.  The wait
                RTI_CUDA(cudaEventSynchronize(c.ev_d2h[b])); // keeps this thread one tile ahead of the copies, no more,
                wide->publish({nullptr, 0, gg, pix_off, tp, Wideners::SKIP});   // so the next tile's choice sees the wideners
                return RTI_OK;
            }
            RTI_CUDA(cudaEventSynchronize(c.ev_stage[sb]));
            const bool over = c.h_flag[sb] != 0;
            if (over) {            // a value outside int16: this tile again, as int32 (its device tile is still intact)
                RTI_CUDA(cudaMemcpy2DAsync(h_table + pix_off, (size_t)slab_pixels * sizeof(int32_t), c.d_table[b],
                                           (size_t)tile_pixels * sizeof(int32_t), (size_t)tp * sizeof(int32_t),
                                           (size_t)gg, cudaMemcpyDeviceToHost, c.s_d2h));
                RTI_CUDA(cudaStreamSynchronize(c.s_d2h));
                d2h_bytes += gg * tp * (int64_t)sizeof(int32_t);
            }
            wide->publish({c.h_stage + (size_t)sb * stage_elems, 0, gg, pix_off, tp, over ? Wideners::SKIP : Wideners::I16});
            return RTI_OK;
        };
        for (int64_t t = 0; t < n_tiles; ++t) {
            const int b = (int)(t & 1), sb = next_slot;
            const int r = (int)(t * tile_rows);
            const int tr = (rows - r < tile_rows) ? rows - r : tile_rows;
            const int64_t tp = (int64_t)tr * cols;
            const int64_t pix_off = (int64_t)r * cols;                    // within the slab
            bool as_int16 = narrow;
            if (mixed) {
                if (mix_every > 0) as_int16 = t % mix_every != mix_every - 1;
                else as_int16 = slot_last[sb] < 0 || wide->done() > slot_last[sb];   // ring full: do not wait, go wide
            }
            // ---- H2D: n rows of tp bytes, host pitch = one full frame
            if (t >= 2) RTI_CUDA(cudaStreamWaitEvent(c.s_h2d, c.ev_fit[b], 0));   // frames[b] free again
            RTI_CUDA(cudaMemcpy2DAsync(c.d_frames[b], (size_t)tile_pixels,
                                       h_frames + (int64_t)(row0 + r) * cols, (size_t)frame_pitch,
                                       (size_t)tp, (size_t)n, cudaMemcpyHostToDevice, c.s_h2d));
            RTI_CUDA(cudaEventRecord(c.ev_h2d[b], c.s_h2d));
            // ---- compute
            RTI_CUDA(cudaStreamWaitEvent(c.s_comp, c.ev_h2d[b], 0));
            if (t >= 2) RTI_CUDA(cudaStreamWaitEvent(c.s_comp, c.ev_d2h[b], 0));  // coef[b]/table[b]/narrow[b] drained
            if (h_table) {
                rc = rti_ptm_fit_table_frames(c.d_frames[b], tile_pixels, c.d_cams, n, row0 + r, tr, cols, c.d_axis, g,
                                              c.d_coef[b], tile_pixels, c.d_table[b], tile_pixels, c.d_bad, c.s_comp);
            } else {
                rc = rti_ptm_fit_frames(c.d_frames[b], tile_pixels, c.d_cams, n, row0 + r, tr, cols, c.d_coef[b],
                                        tile_pixels, c.d_bad, c.s_comp);
            }
            if (rc != RTI_OK) return rc;
            RTI_CUDA(cudaEventRecord(c.ev_fit[b], c.s_comp));
            if (as_int16) {
                RTI_CUDA(cudaMemsetAsync(c.d_flag + b, 0, sizeof(int32_t), c.s_comp));
                if ((rc = launch_narrow(c.d_table[b], tile_pixels, tp, gg, limit, c.d_narrow[b], c.d_flag + b, c.s_comp)) != RTI_OK)
                    return rc;
            }
            RTI_CUDA(cudaEventRecord(c.ev_comp[b], c.s_comp));
            // ---- D2H
            RTI_CUDA(cudaStreamWaitEvent(c.s_d2h, c.ev_comp[b], 0));
            if (as_int16) {
                wide->wait_done(slot_last[sb] + 1);                       // staging tile sb has been widened
                slot_of[(size_t)t] = sb;
                slot_last[sb] = t;
                next_slot = (sb + 1) % N_STAGE;
                RTI_CUDA(cudaMemcpyAsync(c.h_stage + (size_t)sb * stage_elems, c.d_narrow[b],
                                         (size_t)gg * tp * sizeof(int16_t), cudaMemcpyDeviceToHost, c.s_d2h));
                RTI_CUDA(cudaMemcpyAsync(c.h_flag + sb, c.d_flag + b, sizeof(int32_t), cudaMemcpyDeviceToHost, c.s_d2h));
                d2h_bytes += gg * tp * (int64_t)sizeof(int16_t);
            } else if (h_table) {
                RTI_CUDA(cudaMemcpy2DAsync(h_table + pix_off, (size_t)slab_pixels * sizeof(int32_t), c.d_table[b],
                                           (size_t)tile_pixels * sizeof(int32_t), (size_t)tp * sizeof(int32_t),
                                           (size_t)gg, cudaMemcpyDeviceToHost, c.s_d2h));
                d2h_bytes += gg * tp * (int64_t)sizeof(int32_t);
            }
            if (h_coef) {
                RTI_CUDA(cudaMemcpy2DAsync(h_coef + pix_off, (size_t)slab_pixels * sizeof(double), c.d_coef[b],
                                           (size_t)tile_pixels * sizeof(double), (size_t)tp * sizeof(double), 6,
                                           cudaMemcpyDeviceToHost, c.s_d2h));
                d2h_bytes += 6 * tp * (int64_t)sizeof(double);
            }
            RTI_CUDA(cudaEventRecord(c.ev_d2h[b], c.s_d2h));
            if (as_int16) RTI_CUDA(cudaEventRecord(c.ev_stage[sb], c.s_d2h));
            if (narrow && t >= 1 && (rc = hand_over(t - 1)) != RTI_OK) return rc;
        }
        if (narrow) {
            if ((rc = hand_over(n_tiles - 1)) != RTI_OK) return rc;
            wide->wait_done(n_tiles);
        }
    }
    t_last_d2h_bytes = d2h_bytes;
    int32_t bad[2] = {0, 0};
    RTI_CUDA(cudaStreamSynchronize(c.s_d2h));
    RTI_CUDA(cudaMemcpyAsync(bad, c.d_bad, sizeof(bad), cudaMemcpyDeviceToHost, c.s_comp));
    RTI_CUDA(cudaStreamSynchronize(c.s_comp));
    RTI_CUDA(cudaStreamSynchronize(c.s_h2d));
    if (bad[0] > 0) {
        set_error("rti_ptm_compute_host: %d pixel(s) have a singular PTM system", bad[0]);
        return RTI_ERR_SINGULAR;
    }
    if (bad[1] > 0) {
        set_error("rti_ptm_compute_host: table values not representable as int32 (NaN/inf/overflow)");
        return RTI_ERR_NONFINITE;
    }
    return RTI_OK;
}


// One device table tile -> the caller's host table through the same staging ring and host threads, in chunks
// of whole images: the D2H copy of chunk c + 1 runs while chunk c is widened (int16 crossing) or copied
// (int32 crossing: wide mode, or a value of the tile is outside int16).
// Destination: rows of a host table (h_table) or of the image payloads of a result file (fd + per-image offsets).
static int tile_to_sink(const int32_t *d_tile, int64_t image_stride, int64_t n_images, int64_t tile_pixels,
                        int32_t *h_table, int64_t host_image_stride, int fd, const int64_t *file_offsets,
                        int64_t pix_off, void *ready_event, rti_stream_t stream) {
    int dev = 0;
    RTI_CUDA(cudaGetDevice(&dev));
    RTI_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "rti_table_tile_to_host/file: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_pipe_mutex[dev]);
    HostPipe &c = g_pipes[dev];
    int rc = pipe_init(c, dev);
    if (rc != RTI_OK) return rc;
    // order behind the producer of the tile: the caller's event, or everything queued on its stream so far
    if (ready_event != nullptr) {
        RTI_CUDA(cudaStreamWaitEvent(c.s_d2h, static_cast<cudaEvent_t>(ready_event), 0));
    } else {
        RTI_CUDA(cudaEventRecord(c.ev_comp[0], as_stream(stream)));
        RTI_CUDA(cudaStreamWaitEvent(c.s_d2h, c.ev_comp[0], 0));
    }
    bool narrow = transfer_mode() != TRANSFER_WIDE;
    const char *env_limit = getenv("RTI_HOST_NARROW_LIMIT");
    const int limit = env_limit && atoi(env_limit) > 0 && atoi(env_limit) < 32767 ? atoi(env_limit) : 32767;
    if (narrow) {
        if ((rc = grow((void **)&c.d_narrow_tile, &c.narrow_tile_bytes, (size_t)n_images * tile_pixels * sizeof(int16_t))) != RTI_OK)
            return rc;
        RTI_CUDA(cudaMemsetAsync(c.d_flag, 0, sizeof(int32_t), c.s_d2h));
        if ((rc = launch_narrow(d_tile, image_stride, tile_pixels, n_images, limit, c.d_narrow_tile, c.d_flag, c.s_d2h)) != RTI_OK)
            return rc;
        RTI_CUDA(cudaMemcpyAsync(c.h_flag, c.d_flag, sizeof(int32_t), cudaMemcpyDeviceToHost, c.s_d2h));
        RTI_CUDA(cudaStreamSynchronize(c.s_d2h));
        if (c.h_flag[0] != 0) narrow = false;                 // a value outside int16: this tile crosses as int32
    }
    const size_t esize = narrow ? sizeof(int16_t) : sizeof(int32_t);
    const size_t chunk_bytes = (size_t)64 << 20;
    int64_t per = (int64_t)(chunk_bytes / ((size_t)tile_pixels * esize));
    per = per < 1 ? 1 : (per > n_images ? n_images : per);
    const size_t slot_bytes = (size_t)per * tile_pixels * esize;
    if (c.stage_bytes < N_STAGE * slot_bytes) {
        if (c.h_stage) RTI_CUDA(cudaFreeHost(c.h_stage));
        c.h_stage = nullptr;
        c.stage_bytes = 0;
        const size_t want = N_STAGE * (slot_bytes > chunk_bytes ? slot_bytes : chunk_bytes);
        RTI_CUDA(cudaHostAlloc(&c.h_stage, want, cudaHostAllocDefault));
        c.stage_bytes = want;
    }
    const int64_t n_chunks = (n_images + per - 1) / per;
    int64_t d2h_bytes = 0;
    {
        std::unique_ptr<Wideners> wide;
        try {
            wide.reset(new Wideners(widen_threads(), n_chunks, host_image_stride, h_table, fd, file_offsets));
        } catch (const std::exception &e) {
            set_error("rti_table_tile_to_host: cannot start the host threads (%s)", e.what());
            return RTI_ERR_ARG;
        }
        auto slot = [&](int64_t k) { return reinterpret_cast<unsigned char *>(c.h_stage) + (size_t)(k % N_STAGE) * slot_bytes; };
        auto hand_over = [&](int64_t k) -> int {
            const int64_t i0 = k * per, cnt = (n_images - i0 < per) ? n_images - i0 : per;
            RTI_CUDA(cudaEventSynchronize(c.ev_stage[k % N_STAGE]));
            wide->publish({slot(k), i0, cnt, pix_off, tile_pixels, narrow ? Wideners::I16 : Wideners::I32});
            return RTI_OK;
        };
        for (int64_t k = 0; k < n_chunks; ++k) {
            const int64_t i0 = k * per, cnt = (n_images - i0 < per) ? n_images - i0 : per;
            wide->wait_done(k - N_STAGE + 1);                 // the slot's previous chunk is in the table
            if (narrow)
                RTI_CUDA(cudaMemcpyAsync(slot(k), c.d_narrow_tile + i0 * tile_pixels, (size_t)cnt * tile_pixels * esize,
                                         cudaMemcpyDeviceToHost, c.s_d2h));
            else
                RTI_CUDA(cudaMemcpy2DAsync(slot(k), (size_t)tile_pixels * esize, d_tile + i0 * image_stride,
                                           (size_t)image_stride * esize, (size_t)tile_pixels * esize, (size_t)cnt,
                                           cudaMemcpyDeviceToHost, c.s_d2h));
            d2h_bytes += cnt * tile_pixels * (int64_t)esize;
            RTI_CUDA(cudaEventRecord(c.ev_stage[k % N_STAGE], c.s_d2h));
            if (k >= 1 && (rc = hand_over(k - 1)) != RTI_OK) return rc;
        }
        if ((rc = hand_over(n_chunks - 1)) != RTI_OK) return rc;
        wide->wait_done(n_chunks);
        if (wide->error() != 0) {
            set_error("rti_table_tile_to_file: pwrite failed (errno %d)", wide->error());
            return RTI_ERR_ARG;
        }
    }
    t_last_d2h_bytes = d2h_bytes;
    return RTI_OK;
}

extern "C" int rti_table_tile_to_host(const int32_t *d_tile, int64_t image_stride, int64_t n_images, int64_t tile_pixels,
                                      int32_t *h_table, int64_t host_image_stride, int64_t pix_off, void *ready_event,
                                      rti_stream_t stream) {
    RTI_REQUIRE(d_tile && h_table, "rti_table_tile_to_host: NULL pointer");
    RTI_REQUIRE(n_images > 0 && tile_pixels > 0 && image_stride >= tile_pixels && pix_off >= 0 &&
                host_image_stride >= pix_off + tile_pixels, "rti_table_tile_to_host: bad shape");
    return tile_to_sink(d_tile, image_stride, n_images, tile_pixels, h_table, host_image_stride, -1, nullptr, pix_off,
                        ready_event, stream);
}

// The same into a file: image i's int32 values go to byte h_offsets[i] + 4 pix_off of the open descriptor `fd`
// (pwrite from the host threads; the file's image payloads are laid out by table_io.TableTileWriter).
extern "C" int rti_table_tile_to_file(const int32_t *d_tile, int64_t image_stride, int64_t n_images, int64_t tile_pixels,
                                      int fd, const int64_t *h_offsets, int64_t pix_off, void *ready_event,
                                      rti_stream_t stream) {
    RTI_REQUIRE(d_tile && h_offsets && fd >= 0, "rti_table_tile_to_file: NULL pointer or bad descriptor");
    RTI_REQUIRE(n_images > 0 && tile_pixels > 0 && image_stride >= tile_pixels && pix_off >= 0,
                "rti_table_tile_to_file: bad shape");
    return tile_to_sink(d_tile, image_stride, n_images, tile_pixels, nullptr, 0, fd, h_offsets, pix_off, ready_event,
                        stream);
}

extern "C" int rti_destroy(void) {
    release_tile_counters();
    int cur = 0;
    const bool have_cur = cudaGetDevice(&cur) == cudaSuccess;
    for (int d = 0; d < MAX_DEVICES; ++d) {
        std::lock_guard<std::mutex> lock(g_pipe_mutex[d]);
        pipe_release(g_pipes[d]);
    }
    if (have_cur) cudaSetDevice(cur);
    return RTI_OK;
}
