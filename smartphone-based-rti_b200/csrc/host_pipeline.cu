// Host-buffer pipeline for analysis.compute() steps 2-4 with PTM interpolation
// (analysis.py:466-472): frames on the host in, int32 table / coefficients on the
// host out.  The ROI slab is cut into row tiles; the H2D copy of tile t+1, the
// fit + table kernels of tile t and the D2H copy of tile t-1 run on three streams.
#include <mutex>
#include "common.cuh"

namespace rti {

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
        RTI_CUDA(cudaEventCreateWithFlags(&c.ev_d2h[b], cudaEventDisableTiming));
    }
    RTI_CUDA(cudaMalloc(&c.d_bad, 2 * sizeof(int32_t)));
    c.ready = true;
    return RTI_OK;
}

}  // namespace rti

using namespace rti;

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

    for (int b = 0; b < 2; ++b) {
        size_t fb = c.frames_bytes, cb = c.coef_bytes, tb = c.table_bytes;
        if ((rc = grow((void **)&c.d_frames[b], &fb, (size_t)n * tile_pixels)) != RTI_OK) return rc;
        if ((rc = grow((void **)&c.d_coef[b], &cb, (size_t)6 * tile_pixels * sizeof(double))) != RTI_OK) return rc;
        if (h_table && (rc = grow((void **)&c.d_table[b], &tb, (size_t)gg * tile_pixels * sizeof(int32_t))) != RTI_OK) return rc;
        if (b == 1) { c.frames_bytes = fb; c.coef_bytes = cb; if (h_table) c.table_bytes = tb; }
    }
    if ((rc = grow((void **)&c.d_cams, &c.cams_bytes, (size_t)n * 3 * sizeof(double))) != RTI_OK) return rc;
    if ((rc = grow((void **)&c.d_axis, &c.axis_bytes, (size_t)g * sizeof(double))) != RTI_OK) return rc;

    RTI_CUDA(cudaMemsetAsync(c.d_bad, 0, 2 * sizeof(int32_t), c.s_comp));
    RTI_CUDA(cudaMemcpyAsync(c.d_cams, h_cams, (size_t)n * 3 * sizeof(double), cudaMemcpyHostToDevice, c.s_comp));
    RTI_CUDA(cudaMemcpyAsync(c.d_axis, h_axis, (size_t)g * sizeof(double), cudaMemcpyHostToDevice, c.s_comp));

    int t = 0;
    for (int r = 0; r < rows; r += tile_rows, ++t) {
        const int b = t & 1;
        const int tr = (rows - r < tile_rows) ? rows - r : tile_rows;
        const int64_t tp = (int64_t)tr * cols;
        const int64_t pix_off = (int64_t)r * cols;                    // within the slab
        // ---- H2D: n rows of tp bytes, host pitch = one full frame
        if (t >= 2) RTI_CUDA(cudaStreamWaitEvent(c.s_h2d, c.ev_fit[b], 0));   // frames[b] free again
        RTI_CUDA(cudaMemcpy2DAsync(c.d_frames[b], (size_t)tile_pixels,
                                   h_frames + (int64_t)(row0 + r) * cols, (size_t)frame_pitch,
                                   (size_t)tp, (size_t)n, cudaMemcpyHostToDevice, c.s_h2d));
        RTI_CUDA(cudaEventRecord(c.ev_h2d[b], c.s_h2d));
        // ---- compute
        RTI_CUDA(cudaStreamWaitEvent(c.s_comp, c.ev_h2d[b], 0));
        if (t >= 2) RTI_CUDA(cudaStreamWaitEvent(c.s_comp, c.ev_d2h[b], 0));  // coef[b]/table[b] drained
        if (h_table) {
            rc = rti_ptm_fit_table_frames(c.d_frames[b], tile_pixels, c.d_cams, n, row0 + r, tr, cols, c.d_axis, g,
                                          c.d_coef[b], tile_pixels, c.d_table[b], tile_pixels, c.d_bad, c.s_comp);
        } else {
            rc = rti_ptm_fit_frames(c.d_frames[b], tile_pixels, c.d_cams, n, row0 + r, tr, cols, c.d_coef[b],
                                    tile_pixels, c.d_bad, c.s_comp);
        }
        if (rc != RTI_OK) return rc;
        RTI_CUDA(cudaEventRecord(c.ev_fit[b], c.s_comp));
        RTI_CUDA(cudaEventRecord(c.ev_comp[b], c.s_comp));
        // ---- D2H
        RTI_CUDA(cudaStreamWaitEvent(c.s_d2h, c.ev_comp[b], 0));
        if (h_table)
            RTI_CUDA(cudaMemcpy2DAsync(h_table + pix_off, (size_t)slab_pixels * sizeof(int32_t), c.d_table[b],
                                       (size_t)tile_pixels * sizeof(int32_t), (size_t)tp * sizeof(int32_t),
                                       (size_t)gg, cudaMemcpyDeviceToHost, c.s_d2h));
        if (h_coef)
            RTI_CUDA(cudaMemcpy2DAsync(h_coef + pix_off, (size_t)slab_pixels * sizeof(double), c.d_coef[b],
                                       (size_t)tile_pixels * sizeof(double), (size_t)tp * sizeof(double), 6,
                                       cudaMemcpyDeviceToHost, c.s_d2h));
        RTI_CUDA(cudaEventRecord(c.ev_d2h[b], c.s_d2h));
    }
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
