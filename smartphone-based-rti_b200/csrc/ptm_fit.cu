// PTM fit kernels: per-pixel least squares of the 6 biquadratic coefficients
// (reference: _interpolate_PTM, analysis.py:282-298).
#include "common.cuh"

namespace rti {

// --------------------------------------------------------------------------------------
// Fused gather + fit straight from the uint8 frames (no SoA materialisation).
// One thread per pixel; consecutive threads own consecutive pixels so each frame is
// read as one fully used 32-byte sector per warp.  The camera table (cx, cy, cz^2 per
// frame) is staged in shared memory in chunks.  Bound: fp64 pipe (~50 fp64 issue slots
// per pixel.light against 1 byte of HBM traffic).
// --------------------------------------------------------------------------------------
constexpr int FIT_THREADS = 128;
constexpr int FIT_CAM_CHUNK = 512;      // frames of camera data resident in smem at a time
constexpr int FIT_UNROLL = 4;           // independent byte loads in flight per thread

__global__ void __launch_bounds__(FIT_THREADS)
ptm_fit_frames_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride,
                      const double *__restrict__ cams, int n, int row0, int cols,
                      int64_t n_pixels, double *__restrict__ coef, int64_t plane_stride,
                      int32_t *__restrict__ bad) {
    __shared__ double s_cam[FIT_CAM_CHUNK][3];
    const int64_t p = (int64_t)blockIdx.x * FIT_THREADS + threadIdx.x;
    const bool active = p < n_pixels;
    const int64_t pc = active ? p : n_pixels - 1;           // clamp: tail threads redo the last pixel
    const int yl = (int)(pc / cols);
    const double xd = (double)(int)(pc - (int64_t)yl * cols);
    const double yd = (double)(row0 + yl);
    const uint8_t *src = frames + pc;

    NormalEq ne;
    ne.clear();
    for (int f0 = 0; f0 < n; f0 += FIT_CAM_CHUNK) {
        const int fc = min(FIT_CAM_CHUNK, n - f0);
        __syncthreads();
        for (int i = threadIdx.x; i < fc; i += FIT_THREADS) {
            const double cz = cams[(int64_t)(f0 + i) * 3 + 2];
            s_cam[i][0] = cams[(int64_t)(f0 + i) * 3 + 0];
            s_cam[i][1] = cams[(int64_t)(f0 + i) * 3 + 1];
            s_cam[i][2] = cz * cz;
        }
        __syncthreads();
        int i = 0;
        for (; i + FIT_UNROLL <= fc; i += FIT_UNROLL) {
            uint8_t v[FIT_UNROLL];
#pragma unroll
            for (int j = 0; j < FIT_UNROLL; ++j) v[j] = __ldg(src + (int64_t)(f0 + i + j) * frame_stride);
#pragma unroll
            for (int j = 0; j < FIT_UNROLL; ++j) {
                float lu, lv;
                light_vector(s_cam[i + j][0], s_cam[i + j][1], s_cam[i + j][2], xd, yd, lu, lv);
                ne.add(lu, lv, u8_to_double(v[j]));
            }
        }
        for (; i < fc; ++i) {
            const uint8_t v = __ldg(src + (int64_t)(f0 + i) * frame_stride);
            float lu, lv;
            light_vector(s_cam[i][0], s_cam[i][1], s_cam[i][2], xd, yd, lu, lv);
            ne.add(lu, lv, u8_to_double(v));
        }
    }
    double a[6];
    const bool ok = solve_normal_eq(ne, (double)n, a);
    if (active) {
#pragma unroll
        for (int k = 0; k < 6; ++k) coef[k * plane_stride + p] = a[k];
        if (!ok && bad != nullptr) atomicAdd(bad, 1);
    }
}

// --------------------------------------------------------------------------------------
// Fit from materialised (lx, ly, intensity) arrays [pixel][n], light index fastest
// (the compute_intensities layout, analysis.py:217-219).  One warp per pixel: lanes
// stride the light axis so every load is a coalesced 128-byte line, the 26 partial
// sums are combined with warp shuffles and lane 0 does the in-register Cholesky.
// Bound: HBM (12 bytes per pixel.light).
// --------------------------------------------------------------------------------------
constexpr int SOA_WARPS = 8;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(SOA_WARPS * 32)
ptm_fit_soa_kernel(const float *__restrict__ lx, const float *__restrict__ ly,
                   const int32_t *__restrict__ inten, int64_t n_pixels, int n,
                   double *__restrict__ coef, int64_t plane_stride, int32_t *__restrict__ bad) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * SOA_WARPS + (threadIdx.x >> 5);
    const int64_t n_warps = (int64_t)gridDim.x * SOA_WARPS;
    for (int64_t p = warp; p < n_pixels; p += n_warps) {
        const float *px = lx + p * n;
        const float *py = ly + p * n;
        const int32_t *pi = inten + p * n;
        NormalEq ne;
        ne.clear();
        int i = lane;
        for (; i + 96 < n; i += 128) {                     // 4 independent lines in flight
            float u[4], v[4];
            int w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                u[j] = __ldcs(px + i + 32 * j);
                v[j] = __ldcs(py + i + 32 * j);
                w[j] = __ldcs(pi + i + 32 * j);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) ne.add(u[j], v[j], (double)w[j]);
        }
        for (; i < n; i += 32) ne.add(__ldcs(px + i), __ldcs(py + i), (double)__ldcs(pi + i));
#pragma unroll
        for (int k = 0; k < 15; ++k) ne.g[k] = warp_sum(ne.g[k]);
#pragma unroll
        for (int k = 0; k < 5; ++k) ne.s[k] = warp_sum(ne.s[k]);
#pragma unroll
        for (int k = 0; k < 6; ++k) ne.r[k] = warp_sum(ne.r[k]);
        if (lane == 0) {
            double a[6];
            const bool ok = solve_normal_eq(ne, (double)n, a);
#pragma unroll
            for (int k = 0; k < 6; ++k) coef[k * plane_stride + p] = a[k];
            if (!ok && bad != nullptr) atomicAdd(bad, 1);
        }
    }
}

// --------------------------------------------------------------------------------------
// Shared-light fast path: coef[6][P] = pinv[6][n] . I[n][P] (fp64 accumulate).
// Each thread owns 4 consecutive pixels (one 32-bit load per frame), the pseudo-inverse
// columns are broadcast from shared memory.
// --------------------------------------------------------------------------------------
constexpr int SH_THREADS = 128;
constexpr int SH_CHUNK = 256;

__global__ void __launch_bounds__(SH_THREADS)
ptm_fit_shared_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride,
                      const double *__restrict__ pinv, int n, int64_t n_pixels,
                      double *__restrict__ coef, int64_t plane_stride) {
    __shared__ double s_p[SH_CHUNK][6];
    const int64_t p0 = ((int64_t)blockIdx.x * SH_THREADS + threadIdx.x) * 4;
    const bool vec = (p0 + 3 < n_pixels) && ((frame_stride & 3) == 0) &&
                     ((reinterpret_cast<uintptr_t>(frames) & 3) == 0);
    double acc[4][6];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int k = 0; k < 6; ++k) acc[q][k] = 0.0;
    for (int f0 = 0; f0 < n; f0 += SH_CHUNK) {
        const int fc = min(SH_CHUNK, n - f0);
        __syncthreads();
        for (int i = threadIdx.x; i < fc * 6; i += SH_THREADS)
            s_p[i / 6][i % 6] = pinv[(int64_t)(i % 6) * n + f0 + i / 6];
        __syncthreads();
        if (p0 >= n_pixels) continue;
        for (int i = 0; i < fc; ++i) {
            const uint8_t *src = frames + (int64_t)(f0 + i) * frame_stride + p0;
            uint32_t w;
            if (vec) {
                w = __ldg(reinterpret_cast<const uint32_t *>(src));
            } else {
                w = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (p0 + q < n_pixels) w |= (uint32_t)__ldg(src + q) << (8 * q);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double v = (double)((w >> (8 * q)) & 0xffu);
#pragma unroll
                for (int k = 0; k < 6; ++k) acc[q][k] = fma(s_p[i][k], v, acc[q][k]);
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
        if (p0 + q < n_pixels)
#pragma unroll
            for (int k = 0; k < 6; ++k) coef[k * plane_stride + p0 + q] = acc[q][k];
}

}  // namespace rti

using namespace rti;

extern "C" int rti_ptm_fit_frames(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams,
                                  int n, int row0, int rows, int cols, double *d_coef,
                                  int64_t plane_stride, int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_cams && d_coef, "rti_ptm_fit_frames: NULL pointer");
    RTI_REQUIRE(rows > 0 && cols > 0 && row0 >= 0, "rti_ptm_fit_frames: bad slab %d+%d x %d", row0, rows, cols);
    if (n < 6) {
        set_error("rti_ptm_fit_frames: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    const int64_t n_pixels = (int64_t)rows * cols;
    RTI_REQUIRE(plane_stride >= n_pixels, "rti_ptm_fit_frames: plane_stride < pixels");
    RTI_REQUIRE(frame_stride >= n_pixels, "rti_ptm_fit_frames: frame_stride < slab pixels");
    const int64_t blocks = (n_pixels + FIT_THREADS - 1) / FIT_THREADS;
    ptm_fit_frames_kernel<<<(unsigned)blocks, FIT_THREADS, 0, as_stream(stream)>>>(
        d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_coef, plane_stride, d_bad);
    RTI_LAUNCHED("ptm_fit_frames_kernel");
    return RTI_OK;
}

extern "C" int rti_ptm_fit_soa(const float *d_lx, const float *d_ly, const int32_t *d_inten,
                               int64_t n_pixels, int n, double *d_coef, int64_t plane_stride,
                               int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_inten && d_coef, "rti_ptm_fit_soa: NULL pointer");
    RTI_REQUIRE(n_pixels > 0, "rti_ptm_fit_soa: no pixels");
    if (n < 6) {
        set_error("rti_ptm_fit_soa: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    RTI_REQUIRE(plane_stride >= n_pixels, "rti_ptm_fit_soa: plane_stride < pixels");
    int64_t blocks = (n_pixels + SOA_WARPS - 1) / SOA_WARPS;
    const int64_t cap = 148LL * 8 * 16;                    // persistent-ish: 16 waves of 8 CTAs/SM
    if (blocks > cap) blocks = cap;
    ptm_fit_soa_kernel<<<(unsigned)blocks, SOA_WARPS * 32, 0, as_stream(stream)>>>(
        d_lx, d_ly, d_inten, n_pixels, n, d_coef, plane_stride, d_bad);
    RTI_LAUNCHED("ptm_fit_soa_kernel");
    return RTI_OK;
}

extern "C" int rti_ptm_fit_shared(const uint8_t *d_frames, int64_t frame_stride, const double *d_pinv,
                                  int n, int64_t n_pixels, double *d_coef, int64_t plane_stride,
                                  rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_pinv && d_coef, "rti_ptm_fit_shared: NULL pointer");
    RTI_REQUIRE(n > 0 && n_pixels > 0, "rti_ptm_fit_shared: empty input");
    RTI_REQUIRE(plane_stride >= n_pixels && frame_stride >= n_pixels, "rti_ptm_fit_shared: stride < pixels");
    const int64_t blocks = (n_pixels + SH_THREADS * 4 - 1) / (SH_THREADS * 4);
    ptm_fit_shared_kernel<<<(unsigned)blocks, SH_THREADS, 0, as_stream(stream)>>>(
        d_frames, frame_stride, d_pinv, n, n_pixels, d_coef, plane_stride);
    RTI_LAUNCHED("ptm_fit_shared_kernel");
    return RTI_OK;
}
