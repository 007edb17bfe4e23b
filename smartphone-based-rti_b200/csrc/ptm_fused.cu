// Fused PTM path: uint8 frames -> coefficients -> int32 light-grid table in ONE kernel
// (reference: compute_intensities + _interpolate_PTM + prepare_images_data,
// analysis.py:196-246, 263-317, 375-411).
//
// The fit is fp64-pipe bound and reads 1 byte per pixel.light; the table is HBM-write bound
// (4 bytes per pixel.direction) and needs almost no arithmetic.  Running them as two kernels
// serialises the two resources; here every CTA does both for its own 256 pixels, and because
// CTAs on an SM are in different phases the fit of one hides under the stores of the others.
// The coefficients go from registers to shared memory to the table phase without a trip
// through L2 (they are also written out, 48 B per pixel, because they are a result).
#include "common.cuh"

namespace rti {

constexpr int FU_THREADS = 256;         // = pixels per CTA
constexpr int FU_CAM_CHUNK = 256;       // frames of camera data resident in smem at a time
constexpr int FU_UNROLL = 4;

template <bool VEC4>
__global__ void __launch_bounds__(FU_THREADS, 2)
ptm_fused_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride, const double *__restrict__ cams,
                 int n, int row0, int cols, int64_t n_pixels, const double *__restrict__ axis, int g,
                 double *__restrict__ coef, int64_t plane_stride, int32_t *__restrict__ table,
                 int64_t image_stride, int32_t *__restrict__ bad) {
    extern __shared__ __align__(16) double s_dyn[];     // axis[g]
    __shared__ double s_cam[FU_CAM_CHUNK][3];
    __shared__ double s_coef[6][FU_THREADS];
    double *s_axis = s_dyn;

    const int t = threadIdx.x;
    const int64_t pbase = (int64_t)blockIdx.x * FU_THREADS;
    for (int i = t; i < g; i += FU_THREADS) s_axis[i] = axis[i];

    // ------------------------------------------------------------------ phase 1: fit, thread = pixel
    {
        const int64_t p = pbase + t;
        const bool active = p < n_pixels;
        const int64_t pc = active ? p : n_pixels - 1;
        const int yl = (int)(pc / cols);
        const double xd = (double)(int)(pc - (int64_t)yl * cols);
        const double yd = (double)(row0 + yl);
        const uint8_t *src = frames + pc;
        NormalEq ne;
        ne.clear();
        for (int f0 = 0; f0 < n; f0 += FU_CAM_CHUNK) {
            const int fc = min(FU_CAM_CHUNK, n - f0);
            __syncthreads();
            for (int i = t; i < fc; i += FU_THREADS) {
                const double cz = cams[(int64_t)(f0 + i) * 3 + 2];
                s_cam[i][0] = cams[(int64_t)(f0 + i) * 3 + 0];
                s_cam[i][1] = cams[(int64_t)(f0 + i) * 3 + 1];
                s_cam[i][2] = cz * cz;
            }
            __syncthreads();
            int i = 0;
            for (; i + FU_UNROLL <= fc; i += FU_UNROLL) {
                uint8_t v[FU_UNROLL];
#pragma unroll
                for (int j = 0; j < FU_UNROLL; ++j) v[j] = __ldg(src + (int64_t)(f0 + i + j) * frame_stride);
#pragma unroll
                for (int j = 0; j < FU_UNROLL; ++j) {
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
        double bound = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            s_coef[k][t] = a[k];
            bound += fabs(a[k]);
            if (active && coef != nullptr) coef[k * plane_stride + p] = a[k];
        }
        if (active && bad != nullptr) {
            if (!ok) atomicAdd(bad, 1);
            // |value| <= sum|a_k| on [-1,1]^2; NaN coefficients fail the comparison as well
            else if (!(bound < 2147483648.0)) atomicAdd(bad + 1, 1);
        }
    }
    __syncthreads();

    // ------------------------------------------------------------------ phase 2: table
    // warp w: pixel half (w / 4) of the CTA, lane = 4 consecutive pixels (512 B per warp store);
    // the four warps of a half interleave the grid rows v = (w % 4), +4, ...
    {
        const int lane = t & 31, warp = t >> 5;
        const int q0 = 128 * (warp >> 2) + 4 * lane;
        const int64_t p0 = pbase + q0;
        if (p0 >= n_pixels) return;
        const bool full = p0 + 4 <= n_pixels;
        double a[4][6];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int k = 0; k < 6; ++k) a[q][k] = s_coef[k][q0 + q];
        for (int v = warp & 3; v < g; v += 4) {
            const double lv = s_axis[v];
            double bv[4], cv[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                bv[q] = ptm_bv(a[q], lv);
                cv[q] = ptm_cv(a[q], lv);
            }
            int32_t *dst = table + ((int64_t)v * g) * image_stride + p0;
#pragma unroll 4
            for (int u = 0; u < g; ++u) {
                const double lu = s_axis[u];
                int val[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) val[q] = trunc_i32(ptm_eval(a[q][0], bv[q], cv[q], lu));
                if (VEC4 && full) {
                    __stcs(reinterpret_cast<int4 *>(dst), make_int4(val[0], val[1], val[2], val[3]));
                } else {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (p0 + q < n_pixels) __stcs(dst + q, val[q]);
                }
                dst += image_stride;
            }
        }
    }
}

}  // namespace rti

using namespace rti;

extern "C" int rti_ptm_fit_table_frames(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams, int n,
                                        int row0, int rows, int cols, const double *d_axis, int g, double *d_coef,
                                        int64_t plane_stride, int32_t *d_table, int64_t image_stride,
                                        int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_cams && d_axis && d_table, "rti_ptm_fit_table_frames: NULL pointer");
    RTI_REQUIRE(rows > 0 && cols > 0 && row0 >= 0, "rti_ptm_fit_table_frames: bad slab");
    RTI_REQUIRE(g > 0 && g <= 2048, "rti_ptm_fit_table_frames: bad grid size %d", g);
    if (n < 6) {
        set_error("rti_ptm_fit_table_frames: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    const int64_t n_pixels = (int64_t)rows * cols;
    RTI_REQUIRE(frame_stride >= n_pixels && image_stride >= n_pixels && (d_coef == nullptr || plane_stride >= n_pixels),
                "rti_ptm_fit_table_frames: stride < slab pixels");
    const int64_t blocks = (n_pixels + FU_THREADS - 1) / FU_THREADS;
    const size_t smem = (size_t)g * sizeof(double);
    const bool vec4 = (image_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_table) & 15) == 0);
    if (vec4)
        ptm_fused_kernel<true><<<(unsigned)blocks, FU_THREADS, smem, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_axis, g, d_coef, plane_stride, d_table,
            image_stride, d_bad);
    else
        ptm_fused_kernel<false><<<(unsigned)blocks, FU_THREADS, smem, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_axis, g, d_coef, plane_stride, d_table,
            image_stride, d_bad);
    RTI_LAUNCHED("ptm_fused_kernel");
    return RTI_OK;
}
