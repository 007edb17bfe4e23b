// Light-grid evaluation of the PTM coefficient map (reference: the evaluation loop
// of _interpolate_PTM, analysis.py:301-315, and prepare_images_data, :375-411) and
// per-direction relighting (interactive_relighting.py:25-38).
#include "common.cuh"

namespace rti {

// --------------------------------------------------------------------------------------
// Table kernel: table[v][u][p] = trunc(poly_p(axis[u], axis[v])), int32.
// Bound: HBM write, 4 bytes per pixel.direction (+ 48 B per pixel per v-chunk of
// coefficient reads, mostly L2 hits).  Each thread owns PPT consecutive pixels and
// walks TABLE_VC rows of the light grid; for every direction a warp writes one
// contiguous 32*PPT*4-byte run of that direction's image with a streaming store.
// blockIdx.x walks pixel chunks (fastest) so that co-resident CTAs write adjacent
// runs of the same TABLE_VC*g images.
// --------------------------------------------------------------------------------------
constexpr int TABLE_THREADS = 256;
constexpr int TABLE_VC = 4;             // grid rows (lv values) per CTA
constexpr int MAX_G = 2048;             // axis length limit (smem copy of the axis)

template <int PPT>
__global__ void __launch_bounds__(TABLE_THREADS)
ptm_table_kernel(const double *__restrict__ coef, int64_t plane_stride, int64_t n_pixels,
                 const double *__restrict__ axis, int g, int32_t *__restrict__ table,
                 int64_t image_stride, int32_t *__restrict__ bad) {
    extern __shared__ double s_axis[];
    for (int i = threadIdx.x; i < g; i += TABLE_THREADS) s_axis[i] = axis[i];
    __syncthreads();

    const int64_t p0 = ((int64_t)blockIdx.x * TABLE_THREADS + threadIdx.x) * PPT;
    if (p0 >= n_pixels) return;
    const bool full = p0 + PPT <= n_pixels;

    double a[PPT][6];
    bool finite_ok = true;
#pragma unroll
    for (int q = 0; q < PPT; ++q) {
        const int64_t p = (p0 + q < n_pixels) ? p0 + q : n_pixels - 1;
        double bound = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            a[q][k] = __ldg(coef + k * plane_stride + p);
            bound += fabs(a[q][k]);
        }
        // |value| <= sum|a_k| on [-1,1]^2; NaN/inf coefficients fail the comparison too
        if (!(bound < 2147483648.0) && p0 + q < n_pixels) finite_ok = false;
    }
    if (!finite_ok && bad != nullptr && blockIdx.y == 0) atomicAdd(bad, 1);

    const int v0 = blockIdx.y * TABLE_VC;
    const int v1 = min(v0 + TABLE_VC, g);
    for (int v = v0; v < v1; ++v) {
        const double lv = s_axis[v];
        double bv[PPT], cv[PPT];
#pragma unroll
        for (int q = 0; q < PPT; ++q) {
            bv[q] = ptm_bv(a[q], lv);
            cv[q] = ptm_cv(a[q], lv);
        }
        int32_t *dst = table + ((int64_t)v * g) * image_stride + p0;
#pragma unroll 4
        for (int u = 0; u < g; ++u) {
            const double lu = s_axis[u];
            int val[PPT];
#pragma unroll
            for (int q = 0; q < PPT; ++q) val[q] = trunc_i32(ptm_eval(a[q][0], bv[q], cv[q], lu));
            if (PPT == 4 && full) {
                __stcs(reinterpret_cast<int4 *>(dst), make_int4(val[0], val[1], val[2], val[3]));
            } else {
#pragma unroll
                for (int q = 0; q < PPT; ++q)
                    if (p0 + q < n_pixels) __stcs(dst + q, val[q]);
            }
            dst += image_stride;
        }
    }
}

// --------------------------------------------------------------------------------------
// fp64 grids, pixel-major [p][v][u] (interpolate_intensities layout): one CTA per
// pixel-group, threads stride the g*g directions so stores are coalesced.
// Bound: HBM write, 8 bytes per pixel.direction.
// --------------------------------------------------------------------------------------
constexpr int GRIDS_THREADS = 256;

__global__ void __launch_bounds__(GRIDS_THREADS)
ptm_grids_kernel(const double *__restrict__ coef, int64_t plane_stride, int64_t n_pixels,
                 const double *__restrict__ axis, int g, double *__restrict__ grids) {
    extern __shared__ double s_axis[];
    for (int i = threadIdx.x; i < g; i += GRIDS_THREADS) s_axis[i] = axis[i];
    __syncthreads();
    // direction d = v g + u walked by the thread in steps of the CTA size, as (v, u) with a carry:
    // a 64-bit d / g per table entry was 46 instructions per entry and made the kernel issue-bound
    const int gg = g * g;
    const int u0 = threadIdx.x % g, v0 = threadIdx.x / g, du = GRIDS_THREADS % g, dv = GRIDS_THREADS / g;
    for (int64_t p = blockIdx.x; p < n_pixels; p += gridDim.x) {
        double a[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) a[k] = __ldg(coef + k * plane_stride + p);
        double *dst = grids + p * gg;
        int u = u0, v = v0;
        for (int d = threadIdx.x; d < gg; d += GRIDS_THREADS) {
            const double lv = s_axis[v], lu = s_axis[u];
            __stcs(dst + d, ptm_eval(a[0], ptm_bv(a, lv), ptm_cv(a, lv), lu));
            u += du;
            v += dv;
            if (u >= g) { u -= g; ++v; }
        }
    }
}

// --------------------------------------------------------------------------------------
// Relighting: one light direction from the coefficient map.  Bound: HBM read
// (48 B/pixel fp64 map, 24 B/pixel fp32 map) + 1 or 6 bytes of image traffic.
// --------------------------------------------------------------------------------------
constexpr int RELIGHT_THREADS = 256;

__device__ __forceinline__ int clamp_u8(int v) { return min(max(v, 0), 255); }   // interactive_relighting.py:35-36

// HSV -> BGR of up to 4 consecutive pixels p0 .. p0 + cnt - 1 (p0 % 4 == 0) whose V channel is replaced
// by val[] (interactive_relighting.py:37-38).  The 12 bytes of a full quad are 4-byte aligned in both byte
// streams whenever their base pointers are: three word loads and three word stores instead of 8 + 12
// byte accesses; the row position (for cvtColor's row tail, see hsv_row_tail) is derived once per quad.
__device__ __forceinline__ void relight_bgr4(const uint8_t *__restrict__ hsv, uint8_t *__restrict__ out, int64_t p0,
                                             int cnt, const int (&val)[4], int row_width) {
    const bool words = cnt == 4 && (((reinterpret_cast<uintptr_t>(hsv) | reinterpret_cast<uintptr_t>(out)) & 3) == 0);
    uint32_t w[3] = {0u, 0u, 0u};
    if (words) {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(hsv + p0 * 3);
#pragma unroll
        for (int i = 0; i < 3; ++i) w[i] = __ldcs(src + i);
    } else {
#pragma unroll
        for (int i = 0; i < 12; ++i)
            if (i < 3 * cnt) w[i >> 2] |= (uint32_t)hsv[p0 * 3 + i] << (8 * (i & 3));
    }
    int x = 0, tail_from = 0;
    if (row_width > 0) {
        x = (int)(p0 % row_width);
        tail_from = row_width - (row_width & 31);
    }
    uint32_t o[3] = {0u, 0u, 0u};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int h8 = (w[(3 * q) >> 2] >> (8 * ((3 * q) & 3))) & 255;
        const int s8 = (w[(3 * q + 1) >> 2] >> (8 * ((3 * q + 1) & 3))) & 255;
        uint8_t c[3];
        hsv2bgr_u8(h8, s8, val[q], row_width > 0 && x >= tail_from, c[0], c[1], c[2]);
#pragma unroll
        for (int k = 0; k < 3; ++k) o[(3 * q + k) >> 2] |= (uint32_t)c[k] << (8 * ((3 * q + k) & 3));
        if (++x == row_width) x = 0;
    }
    if (words) {
        uint32_t *dst = reinterpret_cast<uint32_t *>(out + p0 * 3);
#pragma unroll
        for (int i = 0; i < 3; ++i) dst[i] = o[i];
    } else {
#pragma unroll
        for (int i = 0; i < 12; ++i)
            if (i < 3 * cnt) out[p0 * 3 + i] = (uint8_t)(o[i >> 2] >> (8 * (i & 3)));
    }
}

template <typename T, bool HSV>
__global__ void __launch_bounds__(RELIGHT_THREADS)
relight_kernel(const T *__restrict__ coef, int64_t plane_stride, int64_t n_pixels, int row_width, T lu, T lv,
               const uint8_t *__restrict__ hsv, uint8_t *__restrict__ out) {
    const int64_t p0 = ((int64_t)blockIdx.x * RELIGHT_THREADS + threadIdx.x) * 4;
    if (p0 >= n_pixels) return;
    const T uu = lu, vv = lv;
    int val[4];
    const bool vec = (p0 + 4 <= n_pixels) && ((plane_stride & 3) == 0) &&
                     ((reinterpret_cast<uintptr_t>(coef) & 31) == 0);
    T a[6][4];
    if (vec) {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            if (sizeof(T) == 8) {
                const double2 lo = __ldcs(reinterpret_cast<const double2 *>(coef + k * plane_stride + p0));
                const double2 hi = __ldcs(reinterpret_cast<const double2 *>(coef + k * plane_stride + p0 + 2));
                a[k][0] = (T)lo.x; a[k][1] = (T)lo.y; a[k][2] = (T)hi.x; a[k][3] = (T)hi.y;
            } else {
                const float4 w = __ldcs(reinterpret_cast<const float4 *>(coef + k * plane_stride + p0));
                a[k][0] = (T)w.x; a[k][1] = (T)w.y; a[k][2] = (T)w.z; a[k][3] = (T)w.w;
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < 6; ++k)
#pragma unroll
            for (int q = 0; q < 4; ++q)
                a[k][q] = coef[k * plane_stride + ((p0 + q < n_pixels) ? p0 + q : n_pixels - 1)];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        // same grouping as the table kernel so that the fp64 path reproduces the table
        // bit for bit at grid nodes
        const T bv = fma(a[2][q], vv, a[3][q]);
        const T cv = fma(fma(a[1][q], vv, a[4][q]), vv, a[5][q]);
        const T L = fma(fma(a[0][q], uu, bv), uu, cv);
        int iv;
        if (sizeof(T) == 8) iv = __double2int_rz((double)L);
        else iv = __float2int_rz((float)L);
        val[q] = clamp_u8(iv);
    }
    if (!HSV) {
        if (p0 + 4 <= n_pixels && (reinterpret_cast<uintptr_t>(out) & 3) == 0) {
            const uint32_t w = (uint32_t)val[0] | ((uint32_t)val[1] << 8) | ((uint32_t)val[2] << 16) |
                               ((uint32_t)val[3] << 24);
            *reinterpret_cast<uint32_t *>(out + p0) = w;
        } else {
            for (int q = 0; q < 4 && p0 + q < n_pixels; ++q) out[p0 + q] = (uint8_t)val[q];
        }
    } else {
        relight_bgr4(hsv, out, p0, (int)min((int64_t)4, n_pixels - p0), val, row_width);
    }
}

template <bool HSV>
__global__ void __launch_bounds__(RELIGHT_THREADS)
relight_image_kernel(const int32_t *__restrict__ image, int64_t n_pixels, int row_width,
                     const uint8_t *__restrict__ hsv, uint8_t *__restrict__ out) {
    const int64_t p0 = ((int64_t)blockIdx.x * RELIGHT_THREADS + threadIdx.x) * 4;
    if (p0 >= n_pixels) return;
    const int cnt = (int)min((int64_t)4, n_pixels - p0);
    int val[4] = {0, 0, 0, 0};
    if (cnt == 4 && (reinterpret_cast<uintptr_t>(image) & 15) == 0) {
        const int4 v = __ldcs(reinterpret_cast<const int4 *>(image + p0));
        val[0] = clamp_u8(v.x); val[1] = clamp_u8(v.y); val[2] = clamp_u8(v.z); val[3] = clamp_u8(v.w);
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (q < cnt) val[q] = clamp_u8(__ldcs(image + p0 + q));
    }
    if (!HSV) {
        if (cnt == 4 && (reinterpret_cast<uintptr_t>(out) & 3) == 0) {
            *reinterpret_cast<uint32_t *>(out + p0) = (uint32_t)val[0] | ((uint32_t)val[1] << 8) |
                                                      ((uint32_t)val[2] << 16) | ((uint32_t)val[3] << 24);
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q < cnt) out[p0 + q] = (uint8_t)val[q];
        }
    } else {
        relight_bgr4(hsv, out, p0, cnt, val, row_width);
    }
}

__global__ void coef_to_f32_kernel(const double *__restrict__ src, int64_t n, float *__restrict__ dst) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = (float)src[i];
}

// --------------------------------------------------------------------------------------
// prepare_images_data on arbitrary input (analysis.py:400-409): fp64 grids [p][d]
// (d = v*g + u) -> int32 table [d][p], truncating toward zero.  32x32 shared-memory
// tile transpose so that both the 8-byte loads and the 4-byte stores are coalesced.
// --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
grids_to_table_kernel(const double *__restrict__ grids, int64_t n_pixels, int64_t n_dirs,
                      int32_t *__restrict__ table, int64_t image_stride, int32_t *__restrict__ bad) {
    __shared__ int s_tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;        // 32 x 8
    const int64_t d0 = (int64_t)blockIdx.x * 32, p0 = (int64_t)blockIdx.y * 32;
    bool any_bad = false;
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        const int64_t p = p0 + ty + j, d = d0 + tx;
        if (p < n_pixels && d < n_dirs) {
            const double v = __ldcs(grids + p * n_dirs + d);
            if (!(fabs(v) < 2147483648.0)) any_bad = true;         // NaN, inf, out of int32
            s_tile[ty + j][tx] = trunc_i32(v);
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        const int64_t d = d0 + ty + j, p = p0 + tx;
        if (p < n_pixels && d < n_dirs) __stcs(table + d * image_stride + p, s_tile[tx][ty + j]);
    }
    if (any_bad && bad != nullptr) atomicAdd(bad, 1);
}

}  // namespace rti

using namespace rti;

extern "C" int rti_grids_to_table(const double *d_grids, int64_t n_pixels, int g, int32_t *d_table,
                                  int64_t image_stride, int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_grids && d_table, "rti_grids_to_table: NULL pointer");
    RTI_REQUIRE(n_pixels > 0 && g > 0 && image_stride >= n_pixels, "rti_grids_to_table: bad sizes");
    const int64_t n_dirs = (int64_t)g * g;
    const int64_t by = (n_pixels + 31) / 32;
    RTI_REQUIRE(by <= 65535, "rti_grids_to_table: more than 2M pixels per call; split the call");
    grids_to_table_kernel<<<dim3((unsigned)((n_dirs + 31) / 32), (unsigned)by), 256, 0, as_stream(stream)>>>(
        d_grids, n_pixels, n_dirs, d_table, image_stride, d_bad);
    RTI_LAUNCHED("grids_to_table_kernel");
    return RTI_OK;
}

extern "C" int rti_ptm_table(const double *d_coef, int64_t plane_stride, int64_t n_pixels,
                             const double *d_axis, int g, int32_t *d_table, int64_t image_stride,
                             int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_coef && d_axis && d_table, "rti_ptm_table: NULL pointer");
    RTI_REQUIRE(n_pixels > 0 && g > 0 && g <= MAX_G, "rti_ptm_table: bad sizes (pixels %lld, g %d)",
                (long long)n_pixels, g);
    RTI_REQUIRE(plane_stride >= n_pixels && image_stride >= n_pixels, "rti_ptm_table: stride < pixels");
    const size_t smem = (size_t)g * sizeof(double);
    const bool vec4 = (image_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_table) & 15) == 0);
    const unsigned gy = (unsigned)((g + TABLE_VC - 1) / TABLE_VC);
    if (vec4) {
        const int64_t bx = (n_pixels + TABLE_THREADS * 4 - 1) / (TABLE_THREADS * 4);
        ptm_table_kernel<4><<<dim3((unsigned)bx, gy), TABLE_THREADS, smem, as_stream(stream)>>>(
            d_coef, plane_stride, n_pixels, d_axis, g, d_table, image_stride, d_bad);
    } else {
        const int64_t bx = (n_pixels + TABLE_THREADS - 1) / TABLE_THREADS;
        ptm_table_kernel<1><<<dim3((unsigned)bx, gy), TABLE_THREADS, smem, as_stream(stream)>>>(
            d_coef, plane_stride, n_pixels, d_axis, g, d_table, image_stride, d_bad);
    }
    RTI_LAUNCHED("ptm_table_kernel");
    return RTI_OK;
}

extern "C" int rti_ptm_grids(const double *d_coef, int64_t plane_stride, int64_t n_pixels,
                             const double *d_axis, int g, double *d_grids, rti_stream_t stream) {
    RTI_REQUIRE(d_coef && d_axis && d_grids, "rti_ptm_grids: NULL pointer");
    RTI_REQUIRE(n_pixels > 0 && g > 0 && g <= MAX_G, "rti_ptm_grids: bad sizes");
    RTI_REQUIRE(plane_stride >= n_pixels, "rti_ptm_grids: stride < pixels");
    int64_t blocks = n_pixels < 148 * 16 ? n_pixels : 148 * 16;
    ptm_grids_kernel<<<(unsigned)blocks, GRIDS_THREADS, (size_t)g * sizeof(double), as_stream(stream)>>>(
        d_coef, plane_stride, n_pixels, d_axis, g, d_grids);
    RTI_LAUNCHED("ptm_grids_kernel");
    return RTI_OK;
}

extern "C" int rti_relight_ptm(const double *d_coef, int64_t plane_stride, int64_t n_pixels, int row_width,
                               double lu, double lv, const uint8_t *d_hsv, uint8_t *d_out, rti_stream_t stream) {
    RTI_REQUIRE(d_coef && d_out, "rti_relight_ptm: NULL pointer");
    RTI_REQUIRE(n_pixels > 0 && plane_stride >= n_pixels && row_width >= 0, "rti_relight_ptm: bad sizes");
    const int64_t blocks = (n_pixels + RELIGHT_THREADS * 4 - 1) / (RELIGHT_THREADS * 4);
    if (d_hsv)
        relight_kernel<double, true><<<(unsigned)blocks, RELIGHT_THREADS, 0, as_stream(stream)>>>(
            d_coef, plane_stride, n_pixels, row_width, lu, lv, d_hsv, d_out);
    else
        relight_kernel<double, false><<<(unsigned)blocks, RELIGHT_THREADS, 0, as_stream(stream)>>>(
            d_coef, plane_stride, n_pixels, row_width, lu, lv, nullptr, d_out);
    RTI_LAUNCHED("relight_kernel<double>");
    return RTI_OK;
}

extern "C" int rti_relight_ptm_f32(const float *d_coef, int64_t plane_stride, int64_t n_pixels, int row_width,
                                   float lu, float lv, const uint8_t *d_hsv, uint8_t *d_out, rti_stream_t stream) {
    RTI_REQUIRE(d_coef && d_out, "rti_relight_ptm_f32: NULL pointer");
    RTI_REQUIRE(n_pixels > 0 && plane_stride >= n_pixels && row_width >= 0, "rti_relight_ptm_f32: bad sizes");
    const int64_t blocks = (n_pixels + RELIGHT_THREADS * 4 - 1) / (RELIGHT_THREADS * 4);
    if (d_hsv)
        relight_kernel<float, true><<<(unsigned)blocks, RELIGHT_THREADS, 0, as_stream(stream)>>>(
            d_coef, plane_stride, n_pixels, row_width, lu, lv, d_hsv, d_out);
    else
        relight_kernel<float, false><<<(unsigned)blocks, RELIGHT_THREADS, 0, as_stream(stream)>>>(
            d_coef, plane_stride, n_pixels, row_width, lu, lv, nullptr, d_out);
    RTI_LAUNCHED("relight_kernel<float>");
    return RTI_OK;
}

extern "C" int rti_coef_to_f32(const double *d_coef, int64_t n_values, float *d_out, rti_stream_t stream) {
    RTI_REQUIRE(d_coef && d_out && n_values > 0, "rti_coef_to_f32: bad arguments");
    coef_to_f32_kernel<<<148 * 8, 256, 0, as_stream(stream)>>>(d_coef, n_values, d_out);
    RTI_LAUNCHED("coef_to_f32_kernel");
    return RTI_OK;
}

extern "C" int rti_relight_from_image(const int32_t *d_image, int64_t n_pixels, int row_width,
                                      const uint8_t *d_hsv, uint8_t *d_out, rti_stream_t stream) {
    RTI_REQUIRE(d_image && d_out && n_pixels > 0 && row_width >= 0, "rti_relight_from_image: bad arguments");
    const int64_t blocks = (n_pixels + RELIGHT_THREADS * 4 - 1) / (RELIGHT_THREADS * 4);
    if (d_hsv)
        relight_image_kernel<true><<<(unsigned)blocks, RELIGHT_THREADS, 0, as_stream(stream)>>>(
            d_image, n_pixels, row_width, d_hsv, d_out);
    else
        relight_image_kernel<false><<<(unsigned)blocks, RELIGHT_THREADS, 0, as_stream(stream)>>>(
            d_image, n_pixels, row_width, nullptr, d_out);
    RTI_LAUNCHED("relight_image_kernel");
    return RTI_OK;
}
