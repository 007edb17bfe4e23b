// Shared helpers for the sm_100a kernels of librti_b200.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/rti_b200.h"

namespace rti {

// ---- error plumbing ----------------------------------------------------------------
void set_error(const char *fmt, ...);
int check_cuda(cudaError_t e, const char *what);
void count_launch(int n = 1);
void release_tile_counters();      // ptm_fused.cu: frees the dynamic-scheduler counters (rti_destroy)

#define RTI_REQUIRE(cond, ...)                 \
    do {                                       \
        if (!(cond)) {                         \
            rti::set_error(__VA_ARGS__);       \
            return RTI_ERR_ARG;                \
        }                                      \
    } while (0)

#define RTI_CUDA(call)                                         \
    do {                                                       \
        int _rc = rti::check_cuda((call), #call);              \
        if (_rc != RTI_OK) return _rc;                         \
    } while (0)

// check the launch itself (configuration errors), not completion
#define RTI_LAUNCHED(name)                                     \
    do {                                                       \
        rti::count_launch();                                   \
        int _rc = rti::check_cuda(cudaGetLastError(), name);   \
        if (_rc != RTI_OK) return _rc;                         \
    } while (0)

static inline cudaStream_t as_stream(rti_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- device math --------------------------------------------------------------------

// 1/sqrt(d2) in fp64 for finite d2 > 0: MUFU.RSQ64H seed (rsqrt.approx.f64, relative error
// 2^-22.9, no float conversions) + one third-order step r(1 + e/2 + 3e^2/8), e = 1 - d2 r^2.
// Truncation error 5e^3/16 < 2^-70, i.e. the result carries only fp64 rounding (~2 ulp).  The
// light vector is rounded to float32 afterwards (analysis.py:230-231), so this matches the
// reference's sqrt + divide except when the fp64 value sits within ~1e-16 relative of a
// float32 rounding boundary.
__device__ __forceinline__ double rsqrt_pos(double d2) {
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d2));
    const double t = d2 * r;
    const double e = fma(-t, r, 1.0);
    const double q = fma(0.375, e, 0.5);
    return fma(r * e, q, r);
}

// uint8 -> double without the conversion pipe: 2^52 + v has v in its low mantissa bits.
__device__ __forceinline__ double u8_to_double(unsigned v) {
    return __hiloint2double(0x43300000, (int)v) - 4503599627370496.0;
}

// Unit light vector of pixel (xd, yd, 0) for camera (cx, cy, cz), cz2 = cz*cz:
// l = (cam - p)/||cam - p|| in fp64, (l[0], l[1]) rounded to float32
// (analysis.py:228-231).
__device__ __forceinline__ void light_vector(double cx, double cy, double cz2, double xd,
                                             double yd, float &lx, float &ly) {
    const double dx = cx - xd;
    const double dy = cy - yd;
    const double d2 = fma(dx, dx, fma(dy, dy, cz2));
    const double r = rsqrt_pos(d2);
    lx = (float)(dx * r);
    ly = (float)(dy * r);
}

// Note on the float32 roundings of the fit (two f64->f32 and five f32->f64 conversions per light
// sample, all XU-pipe instructions at 16 lanes/clk/SM): doing them on the integer pipe instead
// (round-to-nearest-even on the fp64 bit pattern, 7 ALU instructions each) was measured and is
// SLOWER on sm_100 (fit 2.05 -> 2.22 ms at 1M pixels x 500 lights): the ALU pipe has the same
// 16 lanes/clk/SMSP as the fp64 pipe and the instruction count doubles.  The conversions stay.
// Likewise the table's F2I: replacing it by an add.rz.f64 magic-number truncation + 3 integer
// instructions was neutral for the table kernel and slower inside the fused kernels.

// The 26 running sums of one pixel's normal equations for the design row
// m = (u2, v2, uv, u, v, 1) (analysis.py:285) and right-hand side I:
//   g[0..14]  sum m_a*m_b for a <= b < 5 (row-major upper triangle of the 5x5 block)
//   s[0..4]   sum m_a            (the column against the constant term)
//   r[0..4]   sum m_a * I,  r[5] = sum I;   the (5,5) entry is the sample count.
struct NormalEq {
    double g[15];
    double s[5];
    double r[6];
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int i = 0; i < 15; ++i) g[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 5; ++i) s[i] = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) r[i] = 0.0;
    }
    // lu, lv: float32 light components; the three products are formed in float32
    // exactly as numpy does for float32 scalars before the row is promoted to fp64.
    __device__ __forceinline__ void add(float lu, float lv, double inten) {
        double m[5];
        m[0] = (double)__fmul_rn(lu, lu);
        m[1] = (double)__fmul_rn(lv, lv);
        m[2] = (double)__fmul_rn(lu, lv);
        m[3] = (double)lu;
        m[4] = (double)lv;
        accumulate(m, inten);
    }
    __device__ __forceinline__ void accumulate(const double m[5], double inten) {
        int k = 0;
#pragma unroll
        for (int a = 0; a < 5; ++a) {
#pragma unroll
            for (int b = a; b < 5; ++b) {
                g[k] = fma(m[a], m[b], g[k]);
                ++k;
            }
            s[a] += m[a];
            r[a] = fma(m[a], inten, r[a]);
        }
        r[5] += inten;
    }
};

// Solve the 6x6 normal equations by Cholesky (fp64, fully unrolled, in registers).
// Returns false when a pivot is not positive (rank-deficient pixel) or has lost more than 12
// digits against its diagonal entry: pivot_j / G_jj <= 1e-12 means cond(G) >~ 1e12, i.e. a design
// matrix with cond >~ 1e6 (light directions nearly on a line / a conic), where the normal
// equations can no longer deliver north_star's 1e-4 against the reference's SVD
// (analysis.py:295-298).  Such pixels are reported as singular instead of returned silently off.
constexpr double PTM_PIVOT_RTOL = 1e-12;
__device__ __forceinline__ bool solve_normal_eq(const NormalEq &ne, double count, double a[6]) {
    double L[6][6];
    int k = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
#pragma unroll
        for (int j = i; j < 5; ++j) L[j][i] = ne.g[k++];     // lower triangle (symmetric)
        L[5][i] = ne.s[i];
    }
    L[5][5] = count;
    bool ok = true;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        const double d0 = L[j][j];
        double d = d0;
#pragma unroll
        for (int p = 0; p < j; ++p) d = fma(-L[j][p], L[j][p], d);
        ok = ok && (d > PTM_PIVOT_RTOL * d0);
        const double inv = rsqrt(d);
        L[j][j] = d * inv;                                   // sqrt(d)
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            double v = L[i][j];
#pragma unroll
            for (int p = 0; p < j; ++p) v = fma(-L[i][p], L[j][p], v);
            L[i][j] = v * inv;
        }
    }
    // forward: L y = r ; backward: L^T a = y
    double y[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double v = ne.r[i];
#pragma unroll
        for (int p = 0; p < i; ++p) v = fma(-L[i][p], y[p], v);
        y[i] = v / L[i][i];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double v = y[i];
#pragma unroll
        for (int p = i + 1; p < 6; ++p) v = fma(-L[p][i], a[p], v);
        a[i] = v / L[i][i];
    }
    if (!ok) {                                               // flagged pixels carry NaN, never a plausible-looking value
#pragma unroll
        for (int i = 0; i < 6; ++i) a[i] = __longlong_as_double(0x7FF8000000000000LL);
    }
    return ok;
}

// a0*lu^2 + a1*lv^2 + a2*lu*lv + a3*lu + a4*lv + a5 (analysis.py:307-312) regrouped
// as (a0*lu + (a3 + a2*lv))*lu + ((a1*lv + a4)*lv + a5): callers hoist the two
// lv-only terms out of the lu loop.
__device__ __forceinline__ double ptm_bv(const double a[6], double lv) { return fma(a[2], lv, a[3]); }
__device__ __forceinline__ double ptm_cv(const double a[6], double lv) {
    return fma(fma(a[1], lv, a[4]), lv, a[5]);
}
__device__ __forceinline__ double ptm_eval(double a0, double bv, double cv, double lu) {
    return fma(fma(a0, lu, bv), lu, cv);
}

// float64 -> int32 like numpy's element assignment: truncate toward zero.
__device__ __forceinline__ int trunc_i32(double v) { return __double2int_rz(v); }

// OpenCV 8-bit COLOR_HSV2BGR (see oracle/rti_oracle.py hsv2bgr_u8 for the derivation).  cvtColor
// works row by row: the vector body truncates x*255, the scalar tail of each row (the last
// cols % 32 pixels on an AVX2 host) rounds half to even; `tail` selects the latter.
__device__ __forceinline__ bool hsv_row_tail(int64_t p, int row_width) {
    if (row_width <= 0) return false;
    const int x = (int)(p % row_width);
    return x >= row_width - (row_width & 31);
}
__device__ __forceinline__ void hsv2bgr_u8(int h8, int s8, int v8, bool tail, uint8_t &b, uint8_t &g, uint8_t &r) {
    const float h = __fmul_rn((float)h8, (float)(6.0 / 180.0));
    const float s = __fmul_rn((float)s8, (float)(1.0 / 255.0));
    const float v = __fmul_rn((float)v8, (float)(1.0 / 255.0));
    int sector = (int)h;                      // h >= 0: truncation
    const float f = __fsub_rn(h, (float)sector);
    sector = sector % 6;
    const float t0 = v;
    const float t1 = __fmul_rn(v, __fsub_rn(1.0f, s));
    const float t2 = __fmul_rn(v, __fmaf_rn(-s, f, 1.0f));
    const float t3 = __fmul_rn(v, __fmaf_rn(-s, __fsub_rn(1.0f, f), 1.0f));
    float fb, fg, fr;
    switch (sector) {
        case 0: fb = t1; fg = t3; fr = t0; break;
        case 1: fb = t1; fg = t0; fr = t2; break;
        case 2: fb = t3; fg = t0; fr = t1; break;
        case 3: fb = t0; fg = t2; fr = t1; break;
        case 4: fb = t0; fg = t1; fr = t3; break;
        default: fb = t2; fg = t1; fr = t0; break;
    }
    const float xb = __fmul_rn(fb, 255.0f), xg = __fmul_rn(fg, 255.0f), xr = __fmul_rn(fr, 255.0f);
    const int ib = tail ? __float2int_rn(xb) : (int)xb, ig = tail ? __float2int_rn(xg) : (int)xg,
              ir = tail ? __float2int_rn(xr) : (int)xr;
    b = (uint8_t)min(max(ib, 0), 255);
    g = (uint8_t)min(max(ig, 0), 255);
    r = (uint8_t)min(max(ir, 0), 255);
}

}  // namespace rti
