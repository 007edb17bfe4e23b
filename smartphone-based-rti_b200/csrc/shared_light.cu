// Shared-light helpers: detection (do all pixels carry the same light vectors?) and the
// 6 x n pseudo-inverse of the common PTM design matrix (analysis.py:282-298 collapses to
// coef[6][P] = pinv . I[n][P], see rti_ptm_fit_shared).
#include "common.cuh"

namespace rti {

// flag += 1 if any pixel's (lx, ly) row differs bitwise from pixel 0's.  One thread = 4 lights
// of one pixel when the rows are 16-byte aligned, else 1 light.
template <int VEC>
__global__ void lights_shared_kernel(const float *__restrict__ lx, const float *__restrict__ ly,
                                     long long n_pixels, int n, int32_t *__restrict__ flag) {
    const int per_row = n / VEC;
    const long long total = n_pixels * per_row;
    bool differs = false;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const long long p = t / per_row;
        const int c = (int)(t - p * per_row);
        if (VEC == 4) {
            const uint4 a = __ldg(reinterpret_cast<const uint4 *>(lx + p * n) + c);
            const uint4 a0 = __ldg(reinterpret_cast<const uint4 *>(lx) + c);
            const uint4 b = __ldg(reinterpret_cast<const uint4 *>(ly + p * n) + c);
            const uint4 b0 = __ldg(reinterpret_cast<const uint4 *>(ly) + c);
            differs |= (a.x != a0.x) | (a.y != a0.y) | (a.z != a0.z) | (a.w != a0.w) | (b.x != b0.x) |
                       (b.y != b0.y) | (b.z != b0.z) | (b.w != b0.w);
        } else {
            differs |= (__float_as_uint(__ldg(lx + p * n + c)) != __float_as_uint(__ldg(lx + c))) |
                       (__float_as_uint(__ldg(ly + p * n + c)) != __float_as_uint(__ldg(ly + c)));
        }
    }
    if (__any_sync(0xffffffffu, differs) && (threadIdx.x & 31) == 0) atomicAdd(flag, 1);
}

// One CTA: normal equations of the n design rows (fp64 sums of the float32-rounded monomials,
// exactly the matrix the per-pixel kernels accumulate), then pinv[:, i] = (A^T A)^-1 m_i by the
// same in-register Cholesky.  pinv is [6][n].
constexpr int PINV_THREADS = 256;

__global__ void __launch_bounds__(PINV_THREADS)
ptm_shared_pinv_kernel(const float *__restrict__ lx, const float *__restrict__ ly, int n,
                       double *__restrict__ pinv, int32_t *__restrict__ bad) {
    __shared__ double s_part[PINV_THREADS / 32][20];
    __shared__ double s_sum[20];
    NormalEq ne;
    ne.clear();
    for (int i = threadIdx.x; i < n; i += PINV_THREADS) ne.add(lx[i], ly[i], 0.0);
    double v[20];
#pragma unroll
    for (int k = 0; k < 15; ++k) v[k] = ne.g[k];
#pragma unroll
    for (int k = 0; k < 5; ++k) v[15 + k] = ne.s[k];
#pragma unroll
    for (int k = 0; k < 20; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 20; ++k) s_part[warp][k] = v[k];
    }
    __syncthreads();
    if (threadIdx.x < 20) {
        double t = 0.0;
        for (int w = 0; w < PINV_THREADS / 32; ++w) t += s_part[w][threadIdx.x];
        s_sum[threadIdx.x] = t;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 15; ++k) ne.g[k] = s_sum[k];
#pragma unroll
    for (int k = 0; k < 5; ++k) ne.s[k] = s_sum[15 + k];
    bool all_ok = true;
    for (int i = threadIdx.x; i < n; i += PINV_THREADS) {
        const float lu = lx[i], lv = ly[i];
        ne.r[0] = (double)__fmul_rn(lu, lu);
        ne.r[1] = (double)__fmul_rn(lv, lv);
        ne.r[2] = (double)__fmul_rn(lu, lv);
        ne.r[3] = (double)lu;
        ne.r[4] = (double)lv;
        ne.r[5] = 1.0;
        double a[6];
        all_ok = solve_normal_eq(ne, (double)n, a) && all_ok;
#pragma unroll
        for (int k = 0; k < 6; ++k) pinv[(long long)k * n + i] = a[k];
    }
    if (threadIdx.x == 0 && !all_ok && bad != nullptr) atomicAdd(bad, 1);
}

}  // namespace rti

using namespace rti;

extern "C" int rti_lights_shared(const float *d_lx, const float *d_ly, int64_t n_pixels, int n, int32_t *d_flag,
                                 rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_flag, "rti_lights_shared: NULL pointer");
    RTI_REQUIRE(n_pixels > 0 && n > 0, "rti_lights_shared: empty input");
    const bool vec = (n % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_lx) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(d_ly) & 15) == 0);
    const long long total = n_pixels * (vec ? n / 4 : n);
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (vec)
        lights_shared_kernel<4><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(d_lx, d_ly, n_pixels, n, d_flag);
    else
        lights_shared_kernel<1><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(d_lx, d_ly, n_pixels, n, d_flag);
    RTI_LAUNCHED("lights_shared_kernel");
    return RTI_OK;
}

extern "C" int rti_ptm_shared_pinv(const float *d_lx, const float *d_ly, int n, double *d_pinv, int32_t *d_bad,
                                   rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_pinv, "rti_ptm_shared_pinv: NULL pointer");
    if (n < 6) {
        set_error("rti_ptm_shared_pinv: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    ptm_shared_pinv_kernel<<<1, PINV_THREADS, 0, as_stream(stream)>>>(d_lx, d_ly, n, d_pinv, d_bad);
    RTI_LAUNCHED("ptm_shared_pinv_kernel");
    return RTI_OK;
}
