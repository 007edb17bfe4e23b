// Microbenchmark: sustained special-function-unit throughput of sm_100a, the roofline denominator of the
// per-pixel linear-RBF kernels (one MUFU.SQRT per distance):
//   sqrt   : chains of sqrt.approx.ftz.f32 only (16 independent chains per thread)
//   rsq    : rsqrt.approx.ftz.f32 only
//   mix    : the evaluation kernel's instruction mix per term (FADD, FFMA, MUFU.SQRT, two integer-pipe
//            widening ops, DFMA) on register operands: what the SFU sustains next to its feeders
// Prints G ops/s, ops per clock per SM at the clock measured with clock64(), for 8 / 16 / 32 / 64 warps per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peak_sfu peak_sfu.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ double widen(float r) {
    const uint32_t b = __float_as_uint(r);
    return __hiloint2double((int)((b >> 3) + 0x38000000u), (int)(b << 29));
}

template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, long long *clk, int iters, float seed) {
    float v[16];
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { v[i] = seed * (i + 1) + threadIdx.x * 1e-3f; acc[i] = 0.0; }
    const float gx = seed * 0.37f, dy2 = seed * 0.11f;
    const double w = 1.0 + seed;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                if (MODE == 0) asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
                else if (MODE == 1) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
                else {
                    const float dx = gx - v[i];
                    float rr;
                    asm volatile("sqrt.approx.ftz.f32 %0, %1;" : "=f"(rr) : "f"(fmaf(dx, dx, dy2)));
                    acc[i] = fma(widen(rr), w, acc[i]);
                    v[i] = rr;
                }
            }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += v[i] + (float)acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int MODE>
static int run(const char *name, int ctas_per_sm, int sms) {
    const int iters = 2000, grid = sms * ctas_per_sm;
    float *out;
    long long *clk;
    CHECK(cudaMalloc(&out, (size_t)grid * 256 * 4));
    CHECK(cudaMalloc(&clk, (size_t)grid * 8));
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k<MODE><<<grid, 256>>>(out, clk, iters, 1.5f);
    CHECK(cudaDeviceSynchronize());
    cudaEventRecord(a);
    k<MODE><<<grid, 256>>>(out, clk, iters, 1.5f);
    cudaEventRecord(b);
    CHECK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    long long h;
    CHECK(cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost));
    const double ops = (double)grid * 256 * iters * 64;
    const double mhz = h / (ms * 1e3);
    printf("{\"mode\": \"%s\", \"warps_per_sm\": %d, \"ms\": %.4f, \"gops\": %.1f, \"sm_mhz_from_clock64\": %.0f, "
           "\"ops_per_clk_per_sm\": %.2f}\n",
           name, ctas_per_sm * 8, ms, ops / ms / 1e6, mhz, ops / sms / (double)h);
    cudaFree(out);
    cudaFree(clk);
    return 0;
}

int main() {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int c : {1, 2, 4, 8}) {
        if (run<0>("sqrt", c, sms)) return 1;
        if (run<1>("rsq", c, sms)) return 1;
        if (run<2>("mix", c, sms)) return 1;
    }
    return 0;
}
