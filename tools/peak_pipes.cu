// Microbenchmark: do the 16-lane pipes of an sm_100a SM sub-partition (fp64, XU conversions, integer ALU)
// run next to each other or one after the other?  The PTM fit kernels issue 26 DFMA/DADD and 6-8
// conversions per light sample and measure ~the SUM of the two pipe times; this tool times each class of
// instruction alone and mixed, on register operands, for 8 / 16 / 32 warps per SM:
//   dfma      24 independent DFMA chains
//   widen     6 cvt.f64.f32            narrow   6 cvt.rn.f32.f64       f2i   8 cvt.rzi.s32.f64
//   alu       24 integer LOP3/IADD     ffma     48 FFMA
//   dfma+X    the two instruction streams interleaved in one loop body
// Prints SM clocks per loop iteration per warp scheduler (elapsed clocks / iterations / warps per
// scheduler): "a+b" equal to a + b means the pipes serialise, max(a, b) means they overlap.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peak_pipes peak_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

enum { DFMA = 1, WIDEN = 2, NARROW = 4, F2I = 8, ALU = 16, FFMA = 32 };

template <int MASK>
__global__ void __launch_bounds__(512, 1) k(float *out, long long *clk, int iters, float seed) {
    double acc[24], dsrc[8], wide[6];
    float fsrc[6], facc[48], narrow[6];
    int isrc[24], ires[8];
#pragma unroll
    for (int i = 0; i < 24; ++i) { acc[i] = seed * i; isrc[i] = threadIdx.x + i; }
#pragma unroll
    for (int i = 0; i < 48; ++i) facc[i] = seed + i;
#pragma unroll
    for (int i = 0; i < 8; ++i) { dsrc[i] = seed * 1000.0 + i + threadIdx.x; ires[i] = 0; }
#pragma unroll
    for (int i = 0; i < 6; ++i) { fsrc[i] = seed + i * 0.25f; wide[i] = 0.0; narrow[i] = 0.f; }
    const double m = 1.0 + 1e-9 * seed, c = 1e-7 * seed;
    const float fm = 1.0f + 1e-6f * seed, fc = 1e-6f * seed;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 24; ++i) {
            if (MASK & DFMA) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(acc[i]) : "d"(m), "d"(c));
            // conversions: the operand changes every iteration (one integer XOR with the loop counter) and
            // the result is folded into an integer (one more XOR), so that nothing is loop invariant
            if ((MASK & WIDEN) && i % 4 == 0) {
                const float x = __int_as_float(__float_as_int(fsrc[i / 4]) ^ (it & 0xffff));
                double w;
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(w) : "f"(x));
                ires[i / 4] ^= __double2hiint(w);
            }
            if ((MASK & NARROW) && i % 4 == 1) {
                const double x = __hiloint2double(__double2hiint(dsrc[i / 4]), it);
                float w;
                asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(w) : "d"(x));
                ires[i / 4] ^= __float_as_int(w);
            }
            if ((MASK & F2I) && i % 3 == 0) {
                const double x = __hiloint2double(__double2hiint(dsrc[i / 3]), it);
                int w;
                asm volatile("cvt.rzi.s32.f64 %0, %1;" : "=r"(w) : "d"(x));
                ires[i / 3] ^= w;
            }
            if (MASK & ALU) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(isrc[i]) : "r"(it), "r"(i));
            if (MASK & FFMA) {
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(facc[2 * i]) : "f"(fm), "f"(fc));
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(facc[2 * i + 1]) : "f"(fm), "f"(fc));
            }
        }
    }
    const long long t1 = clock64();
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 24; ++i) s += acc[i] + isrc[i];
#pragma unroll
    for (int i = 0; i < 48; ++i) s += facc[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) s += wide[i] + narrow[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) s += ires[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int MASK>
static int run(const char *name, int warps, int sms) {
    const int iters = 4000, grid = sms, threads = warps * 32;     // one CTA per SM
    float *out;
    long long *clk;
    CHECK(cudaMalloc(&out, (size_t)grid * 512 * 4));
    CHECK(cudaMalloc(&clk, (size_t)grid * 8));
    k<MASK><<<grid, threads>>>(out, clk, iters, 1.5f);
    CHECK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MASK><<<grid, threads>>>(out, clk, iters, 1.5f);
    cudaEventRecord(e1);
    CHECK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h;
    CHECK(cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost));
    // clocks of CTA 0 per iteration, per warp of a scheduler; the kernel's wall time says whether all CTAs
    // ran side by side (kernel_ms * clock ~ cta0 clocks) or in waves
    printf("{\"mode\": \"%s\", \"warps_per_sm\": %d, \"clk_per_iter_per_scheduler\": %.1f, \"cta0_clk_per_iter\": %.1f, "
           "\"kernel_ms\": %.4f, \"kernel_clk_per_iter_at_1920mhz\": %.1f}\n", name, warps,
           (double)h / iters / (warps / 4), (double)h / iters, ms, ms * 1e-3 * 1.92e9 / iters);
    cudaFree(out);
    cudaFree(clk);
    return 0;
}

int main() {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int c : {8, 16}) {      // warps per SM
        if (run<DFMA>("dfma24", c, sms)) return 1;
        if (run<WIDEN>("widen6", c, sms)) return 1;
        if (run<NARROW>("narrow6", c, sms)) return 1;
        if (run<F2I>("f2i8", c, sms)) return 1;
        if (run<ALU>("alu24", c, sms)) return 1;
        if (run<FFMA>("ffma48", c, sms)) return 1;
        if (run<DFMA | WIDEN>("dfma24+widen6", c, sms)) return 1;
        if (run<DFMA | NARROW>("dfma24+narrow6", c, sms)) return 1;
        if (run<DFMA | F2I>("dfma24+f2i8", c, sms)) return 1;
        if (run<DFMA | ALU>("dfma24+alu24", c, sms)) return 1;
        if (run<DFMA | FFMA>("dfma24+ffma48", c, sms)) return 1;
        if (run<WIDEN | ALU>("widen6+alu24", c, sms)) return 1;
        if (run<F2I | ALU>("f2i8+alu24", c, sms)) return 1;
        if (run<DFMA | WIDEN | ALU>("dfma24+widen6+alu24", c, sms)) return 1;
    }
    return 0;
}
