// Microbenchmark: sustained fp64 throughput of sm_100a for the roofline denominators of the
// fp64-bound kernels (PTM fit from frames, RBF fit / eval):
//   dfma   : independent DFMA chains (vector fp64 pipe)
//   dmma   : mma.sync.m8n8k4.f64 chains (fp64 tensor path)
//   mixed  : both interleaved in every warp (do they share a pipe?)
//   split  : half of the warps DFMA, half DMMA
//   rsq    : MUFU.RSQ64H (rsqrt.approx.ftz.f64)
//   f2f    : F2F.F32.F64 + F2F.F64.F32 round trips
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peak_fp64 peak_fp64.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int MODE>
__global__ void __launch_bounds__(256) k(double *out, int iters, double seed) {
    double acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = seed * (i + 1) + threadIdx.x;
    double a = seed + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
    const bool tensor_warp = (threadIdx.x >> 5) & 1;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
        } else if (MODE == 1) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int i = 0; i < 8; ++i) dmma(acc[2 * i], acc[2 * i + 1], a, b);
        } else if (MODE == 2) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
#pragma unroll
                for (int i = 0; i < 4; ++i) dmma(acc[2 * i], acc[2 * i + 1], a, b);
#pragma unroll
                for (int i = 8; i < 16; ++i) acc[i] = fma(acc[i], a, b);
#pragma unroll
                for (int i = 8; i < 16; ++i) acc[i] = fma(acc[i], a, b);
#pragma unroll
                for (int i = 8; i < 16; ++i) acc[i] = fma(acc[i], a, b);
#pragma unroll
                for (int i = 8; i < 16; ++i) acc[i] = fma(acc[i], a, b);
            }
        } else if (MODE == 3) {
            if (tensor_warp) {
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int i = 0; i < 8; ++i) dmma(acc[2 * i], acc[2 * i + 1], a, b);
            } else {
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
            }
        } else if (MODE == 4) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                double r;
                asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(acc[i]));
                acc[i] = r;
            }
        } else if (MODE == 5) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                float f;
                asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(f) : "d"(acc[i]));
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(acc[i]) : "f"(f));
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    if (s == 123.456) out[0] = s;
}

template <int MODE>
static int run(const char *name, double fma_per_thread_iter, double *d_out, int sms) {
    const int iters = 4096, blocks = sms * 8;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, 256>>>(d_out, 16, 0.5);
    CHECK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k<MODE><<<blocks, 256>>>(d_out, iters, 0.5);
        cudaEventRecord(e1);
        CHECK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double ops = (double)blocks * 256 * iters * fma_per_thread_iter;
    printf("{\"mode\": \"%s\", \"ms\": %.4f, \"gops_per_s\": %.1f}\n", name, best, ops / best * 1e-6);
    return 0;
}

int main() {
    cudaDeviceProp p;
    CHECK(cudaGetDeviceProperties(&p, 0));
    double *d_out;
    CHECK(cudaMalloc(&d_out, 8));
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_mhz\": %d}\n", p.name, p.multiProcessorCount, p.clockRate / 1000);
    // "ops" = fp64 FMAs (dmma m8n8k4 = 256 FMA per warp = 8 per thread)
    run<0>("dfma (G FMA/s)", 64, d_out, p.multiProcessorCount);
    run<1>("dmma m8n8k4 (G FMA/s)", 32 * 8, d_out, p.multiProcessorCount);
    run<2>("mixed dmma+dfma per warp (G FMA/s)", 4 * (4 * 8 + 32), d_out, p.multiProcessorCount);
    run<3>("split warps dmma|dfma (G FMA/s)", (32 * 8 + 64) / 2.0, d_out, p.multiProcessorCount);
    run<4>("rsqrt.approx.f64 (G op/s)", 16, d_out, p.multiProcessorCount);
    run<5>("cvt f64->f32->f64 (G pairs/s)", 16, d_out, p.multiProcessorCount);
    return 0;
}
