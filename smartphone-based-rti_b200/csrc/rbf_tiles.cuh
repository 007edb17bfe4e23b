// Building blocks shared by the per-pixel linear-RBF kernels (rbf.cu: fp64 Cholesky fit of one
// pixel per CTA; rbf_pcg.cu: fp64 inverse of a tile-centre matrix as the preconditioner of the
// tile's conjugate-gradient solves): tile constants, cp.async staging and the fp64
// tensor-core (DMMA) micro-kernel.
#pragma once
#include "common.cuh"

namespace rti {

constexpr int TS = 64;                  // tile edge
constexpr int KC = 32;                  // k-rows per 16 KB block of the L workspace
constexpr int BLK = KC * TS;            // doubles per block (16 KB)
constexpr int KS = 16;                  // k-rows staged per pipeline stage
constexpr int LDS = TS + 4;             // leading dimension of staged operands: the 4 k-rows of a DMMA
                                        // fragment land in disjoint bank groups (68*2 mod 32 = 8)
constexpr int STG = KS * LDS;           // doubles per staged operand chunk
constexpr int RBF_THREADS = 256;
constexpr int SD_LD = TS + 1;           // padded leading dimension of the potrf scratch

__host__ __device__ inline int rbf_mpad(int n) { return ((n - 1 + 2 + TS - 1) / TS) * TS; }

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// |d| for d2 >= 0 in fp64 (MUFU seed + Newton, ~2 ulp); 0 for d2 == 0
__device__ __forceinline__ double sqrt_pos(double d2) { return d2 > 0.0 ? d2 * rsqrt_pos(d2) : 0.0; }

// fp64 tensor-core step: D(8x8) += A(8x4) . B(4x8), mma.sync.m8n8k4 (DMMA).  Lane l holds
// A[l>>2][l&3], B[l&3][l>>2] and D[l>>2][2*(l&3) + {0,1}].  Measured on B200: 18.5 T FMA/s
// against 16.8 T for DFMA chains (tools/peak_fp64.cu), with 1/8 of the issue slots and 1/8 of
// the shared-memory operand traffic of a 4x4 register-tiled DFMA micro-kernel.
__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// A warp owns a 16 x 32 patch of the 64 x 64 tile: 2 x 4 DMMA tiles, acc[mi][ni][e] is element
// (rw + 8*mi + (lane>>2), cw + 8*ni + 2*(lane&3) + e).
// acc -= sum_k A[k][row] * B[k][col] over `depth` staged k-rows (operands [k][LD], k-major).
template <int LD>
__device__ __forceinline__ void mma_chunk(double (&acc)[2][4][2], const double *__restrict__ sA,
                                          const double *__restrict__ sB, int rw, int cw, int lane, int depth) {
    const double *pa = sA + (lane & 3) * LD + rw + (lane >> 2);
    const double *pb = sB + (lane & 3) * LD + cw + (lane >> 2);
#pragma unroll 4
    for (int k = 0; k < depth; k += 4) {
        double a[2], b[4];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) a[mi] = -pa[k * LD + 8 * mi];
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) b[ni] = pb[k * LD + 8 * ni];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) dmma(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
    }
}

__device__ __forceinline__ double block_sum(double v, double *s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < RBF_THREADS / 32; ++w) t += s_red[w];
    return t;
}

}  // namespace rti
