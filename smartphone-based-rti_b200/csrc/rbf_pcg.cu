// Linear RBF fit with per-pixel lights, tile-preconditioned (reference: _interpolate_RBF,
// analysis.py:249-260 -> scipy Rbf('linear'): w = solve(A, d), A[j][k] = |n_j - n_k|).
//
// The fp64 Cholesky of rbf.cu costs n^3/3 DFMA per pixel.  The node sets of neighbouring
// pixels differ by ~1e-3 per pixel step (the light vector moves with p = (x, y, 0),
// analysis.py:228-229), so one factorisation can serve a whole tile of pixels as a
// PRECONDITIONER:
//
//   * per 16 x 16 pixel tile: the reduced SPD matrix B0 of the tile-centre pixel
//     (B[j][k] = a_j + a_k - |n_j - n_k|, a_j = |n_j - n_0|, see rbf.cu) is inverted in fp64
//     on the fp64 tensor cores (left-looking blocked Cholesky with an identity border, then
//     Z Z^T with Z = L^-T) and stored as fp16 hi + lo with power-of-two row scales;
//   * per pixel: two preconditioned conjugate-gradient solves  B p = d_0 1 - d',  B q = a
//     (the same two systems the direct kernel solves), sharing every matrix entry.  B is never
//     stored: a step regenerates its (n-1)^2 distances in float32 (MUFU.SQRT, the same rounded
//     value every step, i.e. a FIXED matrix within 1e-7 of B) and accumulates the products in
//     fp64 on the DFMA pipe; search directions are rounded to float32 once per step (any vector
//     may serve as a search direction), iterates and residuals are fp64;
//   * the preconditioner application z = B0^-1 r of all pixels of all tiles is ONE batched
//     tcgen05 contraction per step (rbf_shared.cu: fp16 hi + lo operator against the residuals
//     split into fp16 hi + lo, fp32 accumulation in TMEM, TMA-fed).
//
// A pixel leaves the iteration when both residuals are below the tolerance (max norm, in the
// units of the intensities); what has not converged after max_iters steps is reported so that the
// caller can hand those pixels to the direct fp64 kernel.
#include <cuda_fp16.h>
#include "rbf_tiles.cuh"

namespace rti {

int rbf_gemm_batched(const uint16_t *d_packed, int k_rows, long long n_cols, const uint16_t *d_op_hi,
                     const uint16_t *d_op_lo, const float *d_inv_scale, int n_ops, int op_rows_pad, int a_k_pad,
                     int blks_per_op, double *d_out, const int32_t *d_tile_active, cudaStream_t stream);

constexpr int PT = 16;                  // tile edge in pixels
constexpr int PTILE = PT * PT;          // pixels per tile = columns of one right-hand side block
constexpr int PCG_THREADS = 256;

struct PcgGeom {
    int rows, cols, n, tiles_x;
    long long tile0;                    // first tile of this batch (row-major tile index)
    int n_tiles;                        // tiles in this batch
    int m;                              // n - 1 unknowns
    int k_pad;                          // m rounded up to 64 (operator columns, vector length in the GEMM)
    int r_pad;                          // m rounded up to 128 (operator rows, stride of the vectors)
};

__device__ __forceinline__ float sqrt_f32(float d2) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d2));
    return r;
}
__device__ __forceinline__ double widen_pos(float r) {          // r >= 0; 0 maps to 2^-127
    const uint32_t b = __float_as_uint(r);
    return __hiloint2double((int)((b >> 3) + 0x38000000u), (int)(b << 29));
}
// entry of the (fixed) float32-rounded distance matrix, as used by every step and by the init
__device__ __forceinline__ float dist_f32(float xj, float yj, float xk, float yk) {
    const float dx = xj - xk, dy = yj - yk;
    return sqrt_f32(fmaf(dx, dx, dy * dy));
}

__device__ __forceinline__ bool tile_pixel(const PcgGeom &g, int tb, int slot, long long *pix) {
    const long long tile = g.tile0 + tb;
    const int ty = (int)(tile / g.tiles_x), tx = (int)(tile - (long long)ty * g.tiles_x);
    const int y = ty * PT + slot / PT, x = tx * PT + slot % PT;
    *pix = (long long)y * g.cols + x;
    return y < g.rows && x < g.cols;
}

// sum / max of up to four values over the CTA (all threads get the result)
template <int NV, bool MAX>
__device__ __forceinline__ void block_reduce(double (&v)[NV], double *s_red) {
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double t = __shfl_xor_sync(0xffffffffu, v[i], o);
            v[i] = MAX ? fmax(v[i], t) : v[i] + t;
        }
    __syncthreads();
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int i = 0; i < NV; ++i) s_red[(threadIdx.x >> 5) * NV + i] = v[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        double t = s_red[i];
#pragma unroll
        for (int w = 1; w < PCG_THREADS / 32; ++w) t = MAX ? fmax(t, s_red[w * NV + i]) : t + s_red[w * NV + i];
        v[i] = t;
    }
}

// residual column -> fp16 hi + lo rows of the packed GEMM operand, scaled by a power of two so that
// the largest entry lies in [2^10, 2^11); returns the unscale factor for z
__device__ __forceinline__ float pack_scale(double rmax, double *scale) {
    int e = (rmax > 0.0 && rmax < 1e300) ? ilogb(rmax) : 10;
    e = max(-900, min(900, e));
    *scale = scalbn(1.0, 10 - e);
    return (float)scalbn(1.0, e - 10);
}
// column `col` of the packed operand: [hi halves of rows 0..k_pad) | lo halves], contiguous (K-major)
__device__ __forceinline__ void pack_store(uint16_t *packed, long long col, int k_rows, int k_pad, int j, double v) {
    const __half h = __double2half(v);
    const __half l = __double2half(v - (double)__half2float(h));
    packed[col * k_rows + j] = __half_as_ushort(h);
    packed[col * k_rows + k_pad + j] = __half_as_ushort(l);
}

// ------------------------------------------------------------------------------------------
// Preconditioner: fp64 inverse of the tile-centre pixel's reduced matrix.
// Bordered matrix [[B], [I]] of 2*mp rows (mp = m rounded up to 64): the left-looking blocked
// Cholesky of rbf.cu turns the identity border into Z = L^-T (upper block triangular: tile
// (iz, kb) is zero for kb < iz and its update sum starts at k = 64*iz), then B^-1 = Z Z^T.
// Workspace per CTA (doubles): L row blocks [2*nT][mp k-rows][64], the D^-T tiles [nT][64][64],
// and B^-1 [mp][mp] before it is scaled and split.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RBF_THREADS, 2)
rbf_tile_inverse_kernel(const float *__restrict__ lx, const float *__restrict__ ly, PcgGeom g,
                        double *__restrict__ work, int64_t work_stride, uint16_t *__restrict__ op_hi,
                        uint16_t *__restrict__ op_lo, float *__restrict__ inv_scale, int32_t *__restrict__ bad) {
    extern __shared__ __align__(16) double smem[];
    double *sA = smem;
    double *sB = smem + 2 * STG;
    double *sD = smem;                          // alias: [64][65]
    double *sCk = smem;                         // alias: [64 k][LDS]
    double *sW = smem + 4 * STG;                // [64 k][LDS]: sW[k][c] = Dinv[c][k]
    double *sRsq = sW + TS * LDS;
    double *sx = sRsq + TS + 8;
    double *sy = sx + g.n;
    double *sa = sy + g.n;

    const int n = g.n;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int rw = 16 * (warp >> 1);
    const int cw = 32 * (warp & 1);
    const int er = rw + (lane >> 2);
    const int ec = cw + 2 * (lane & 3);
    const int M = g.m;
    const int mp = g.k_pad;
    const int nT = mp / TS, nKC = mp / KC;
    double *L = work + (int64_t)blockIdx.x * work_stride;       // row block rb at rb * nKC * BLK
    double *Wg = L + (int64_t)2 * nT * nKC * BLK;
    double *Minv = Wg + (int64_t)nT * TS * TS;

    for (int tb = blockIdx.x; tb < g.n_tiles; tb += gridDim.x) {
        const long long tile = g.tile0 + tb;
        const int ty = (int)(tile / g.tiles_x), tx = (int)(tile - (long long)ty * g.tiles_x);
        const int cy = min(ty * PT + PT / 2, g.rows - 1), cx = min(tx * PT + PT / 2, g.cols - 1);
        const long long pix = (long long)cy * g.cols + cx;
        __syncthreads();
        for (int i = t; i < n; i += RBF_THREADS) {
            sx[i] = (double)lx[pix * n + i];
            sy[i] = (double)ly[pix * n + i];
        }
        __syncthreads();
        for (int i = t; i < n; i += RBF_THREADS) {
            const double dx = sx[i] - sx[0], dy = sy[i] - sy[0];
            sa[i] = sqrt_pos(fma(dx, dx, dy * dy));
        }
        __syncthreads();
        bool ok = true;

        for (int kb = 0; kb < nT; ++kb) {
            for (int it = kb; it < 2 * nT; ++it) {
                const bool border = it >= nT;
                const int iz = it - nT;
                if (border && kb < iz) continue;                              // zero tile of Z
                double acc[2][4][2];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi) {
                    const int lr = (border ? iz : it) * TS + er + 8 * mi;     // row inside B or inside the border
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int gc = kb * TS + ec + 8 * ni + e;
                            double v;
                            if (border || gc >= M) v = (lr == gc) ? 1.0 : 0.0;   // identity border / padding
                            else if (lr < M) {
                                const double dx = sx[lr + 1] - sx[gc + 1], dy = sy[lr + 1] - sy[gc + 1];
                                v = sa[lr + 1] + sa[gc + 1] - sqrt_pos(fma(dx, dx, dy * dy));
                            } else v = 0.0;
                            acc[mi][ni][e] = v;
                        }
                }
                const int kc0 = border ? iz * (TS / KS) : 0;
                const int nk = kb * (TS / KS);
                const bool diag = (it == kb);
                const double *ga = L + (int64_t)it * nKC * BLK;
                const double *gb = L + (int64_t)kb * nKC * BLK;
                auto issue = [&](int buf, int kc) {
#pragma unroll
                    for (int i = 0; i < KS * TS * 8 / 16 / RBF_THREADS; ++i) {
                        const int ch = t + i * RBF_THREADS;
                        const int so = (ch >> 5) * LDS + (ch & 31) * 2;
                        const int64_t go = (int64_t)kc * KS * TS + ch * 2;
                        cp_async16(sA + buf * STG + so, ga + go);
                        if (!diag) cp_async16(sB + buf * STG + so, gb + go);
                    }
                    cp_async_commit();
                };
                __syncthreads();
                if (nk > kc0) issue(0, kc0);
                for (int kc = kc0; kc < nk; ++kc) {
                    const int buf = (kc - kc0) & 1;
                    if (kc + 1 < nk) { issue(buf ^ 1, kc + 1); cp_async_wait<1>(); }
                    else cp_async_wait<0>();
                    __syncthreads();
                    const double *pa = sA + buf * STG;
                    mma_chunk<LDS>(acc, pa, diag ? pa : sB + buf * STG, rw, cw, lane, KS);
                    __syncthreads();
                }

                if (diag) {
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int r = er + 8 * mi, c = ec + 8 * ni + e;
                                const int gr = kb * TS + r, gc = kb * TS + c;
                                sD[r * SD_LD + c] = (gc >= M) ? ((gr == gc) ? 1.0 : 0.0) : acc[mi][ni][e];
                            }
                    // blocked right-looking Cholesky of the diagonal tile (as in rbf_fit_kernel)
                    const int cl = min(TS, M - kb * TS);
                    for (int jb = 0; jb < TS / 16; ++jb) {
                        const int o = 16 * jb;
                        __syncthreads();
                        if (warp == 0) {
                            for (int j = 0; j < 16; ++j) {
                                const int cj = o + j;
                                const double piv = sD[cj * SD_LD + cj];
                                const double ref = (cj < cl) ? 2.0 * sa[kb * TS + cj + 1] : 1.0;
                                if (!(piv > 1e-13 * ref) || !(piv > 0.0)) ok = false;
                                const double rs = rsqrt(piv);
                                __syncwarp();
                                const int r = cj + 1 + lane;
                                if (lane == 0) { sD[cj * SD_LD + cj] = piv * rs; sRsq[cj] = rs; }
                                if (r < o + 16) sD[r * SD_LD + cj] *= rs;
                                __syncwarp();
                                const int r2 = cj + 1 + (lane >> 1);
                                if (r2 < o + 16) {
                                    const double lrj = sD[r2 * SD_LD + cj];
                                    for (int c = cj + 1 + (lane & 1); c <= r2; c += 2)
                                        if (c < cl)
                                            sD[r2 * SD_LD + c] = fma(-lrj, sD[c * SD_LD + cj], sD[r2 * SD_LD + c]);
                                }
                                __syncwarp();
                            }
                        }
                        if (jb == TS / 16 - 1) break;
                        __syncthreads();
                        if (t < TS - o - 16) {
                            const int r = o + 16 + t;
                            double x[16];
#pragma unroll
                            for (int c = 0; c < 16; ++c) x[c] = sD[r * SD_LD + o + c];
#pragma unroll
                            for (int pp = 0; pp < 16; ++pp) {
                                if (o + pp < cl) {
                                    x[pp] *= sRsq[o + pp];
#pragma unroll
                                    for (int c = pp + 1; c < 16; ++c)
                                        if (o + c < cl) x[c] = fma(-x[pp], sD[(o + c) * SD_LD + o + pp], x[c]);
                                }
                            }
#pragma unroll
                            for (int c = 0; c < 16; ++c) sD[r * SD_LD + o + c] = x[c];
                        }
                        __syncthreads();
                        {
                            const int tt = (TS - o - 16) / 4;
                            const int tr = t / tt, tc = t - tr * tt;
                            if (tr < tt && tc <= tr) {
                                const int rb = o + 16 + 4 * tr, cb = o + 16 + 4 * tc;
                                double u[4][4];
#pragma unroll
                                for (int i = 0; i < 4; ++i)
#pragma unroll
                                    for (int j = 0; j < 4; ++j) u[i][j] = 0.0;
#pragma unroll 4
                                for (int pp = 0; pp < 16; ++pp) {
                                    double xr[4], xc[4];
#pragma unroll
                                    for (int i = 0; i < 4; ++i) {
                                        xr[i] = sD[(rb + i) * SD_LD + o + pp];
                                        xc[i] = sD[(cb + i) * SD_LD + o + pp];
                                    }
#pragma unroll
                                    for (int i = 0; i < 4; ++i)
#pragma unroll
                                        for (int j = 0; j < 4; ++j) u[i][j] = fma(xr[i], xc[j], u[i][j]);
                                }
#pragma unroll
                                for (int i = 0; i < 4; ++i)
#pragma unroll
                                    for (int j = 0; j < 4; ++j)
                                        if (cb + j <= rb + i && cb + j < cl) sD[(rb + i) * SD_LD + cb + j] -= u[i][j];
                            }
                        }
                    }
                    __syncthreads();
                    for (int e = t; e < TS * TS; e += RBF_THREADS) {
                        const int r = e >> 6, c = e & 63;
                        if (c > r) sD[r * SD_LD + c] = 0.0;
                    }
                    __syncthreads();
                    {
                        const int k = t >> 2, q = t & 3;
                        for (int c = 0; c < TS; ++c) {
                            double part = 0.0;
                            for (int p = k + q; p < c; p += 4) part = fma(sD[c * SD_LD + p], sW[k * LDS + p], part);
                            part += __shfl_xor_sync(0xffffffffu, part, 1);
                            part += __shfl_xor_sync(0xffffffffu, part, 2);
                            double v;
                            if (c < k) v = 0.0;
                            else if (c == k) v = sRsq[c];
                            else v = -part * sRsq[c];
                            if (q == 0) sW[k * LDS + c] = v;
                            __syncwarp();
                        }
                    }
                    __syncthreads();
                    double *gd = L + ((int64_t)kb * nKC + 2 * kb) * BLK;
                    for (int e = t; e < TS * TS; e += RBF_THREADS) {
                        const int c = e >> 6, r = e & 63;
                        gd[c * TS + r] = sD[r * SD_LD + c];
                        Wg[(int64_t)kb * TS * TS + e] = sW[c * LDS + r];
                    }
                } else {
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int r = er + 8 * mi, c = ec + 8 * ni + e;
                                const int gc = kb * TS + c;
                                const int lr = (border ? iz : it) * TS + r;
                                // padding columns keep the identity: X = C there (D of a padding column is 1)
                                sCk[c * LDS + r] = (!border && gc >= M) ? ((lr == gc) ? 1.0 : 0.0) : acc[mi][ni][e];
                            }
                    __syncthreads();
                    double x[2][4][2];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni) x[mi][ni][0] = x[mi][ni][1] = 0.0;
                    mma_chunk<LDS>(x, sCk, sW, rw, cw, lane, TS);                    // x = -C.D^-T
                    double *gx = L + ((int64_t)it * nKC + 2 * kb) * BLK;
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                            for (int e = 0; e < 2; ++e)
                                gx[(ec + 8 * ni + e) * TS + er + 8 * mi] = -x[mi][ni][e];
                }
            }
        }

        // ---- B^-1 = Z Z^T: tile (I, J <= I) = sum over k >= 64 I of Z[I][k] Z[J][k]
        for (int I = 0; I < nT; ++I) {
            for (int J = 0; J <= I; ++J) {
                double acc[2][4][2];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
                const bool diag = (I == J);
                const double *ga = L + (int64_t)(nT + I) * nKC * BLK;
                const double *gb = L + (int64_t)(nT + J) * nKC * BLK;
                const int kc0 = I * (TS / KS), nk = nT * (TS / KS);
                auto issue = [&](int buf, int kc) {
#pragma unroll
                    for (int i = 0; i < KS * TS * 8 / 16 / RBF_THREADS; ++i) {
                        const int ch = t + i * RBF_THREADS;
                        const int so = (ch >> 5) * LDS + (ch & 31) * 2;
                        const int64_t go = (int64_t)kc * KS * TS + ch * 2;
                        cp_async16(sA + buf * STG + so, ga + go);
                        if (!diag) cp_async16(sB + buf * STG + so, gb + go);
                    }
                    cp_async_commit();
                };
                __syncthreads();
                issue(0, kc0);
                for (int kc = kc0; kc < nk; ++kc) {
                    const int buf = (kc - kc0) & 1;
                    if (kc + 1 < nk) { issue(buf ^ 1, kc + 1); cp_async_wait<1>(); }
                    else cp_async_wait<0>();
                    __syncthreads();
                    const double *pa = sA + buf * STG;
                    mma_chunk<LDS>(acc, pa, diag ? pa : sB + buf * STG, rw, cw, lane, KS);
                    __syncthreads();
                }
#pragma unroll
                for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int r = I * TS + er + 8 * mi, c = J * TS + ec + 8 * ni + e;
                            const double v = (r < M && c < M) ? -acc[mi][ni][e] : 0.0;
                            Minv[(int64_t)r * mp + c] = v;
                            if (!diag) Minv[(int64_t)c * mp + r] = v;
                        }
            }
        }
        __syncthreads();
        // ---- row scale (largest entry into [2^13, 2^14)) and fp16 hi + lo split, K-major rows
        for (int r = warp; r < g.r_pad; r += RBF_THREADS / 32) {
            double mx = 0.0;
            if (r < M)
                for (int c = lane; c < M; c += 32) mx = fmax(mx, fabs(Minv[(int64_t)r * mp + c]));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            const bool fin = mx < 1.0e300;
            int e = (mx > 0.0 && fin) ? ilogb(mx) : 13;
            e = max(-100, min(100, e));
            const double sc = scalbn(1.0, 13 - e);
            if (!fin) ok = false;
            const int64_t orow = ((int64_t)tb * g.r_pad + r) * mp;
            if (lane == 0) inv_scale[(int64_t)tb * g.r_pad + r] = (r < M) ? (float)scalbn(1.0, e - 13) : 0.0f;
            for (int c = lane; c < mp; c += 32) {
                const double v = (r < M && c < M && fin) ? Minv[(int64_t)r * mp + c] * sc : 0.0;
                const __half h = __double2half(v);
                const __half l = __double2half(v - (double)__half2float(h));
                op_hi[orow + c] = __half_as_ushort(h);
                op_lo[orow + c] = __half_as_ushort(l);
            }
        }
        const int any_bad = __syncthreads_or(!ok);
        if (t == 0 && any_bad && bad != nullptr) atomicAdd(bad, 1);
    }
}

// ------------------------------------------------------------------------------------------
// State of the iteration, all indexed by GEMM column col = tb * 2*PTILE + rhs * PTILE + slot:
//   X, R      float64 [col][r_pad]   iterate and residual
//   Pv        float32 [col][r_pad]   search direction (rounded once, used rounded)
//   Z         float64 [col][r_pad]   preconditioned residual of the last contraction, times 1/zscale
//   packed    fp16 [col][2*k_pad], the contraction's vector operand (residual hi halves, then lo halves)
//   rz, zscale per column; done, iters per pixel; tile_active per tile.
// ------------------------------------------------------------------------------------------
struct PcgState {
    double *X, *R, *Z, *rz;
    float *Pv, *zscale;
    uint16_t *packed;
    int32_t *done, *iters, *tile_active;
};

__global__ void __launch_bounds__(PCG_THREADS)
rbf_pcg_init_kernel(const float *__restrict__ lx, const float *__restrict__ ly, const int32_t *__restrict__ inten,
                    PcgGeom g, PcgState st) {
    __shared__ double s_red[PCG_THREADS / 32 * 4];
    const int slot = blockIdx.x, tb = blockIdx.y, t = threadIdx.x;
    const long long pidx = (long long)tb * PTILE + slot;
    const long long col0 = (long long)tb * 2 * PTILE + slot, col1 = col0 + PTILE;
    const int k_rows = 2 * g.k_pad;
    long long pix;
    const bool valid = tile_pixel(g, tb, slot, &pix);
    const int n = g.n, M = g.m;
    float x0 = 0.f, y0 = 0.f;
    double d0 = 0.0;
    if (valid) {
        x0 = lx[pix * n];
        y0 = ly[pix * n];
        d0 = (double)inten[pix * n];
    }
    double r0[4], r1[4];                                          // rows t + 256 i (M <= 1024)
    double mx[2] = {0.0, 0.0};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int j = t + i * PCG_THREADS;
        r0[i] = r1[i] = 0.0;
        if (valid && j < M) {
            r0[i] = d0 - (double)inten[pix * n + j + 1];
            r1[i] = widen_pos(dist_f32(lx[pix * n + j + 1], ly[pix * n + j + 1], x0, y0));
        }
        mx[0] = fmax(mx[0], fabs(r0[i]));
        mx[1] = fmax(mx[1], fabs(r1[i]));
    }
    block_reduce<2, true>(mx, s_red);
    double sc0, sc1;
    const float zs0 = pack_scale(mx[0], &sc0), zs1 = pack_scale(mx[1], &sc1);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int j = t + i * PCG_THREADS;
        if (j < g.r_pad) {
            st.X[col0 * g.r_pad + j] = 0.0;
            st.X[col1 * g.r_pad + j] = 0.0;
            st.R[col0 * g.r_pad + j] = r0[i];
            st.R[col1 * g.r_pad + j] = r1[i];
            st.Pv[col0 * g.r_pad + j] = 0.f;
            st.Pv[col1 * g.r_pad + j] = 0.f;
        }
        if (j < g.k_pad) {
            pack_store(st.packed, col0, k_rows, g.k_pad, j, r0[i] * sc0);
            pack_store(st.packed, col1, k_rows, g.k_pad, j, r1[i] * sc1);
        }
    }
    if (t == 0) {
        st.rz[col0] = 0.0;
        st.rz[col1] = 0.0;
        st.zscale[col0] = zs0;
        st.zscale[col1] = zs1;
        st.iters[pidx] = 0;
        // a pixel outside the image, or with nothing to solve, never enters the iteration
        const bool idle = !valid || (mx[0] == 0.0 && mx[1] == 0.0);
        st.done[pidx] = idle ? (valid ? 1 : 2) : 0;
        if (!idle) atomicAdd(st.tile_active + tb, 1);
    }
}

// One conjugate-gradient step of both systems of one pixel.  ROWS = rows of the matrix per thread.
template <int ROWS, int MINB>
__global__ void __launch_bounds__(PCG_THREADS, MINB)
rbf_pcg_step_kernel(const float *__restrict__ lx, const float *__restrict__ ly, PcgGeom g, PcgState st, int iter,
                    double tol0, double tol1) {
    extern __shared__ __align__(16) unsigned char pcg_smem[];
    const int slot = blockIdx.x, tb = blockIdx.y, t = threadIdx.x;
    const long long pidx = (long long)tb * PTILE + slot;
    if (st.done[pidx] != 0) return;
    const int M = g.m, n = g.n, r_pad = g.r_pad;
    double2 *s_p = reinterpret_cast<double2 *>(pcg_smem);                     // [r_pad] search directions (both systems)
    float2 *s_xy = reinterpret_cast<float2 *>(s_p + r_pad);                   // [r_pad] nodes 1..M
    double *s_red = reinterpret_cast<double *>(s_xy + r_pad);                 // [8 * 4]
    long long pix;
    tile_pixel(g, tb, slot, &pix);
    const long long col0 = (long long)tb * 2 * PTILE + slot, col1 = col0 + PTILE;
    const int k_rows = 2 * g.k_pad;
    const float x0 = lx[pix * n], y0 = ly[pix * n];

    double z0[ROWS], z1[ROWS];
    float xj[ROWS], yj[ROWS];
    const float zs0 = st.zscale[col0], zs1 = st.zscale[col1];
    double red[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
        const int j = t + i * PCG_THREADS;
        z0[i] = z1[i] = 0.0;
        xj[i] = yj[i] = 0.f;
        double r0 = 0.0, r1 = 0.0;
        if (j < M) {
            xj[i] = lx[pix * n + j + 1];
            yj[i] = ly[pix * n + j + 1];
            z0[i] = st.Z[col0 * r_pad + j] * (double)zs0;
            z1[i] = st.Z[col1 * r_pad + j] * (double)zs1;
            r0 = st.R[col0 * r_pad + j];
            r1 = st.R[col1 * r_pad + j];
        }
        if (j < r_pad) s_xy[j] = make_float2(xj[i], yj[i]);
        red[0] = fma(r0, z0[i], red[0]);
        red[1] = fma(r1, z1[i], red[1]);
    }
    block_reduce<2, false>(reinterpret_cast<double(&)[2]>(red), s_red);
    const double rz0 = red[0], rz1 = red[1];
    const double rzo0 = st.rz[col0], rzo1 = st.rz[col1];
    const double beta0 = (iter > 0 && rzo0 != 0.0) ? rz0 / rzo0 : 0.0;
    const double beta1 = (iter > 0 && rzo1 != 0.0) ? rz1 / rzo1 : 0.0;
    red[0] = red[1] = red[2] = red[3] = 0.0;
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
        const int j = t + i * PCG_THREADS;
        double p0 = 0.0, p1 = 0.0, aj = 0.0;
        if (j < M) {
            // the direction is rounded to float32 ONCE and used rounded everywhere
            p0 = (double)(float)fma(beta0, (double)st.Pv[col0 * r_pad + j], z0[i]);
            p1 = (double)(float)fma(beta1, (double)st.Pv[col1 * r_pad + j], z1[i]);
            st.Pv[col0 * r_pad + j] = (float)p0;
            st.Pv[col1 * r_pad + j] = (float)p1;
            aj = widen_pos(dist_f32(xj[i], yj[i], x0, y0));
        }
        if (j < r_pad) s_p[j] = make_double2(p0, p1);
        red[0] += p0;
        red[1] += p1;
        red[2] = fma(aj, p0, red[2]);
        red[3] = fma(aj, p1, red[3]);
    }
    block_reduce<4, false>(red, s_red);                              // (also publishes s_p / s_xy)
    const double S1_0 = red[0], S1_1 = red[1], S2_0 = red[2], S2_1 = red[3];

    // ---- (A' p)_j = sum_k |n_j - n_k| p_k over the pixel's nodes 1..M, both systems at once
    double acc0[ROWS], acc1[ROWS];
#pragma unroll
    for (int i = 0; i < ROWS; ++i) acc0[i] = acc1[i] = 0.0;
    // Where a step's time goes (ncu + timing experiments at n = 500, 16 384 pixels, 1.73 ms per step): the sweep
    // is 88 % of the instructions and issue-bound (stall = not selected; 1.19 ms at one instruction per clock per
    // scheduler, the same as its SFU time at the 78 % the evaluation kernel reaches); everything else alone takes
    // 0.26 ms (the step without the sweep: 3.4 TB/s of vector traffic).  Without the square root 1.57 ms, with
    // one system only 1.56 ms, packed FADD2 / FMUL2 / FFMA2 for the float32 part: no change (they issue at half
    // rate); 5 or 6 CTAs per SM instead of 4: -1.6 %.
#pragma unroll 4
    for (int k = 0; k < M; ++k) {
        const float2 xy = s_xy[k];
        const double2 pk = s_p[k];
#pragma unroll
        for (int i = 0; i < ROWS; ++i) {
            const double r = widen_pos(dist_f32(xj[i], yj[i], xy.x, xy.y));
            acc0[i] = fma(r, pk.x, acc0[i]);
            acc1[i] = fma(r, pk.y, acc1[i]);
        }
    }
    // q = B p = a (1.p) + (a.p) 1 - A' p ; alpha = rz / p.q.  a, p and r are re-derived here (node 0 distance,
    // shared memory, L2) instead of being kept in registers across the sweep: the sweep then needs 40 registers
    // and six CTAs fit on an SM, whose load / reduction phases hide behind one another's sweeps.
    double q0[ROWS], q1[ROWS], pa[ROWS], pb[ROWS];
    red[0] = red[1] = 0.0;
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
        const int j = t + i * PCG_THREADS;
        q0[i] = q1[i] = pa[i] = pb[i] = 0.0;
        if (j < M) {
            const double aj = widen_pos(dist_f32(xj[i], yj[i], x0, y0));
            const double2 pj = s_p[j];
            pa[i] = pj.x;
            pb[i] = pj.y;
            q0[i] = fma(aj, S1_0, S2_0) - acc0[i];
            q1[i] = fma(aj, S1_1, S2_1) - acc1[i];
        }
        red[0] = fma(pa[i], q0[i], red[0]);
        red[1] = fma(pb[i], q1[i], red[1]);
    }
    block_reduce<2, false>(reinterpret_cast<double(&)[2]>(red), s_red);
    const double al0 = (red[0] > 0.0) ? rz0 / red[0] : 0.0;
    const double al1 = (red[1] > 0.0) ? rz1 / red[1] : 0.0;
    red[0] = red[1] = 0.0;
    double r0[ROWS], r1[ROWS];
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
        const int j = t + i * PCG_THREADS;
        r0[i] = r1[i] = 0.0;
        if (j < M) {
            st.X[col0 * r_pad + j] = fma(al0, pa[i], st.X[col0 * r_pad + j]);
            st.X[col1 * r_pad + j] = fma(al1, pb[i], st.X[col1 * r_pad + j]);
            r0[i] = fma(-al0, q0[i], st.R[col0 * r_pad + j]);
            r1[i] = fma(-al1, q1[i], st.R[col1 * r_pad + j]);
            st.R[col0 * r_pad + j] = r0[i];
            st.R[col1 * r_pad + j] = r1[i];
        }
        red[0] = fmax(red[0], fabs(r0[i]));
        red[1] = fmax(red[1], fabs(r1[i]));
    }
    block_reduce<2, true>(reinterpret_cast<double(&)[2]>(red), s_red);
    const bool conv = red[0] < tol0 && red[1] < tol1;
    if (!conv) {
        double sc0, sc1;
        const float nz0 = pack_scale(red[0], &sc0), nz1 = pack_scale(red[1], &sc1);
#pragma unroll
        for (int i = 0; i < ROWS; ++i) {
            const int j = t + i * PCG_THREADS;
            if (j < g.k_pad) {
                pack_store(st.packed, col0, k_rows, g.k_pad, j, r0[i] * sc0);
                pack_store(st.packed, col1, k_rows, g.k_pad, j, r1[i] * sc1);
            }
        }
        if (t == 0) {
            st.zscale[col0] = nz0;
            st.zscale[col1] = nz1;
        }
    }
    if (t == 0) {
        st.rz[col0] = rz0;
        st.rz[col1] = rz1;
        st.iters[pidx] = iter + 1;
        if (conv) {
            st.done[pidx] = 1;
            atomicSub(st.tile_active + tb, 1);
        }
    }
}

// weights from the two solutions: sigma = (d_0 - a.p) / (a.q), w' = p + sigma q, w_0 = sigma - sum(w')
__global__ void __launch_bounds__(PCG_THREADS)
rbf_pcg_finish_kernel(const float *__restrict__ lx, const float *__restrict__ ly, const int32_t *__restrict__ inten,
                      PcgGeom g, PcgState st, double *__restrict__ weights, int32_t *__restrict__ iters_out,
                      int32_t *__restrict__ bad, double wnorm_limit) {
    __shared__ double s_red[PCG_THREADS / 32 * 4];
    const int slot = blockIdx.x, tb = blockIdx.y, t = threadIdx.x;
    const long long pidx = (long long)tb * PTILE + slot;
    long long pix;
    if (!tile_pixel(g, tb, slot, &pix)) return;
    const int n = g.n, M = g.m, r_pad = g.r_pad;
    const long long col0 = (long long)tb * 2 * PTILE + slot, col1 = col0 + PTILE;
    const float x0 = lx[pix * n], y0 = ly[pix * n];
    double red[2] = {0.0, 0.0};
    double xa[4], xb[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int j = t + i * PCG_THREADS;
        xa[i] = xb[i] = 0.0;
        if (j < M) {
            const double a = widen_pos(dist_f32(lx[pix * n + j + 1], ly[pix * n + j + 1], x0, y0));
            xa[i] = st.X[col0 * r_pad + j];
            xb[i] = st.X[col1 * r_pad + j];
            red[0] = fma(a, xa[i], red[0]);
            red[1] = fma(a, xb[i], red[1]);
        }
    }
    block_reduce<2, false>(red, s_red);
    const double sigma = ((double)inten[pix * n] - red[0]) / red[1];
    double sw[2] = {0.0, 0.0};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int j = t + i * PCG_THREADS;
        if (j < M) {
            const double w = fma(sigma, xb[i], xa[i]);
            weights[pix * n + j + 1] = w;
            sw[0] += w;
            sw[1] = fma(w, w, sw[1]);
        }
    }
    block_reduce<2, false>(sw, s_red);
    if (t == 0) {
        const double w0 = sigma - sw[0];
        weights[pix * n] = w0;
        // The matrix the iteration solved is the float32-rounded one: against the exact matrix the node residual
        // is ~1.2e-7 |w|_2 (measured over whole slabs: median 0.013, but 0.38 for the one pixel in 10^4 whose two
        // closest nodes are 1.6e-5 apart and whose weights reach 4e6 -- 1.3e-3 of the grid maximum).  Such pixels
        // are reported like unconverged ones: the caller refits AND evaluates them in fp64.
        const bool imprecise = !(sqrt(fma(w0, w0, sw[1])) <= wnorm_limit);
        const bool unconverged = st.done[pidx] == 0 || imprecise;             // not below tol within max_iters
        if (iters_out != nullptr) iters_out[pix] = unconverged ? -st.iters[pidx] : st.iters[pidx];
        if (unconverged && bad != nullptr) atomicAdd(bad + 1, 1);
    }
}

static inline long long pcg_round_up(long long v, long long m) { return (v + m - 1) / m * m; }

struct PcgLayout {
    int k_pad, r_pad, inv_ctas;
    size_t off_hi, off_lo, off_scale, off_packed, off_Z, off_X, off_R, off_P, off_rz, off_zs, off_done, off_iters,
        off_active, off_inv, inv_stride, total;
};

static PcgLayout pcg_layout(int n, long long tiles) {
    PcgLayout L;
    const int m = n - 1;
    L.k_pad = (int)pcg_round_up(m, 64);
    L.r_pad = (int)pcg_round_up(m, 128);
    L.inv_ctas = (int)(tiles < 296 ? tiles : 296);
    const size_t cols = (size_t)tiles * 2 * PTILE;
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t at = o; o += (bytes + 255) / 256 * 256; return at; };
    L.off_hi = take((size_t)tiles * L.r_pad * L.k_pad * 2);
    L.off_lo = take((size_t)tiles * L.r_pad * L.k_pad * 2);
    L.off_scale = take((size_t)tiles * L.r_pad * 4);
    L.off_packed = take(cols * 2 * L.k_pad * 2);
    L.off_Z = take(cols * L.r_pad * 8);
    L.off_X = take(cols * L.r_pad * 8);
    L.off_R = take(cols * L.r_pad * 8);
    L.off_P = take(cols * L.r_pad * 4);
    L.off_rz = take(cols * 8);
    L.off_zs = take(cols * 4);
    L.off_done = take((size_t)tiles * PTILE * 4);
    L.off_iters = take((size_t)tiles * PTILE * 4);
    L.off_active = take((size_t)tiles * 4);
    const size_t nT = L.k_pad / TS, nKC = L.k_pad / KC;
    L.inv_stride = 2 * nT * nKC * BLK + nT * TS * TS + (size_t)L.k_pad * L.k_pad;       // doubles
    L.off_inv = take((size_t)L.inv_ctas * L.inv_stride * 8);
    L.total = o;
    return L;
}

}  // namespace rti

using namespace rti;

extern "C" int64_t rti_rbf_pcg_workspace_bytes(int n, int64_t tiles) {
    if (n < 3 || n > 1025 || tiles < 1) return 0;
    return (int64_t)pcg_layout(n, tiles).total;
}

extern "C" int rti_rbf_fit_pcg(const float *d_lx, const float *d_ly, const int32_t *d_inten, int rows, int cols, int n,
                               double *d_weights, void *d_work, int64_t work_bytes, int max_iters, double tol,
                               int32_t *d_iters, int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_inten && d_weights && d_work, "rti_rbf_fit_pcg: NULL pointer");
    RTI_REQUIRE(rows > 0 && cols > 0, "rti_rbf_fit_pcg: empty input");
    RTI_REQUIRE(n >= 3 && n <= 1025, "rti_rbf_fit_pcg: 3 <= n <= 1025 nodes supported, got %d", n);
    RTI_REQUIRE(max_iters >= 1 && tol > 0.0, "rti_rbf_fit_pcg: max_iters >= 1 and tol > 0 required");
    RTI_REQUIRE((reinterpret_cast<uintptr_t>(d_work) & 255) == 0, "rti_rbf_fit_pcg: workspace not 256-byte aligned");
    RTI_REQUIRE(work_bytes >= rti_rbf_pcg_workspace_bytes(n, 1), "rti_rbf_fit_pcg: workspace below one tile (%lld bytes)",
                (long long)rti_rbf_pcg_workspace_bytes(n, 1));
    cudaStream_t s = as_stream(stream);
    const int tiles_x = (cols + PT - 1) / PT, tiles_y = (rows + PT - 1) / PT;
    const long long tiles_total = (long long)tiles_x * tiles_y;
    // largest batch of tiles the workspace holds (the layout grows monotonically with the tile count)
    long long lo = 1, hi = tiles_total < 65535 ? tiles_total : 65535;
    while (lo < hi) {
        const long long mid = (lo + hi + 1) / 2;
        if ((int64_t)pcg_layout(n, mid).total <= work_bytes) lo = mid;
        else hi = mid - 1;
    }
    const long long batch = lo;
    uint8_t *w = static_cast<uint8_t *>(d_work);
    const size_t inv_smem = (size_t)(4 * STG + TS * LDS + TS + 8 + 3 * n) * sizeof(double);
    RTI_REQUIRE(inv_smem <= 227 * 1024, "rti_rbf_fit_pcg: n = %d needs %zu bytes of shared memory", n, inv_smem);
    RTI_CUDA(cudaFuncSetAttribute(rbf_tile_inverse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)inv_smem));

    for (long long t0 = 0; t0 < tiles_total; t0 += batch) {
        const int T = (int)((tiles_total - t0) < batch ? (tiles_total - t0) : batch);
        const PcgLayout L = pcg_layout(n, T);
        PcgGeom g;
        g.rows = rows; g.cols = cols; g.n = n; g.tiles_x = tiles_x; g.tile0 = t0; g.n_tiles = T;
        g.m = n - 1; g.k_pad = L.k_pad; g.r_pad = L.r_pad;
        PcgState st;
        st.X = reinterpret_cast<double *>(w + L.off_X);
        st.R = reinterpret_cast<double *>(w + L.off_R);
        st.Z = reinterpret_cast<double *>(w + L.off_Z);
        st.rz = reinterpret_cast<double *>(w + L.off_rz);
        st.Pv = reinterpret_cast<float *>(w + L.off_P);
        st.zscale = reinterpret_cast<float *>(w + L.off_zs);
        st.packed = reinterpret_cast<uint16_t *>(w + L.off_packed);
        st.done = reinterpret_cast<int32_t *>(w + L.off_done);
        st.iters = reinterpret_cast<int32_t *>(w + L.off_iters);
        st.tile_active = reinterpret_cast<int32_t *>(w + L.off_active);
        uint16_t *op_hi = reinterpret_cast<uint16_t *>(w + L.off_hi), *op_lo = reinterpret_cast<uint16_t *>(w + L.off_lo);
        float *inv_scale = reinterpret_cast<float *>(w + L.off_scale);

        RTI_CUDA(cudaMemsetAsync(st.tile_active, 0, (size_t)T * 4, s));
        rbf_tile_inverse_kernel<<<L.inv_ctas, RBF_THREADS, inv_smem, s>>>(
            d_lx, d_ly, g, reinterpret_cast<double *>(w + L.off_inv), (int64_t)L.inv_stride, op_hi, op_lo, inv_scale, d_bad);
        RTI_LAUNCHED("rbf_tile_inverse_kernel");
        rbf_pcg_init_kernel<<<dim3(PTILE, T), PCG_THREADS, 0, s>>>(d_lx, d_ly, d_inten, g, st);
        RTI_LAUNCHED("rbf_pcg_init_kernel");
        const int rows_per_thread = (g.m + PCG_THREADS - 1) / PCG_THREADS;
        const size_t step_smem = (size_t)L.r_pad * (sizeof(double2) + sizeof(float2)) + 8 * 4 * sizeof(double);
        const double tol1 = tol / 1024.0;            // the a-system's solution is multiplied by sigma ~ 1e2
        for (int it = 0; it < max_iters; ++it) {
            int rc = rbf_gemm_batched(st.packed, 2 * L.k_pad, (long long)T * 2 * PTILE, op_hi, op_lo, inv_scale, T, L.r_pad,
                                      L.k_pad, 2, st.Z, st.tile_active, s);
            if (rc != RTI_OK) return rc;
#define RTI_PCG_STEP(R)                                                                                            \
    rbf_pcg_step_kernel<R, MB><<<dim3(PTILE, T), PCG_THREADS, step_smem, s>>>(d_lx, d_ly, g, st, it, tol, tol1)
            switch (rows_per_thread) {
            case 1: { constexpr int MB = 6; RTI_PCG_STEP(1); } break;
            case 2: { constexpr int MB = 6; RTI_PCG_STEP(2); } break;
            case 3: { constexpr int MB = 3; RTI_PCG_STEP(3); } break;
            default: { constexpr int MB = 2; RTI_PCG_STEP(4); } break;
            }
#undef RTI_PCG_STEP
            RTI_LAUNCHED("rbf_pcg_step_kernel");
        }
        // |w|_2 beyond which the float32 distances cost more than 10 tol of node residual (tol = 0.01: 0.1 intensity
        // units, i.e. grid values ~3.5e-4 of their maximum off from the fit and as much again from the float32
        // evaluation; measured: grid error / maximum ~ 3e-3 x 1.2e-7 |w|_2, tools/rbf_tail_study.py)
        const double wnorm_limit = 10.0 * tol / 1.2e-7;
        rbf_pcg_finish_kernel<<<dim3(PTILE, T), PCG_THREADS, 0, s>>>(d_lx, d_ly, d_inten, g, st, d_weights, d_iters, d_bad,
                                                                     wnorm_limit);
        RTI_LAUNCHED("rbf_pcg_finish_kernel");
    }
    return RTI_OK;
}
