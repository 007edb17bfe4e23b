// Linear radial-basis-function path (reference: _interpolate_RBF, analysis.py:249-260,
// i.e. scipy.interpolate.Rbf(lx, ly, values, function='linear') evaluated on the grid).
//
// Fit.  scipy solves A w = d with A[j][k] = |n_j - n_k| (symmetric, zero diagonal,
// indefinite -> pivoted solver).  phi(r) = r is conditionally negative definite, so
// eliminating node 0 gives the (n-1) x (n-1) matrix
//     B[j][k] = a_j + a_k - |n_j - n_k|,   a_j = |n_j - n_0|          (j, k = 1..n-1)
// which is symmetric POSITIVE definite for distinct nodes: Cholesky, no pivoting.
// With B p = d_0*1 - d', B q = a, sigma = (d_0 - a.p)/(a.q):
//     w' = p + sigma*q,   w_0 = sigma - sum(w')
// is exactly the solution of A w = d.  Everything is fp64.
//
// One CTA factorises one pixel's matrix with a left-looking blocked Cholesky on 64 x 64
// tiles.  B is generated on the fly from the node coordinates (never stored); L lives in
// an L2-resident workspace in 16 KB blocks laid out [k][row] so that every operand chunk
// is one contiguous cp.async stream into shared memory.  The two right-hand sides ride
// along as two extra rows of the matrix, which performs the forward substitution inside
// the factorisation; only the backward substitution remains.
#include <mutex>
#include "rbf_tiles.cuh"

namespace rti {

// Shared memory map (doubles):  [0, 2*STG) sA (2 stages)   [2*STG, 4*STG) sB (2 stages)
//   the potrf scratch sD[64][65] and the transposed C tile sCk[64][LDS] alias sA/sB, as do the
//   substitution vectors;  [4*STG, 4*STG + 64*LDS) sW = D^-T of the current block column
//   then sRsq[64], s_red[8], nodes (4 * n).  102 KB at n = 1000: two CTAs per SM.
__global__ void __launch_bounds__(RBF_THREADS, 2)
rbf_fit_kernel(const float *__restrict__ lx, const float *__restrict__ ly, const int32_t *__restrict__ inten,
               int64_t n_pixels, int n, double *__restrict__ weights, double *__restrict__ work,
               int64_t work_stride, int32_t *__restrict__ bad) {
    extern __shared__ __align__(16) double smem[];
    double *sA = smem;
    double *sB = smem + 2 * STG;
    double *sD = smem;                          // alias: [64][65]
    double *sCk = smem;                         // alias: [64 k][LDS]
    double *sW = smem + 4 * STG;                // [64 k][LDS]: sW[k][c] = Dinv[c][k]
    double *sRsq = sW + TS * LDS;
    double *s_red = sRsq + TS;
    double *sx = s_red + 8;
    double *sy = sx + n;
    double *sa = sy + n;
    double *sd = sa + n;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int rw = 16 * (warp >> 1);                            // warp patch origin inside a 64x64 tile
    const int cw = 32 * (warp & 1);
    const int er = rw + (lane >> 2);                            // element (mi, ni, e): row er + 8*mi,
    const int ec = cw + 2 * (lane & 3);                         //                      col ec + 8*ni + e
    const int M = n - 1;
    const int mpad = rbf_mpad(n);
    const int nT = mpad / TS, nKC = mpad / KC;
    double *L = work + (int64_t)blockIdx.x * work_stride;       // blocks (it, kc) at (it*nKC + kc)*BLK
    double *Wg = L + (int64_t)nT * nKC * BLK;                   // nT tiles of 64x64

    for (int64_t pix = blockIdx.x; pix < n_pixels; pix += gridDim.x) {
        __syncthreads();
        for (int i = t; i < n; i += RBF_THREADS) {
            sx[i] = (double)lx[pix * n + i];
            sy[i] = (double)ly[pix * n + i];
            sd[i] = (double)inten[pix * n + i];
        }
        __syncthreads();
        for (int i = t; i < n; i += RBF_THREADS) {
            const double dx = sx[i] - sx[0], dy = sy[i] - sy[0];
            sa[i] = sqrt_pos(fma(dx, dx, dy * dy));
        }
        __syncthreads();
        bool ok = true;

        for (int kb = 0; kb < nT; ++kb) {
            for (int it = kb; it < nT; ++it) {
                // ---- generate the tile of the bordered matrix in registers
                double acc[2][4][2];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi) {
                    const int gr = it * TS + er + 8 * mi;
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int gc = kb * TS + ec + 8 * ni + e;
                            double v;
                            if (gc >= M) v = (gr == gc) ? 1.0 : 0.0;                    // padding: identity
                            else if (gr < M) {
                                const double dx = sx[gr + 1] - sx[gc + 1], dy = sy[gr + 1] - sy[gc + 1];
                                v = sa[gr + 1] + sa[gc + 1] - sqrt_pos(fma(dx, dx, dy * dy));
                            } else if (gr == M) v = sd[0] - sd[gc + 1];                 // rhs 1: d0*1 - d'
                            else if (gr == M + 1) v = sa[gc + 1];                       // rhs 2: a
                            else v = 0.0;
                            acc[mi][ni][e] = v;
                        }
                }
                // ---- acc -= L[it][0:k0] . L[kb][0:k0]^T, operand chunks streamed with cp.async
                const int nk = kb * (TS / KS);
                const bool diag = (it == kb);
                // row block `it` of L is one contiguous [k][64] array of mpad k-rows
                const double *ga = L + (int64_t)it * nKC * BLK;
                const double *gb = L + (int64_t)kb * nKC * BLK;
                auto issue = [&](int buf, int kc) {
#pragma unroll
                    for (int i = 0; i < KS * TS * 8 / 16 / RBF_THREADS; ++i) {
                        const int ch = t + i * RBF_THREADS;              // 16-byte chunk: k-row ch/32, doubles 2*(ch%32)
                        const int so = (ch >> 5) * LDS + (ch & 31) * 2;
                        const int64_t go = (int64_t)kc * KS * TS + ch * 2;
                        cp_async16(sA + buf * STG + so, ga + go);
                        if (!diag) cp_async16(sB + buf * STG + so, gb + go);
                    }
                    cp_async_commit();
                };
                __syncthreads();                 // smem aliases of the previous tile are dead; L stores visible
                if (nk > 0) issue(0, 0);
                for (int kc = 0; kc < nk; ++kc) {
                    if (kc + 1 < nk) { issue((kc + 1) & 1, kc + 1); cp_async_wait<1>(); }
                    else cp_async_wait<0>();
                    __syncthreads();
                    const double *pa = sA + (kc & 1) * STG;
                    mma_chunk<LDS>(acc, pa, diag ? pa : sB + (kc & 1) * STG, rw, cw, lane, KS);
                    __syncthreads();
                }

                if (diag) {
                    // ---- potrf of the 64x64 diagonal tile in shared memory
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
                    // Blocked right-looking Cholesky, four 16-column panels.  Per panel: warp 0
                    // factors the 16x16 diagonal block (warp barriers only), one thread per row
                    // below it does the forward substitution of the panel, then all threads apply
                    // the rank-16 update to the trailing lower triangle in 4x4 register tiles.
                    // Columns >= cl are padding (identity) and are never touched.
                    const int cl = min(TS, M - kb * TS);
                    for (int jb = 0; jb < TS / 16; ++jb) {
                        const int o = 16 * jb;
                        __syncthreads();
                        if (warp == 0) {
                            for (int j = 0; j < 16; ++j) {
                                const int cj = o + j;
                                const double piv = sD[cj * SD_LD + cj];
                                // relative pivot test against the original diagonal 2*a_j (duplicate nodes -> 0)
                                const double ref = (cj < cl) ? 2.0 * sa[kb * TS + cj + 1] : 1.0;
                                if (!(piv > 1e-13 * ref) || !(piv > 0.0)) ok = false;
                                const double rs = rsqrt(piv);
                                __syncwarp();
                                const int r = cj + 1 + lane;
                                if (lane == 0) { sD[cj * SD_LD + cj] = piv * rs; sRsq[cj] = rs; }
                                if (r < o + 16) sD[r * SD_LD + cj] *= rs;
                                __syncwarp();
                                // trailing update inside the block: (r2, c) with cj < c <= r2 < o + 16
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
                        // panel: X[r][o..o+15] = S[r][o..o+15] . L_jj^-T, one thread per row r >= o + 16
                        if (t < TS - o - 16) {
                            const int r = o + 16 + t;
                            double x[16];
#pragma unroll
                            for (int c = 0; c < 16; ++c) x[c] = sD[r * SD_LD + o + c];
#pragma unroll
                            for (int pp = 0; pp < 16; ++pp) {
                                if (o + pp < cl) {                          // padding columns keep their identity entries
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
                        // trailing update: S[r][c] -= sum_p X[r][p] X[c][p] for o + 16 <= c <= r < 64
                        {
                            const int tt = (TS - o - 16) / 4;                // micro tiles per side: 12, 8, 4
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
                    // zero the strict upper triangle (the staged tile was the full symmetric matrix)
                    for (int e = t; e < TS * TS; e += RBF_THREADS) {
                        const int r = e >> 6, c = e & 63;
                        if (c > r) sD[r * SD_LD + c] = 0.0;
                    }
                    __syncthreads();
                    // ---- Dinv by forward substitution, 4 threads per column k: sW[k][c] = Dinv[c][k]
                    {
                        const int k = t >> 2, q = t & 3;
                        for (int c = 0; c < TS; ++c) {
                            double part = 0.0;
                            for (int p = k + q; p < c; p += 4) part = fma(sD[c * SD_LD + p], sW[k * LDS + p], part);
                            part += __shfl_xor_sync(0xffffffffu, part, 1);
                            part += __shfl_xor_sync(0xffffffffu, part, 2);
                            double v;
                            if (c < k) v = 0.0;
                            else if (c == k) v = sRsq[c];                            // 1/D[k][k]
                            else v = -part * sRsq[c];                               // 1/D[c][c] = rsqrt(piv_c)
                            if (q == 0) sW[k * LDS + c] = v;
                            __syncwarp();
                        }
                    }
                    __syncthreads();
                    // ---- store D (block layout [k=c][r]) and D^-T
                    double *gd = L + ((int64_t)kb * nKC + 2 * kb) * BLK;
                    for (int e = t; e < TS * TS; e += RBF_THREADS) {
                        const int c = e >> 6, r = e & 63;
                        gd[c * TS + r] = sD[r * SD_LD + c];                          // chunks 2kb, 2kb+1 are adjacent
                        Wg[(int64_t)kb * TS * TS + e] = sW[c * LDS + r];
                    }
                } else {
                    // ---- X = C . D^-T through the same micro-kernel: A operand = C^T staged [k=c][r]
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int r = er + 8 * mi, c = ec + 8 * ni + e;
                                const int gc = kb * TS + c, gr = it * TS + r;
                                sCk[c * LDS + r] = (gc >= M) ? ((gr == gc) ? 1.0 : 0.0) : acc[mi][ni][e];
                            }
                    __syncthreads();
                    double x[2][4][2];
#pragma unroll
                    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni) x[mi][ni][0] = x[mi][ni][1] = 0.0;
                    mma_chunk<LDS>(x, sCk, sW, rw, cw, lane, TS);                    // x = -C.D^-T
                    double *gx = L + ((int64_t)it * nKC + 2 * kb) * BLK;             // [k = c][r], 64 k-rows
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
        __syncthreads();

        // ---- backward substitution L^T x = z for both right-hand sides (z = rows M, M+1 of L)
        double *x1 = smem, *x2 = smem + mpad, *t1 = smem + 2 * mpad, *t2 = smem + 3 * mpad;   // alias sA/sB
        auto Lat = [&](int r, int c) -> double {
            return L[((int64_t)(r >> 6) * nKC + (c >> 5)) * BLK + (c & 31) * TS + (r & 63)];
        };
        for (int c = t; c < mpad; c += RBF_THREADS) {
            t1[c] = (c < M) ? Lat(M, c) : 0.0;
            t2[c] = (c < M) ? Lat(M + 1, c) : 0.0;
            x1[c] = 0.0;
            x2[c] = 0.0;
        }
        __syncthreads();
        for (int kb = nT - 1; kb >= 0; --kb) {
            if (kb * TS >= M) continue;                       // tile of padding only
            // s[c] = sum_{r >= (kb+1)*64, r < M} L[r][c] x[r]; one warp per column, lanes stride rows
            for (int cl = warp; cl < TS; cl += RBF_THREADS / 32) {
                const int c = kb * TS + cl;
                double s1 = 0.0, s2 = 0.0;
                for (int r = (kb + 1) * TS + lane; r < M; r += 32) {
                    const double l = Lat(r, c);
                    s1 = fma(l, x1[r], s1);
                    s2 = fma(l, x2[r], s2);
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
                    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
                }
                if (lane == 0) { t1[c] -= s1; t2[c] -= s2; }
            }
            __syncthreads();
            // x_kb[c] = sum_{k >= c} Dinv[k][c] * t[k]  (Dinv[k][c] = W[c][k], W stored [c][k] k-fastest)
            if (t < 2 * TS) {
                const int cl = t & 63;
                const double *tv = (t < TS) ? t1 : t2;
                const double *wrow = Wg + (int64_t)kb * TS * TS + cl * TS;
                double s = 0.0;
                for (int k = cl; k < TS; ++k) s = fma(wrow[k], tv[kb * TS + k], s);
                ((t < TS) ? x1 : x2)[kb * TS + cl] = (kb * TS + cl < M) ? s : 0.0;
            }
            __syncthreads();
        }
        // ---- sigma and the weights
        double ap = 0.0, aq = 0.0;
        for (int j = t; j < M; j += RBF_THREADS) {
            ap = fma(sa[j + 1], x1[j], ap);
            aq = fma(sa[j + 1], x2[j], aq);
        }
        ap = block_sum(ap, s_red);
        aq = block_sum(aq, s_red);
        const double sigma = (sd[0] - ap) / aq;
        double sw = 0.0;
        for (int j = t; j < M; j += RBF_THREADS) {
            const double w = fma(sigma, x2[j], x1[j]);
            weights[pix * n + j + 1] = w;
            sw += w;
        }
        sw = block_sum(sw, s_red);
        if (t == 0) {
            weights[pix * n] = sigma - sw;
            if (!ok && bad != nullptr) atomicAdd(bad, 1);
        }
        // `ok` is per thread; combine so that any thread's failure is reported once
        const int any_bad = __syncthreads_or(!ok);
        if (t == 0 && any_bad && ok && bad != nullptr) atomicAdd(bad, 1);
    }
}

// --------------------------------------------------------------------------------------
// Grid evaluation: value(v, u) = sum_k |(axis[u], axis[v]) - n_k| * w_k  (Rbf.__call__).
// A CTA owns PB consecutive pixels and EV_THREADS strips of EV_U consecutive directions of
// one grid row; a thread owns one strip, so (gy - y_k)^2 is shared by its EV_U directions
// and a node costs 3 broadcast LDS per 2 + 6*EV_U fp64 instructions.  Per term:
//     dx = gx - x_k;  d2 = dx*dx + dy2;  r = MUFU.RSQ64H(d2) (2^-22.9);  t = d2*r;
//     q = 3 - t*r;    acc += (t*q) * (w_k/2)          [= w_k sqrt(d2) (1 - 1.5 delta^2)]
// i.e. 6 fp64 issue slots and one MUFU, relative error 2.4e-14 per distance.  dy2 carries
// +1e-300 so that a grid point that coincides with a node gives t = 1e-150 instead of 0*inf.
// Table stores are PB*4 contiguous bytes per direction (one 32 B sector for PB = 8); the
// kernel is fp64-pipe bound: the table write is 4 B per N*6.5 fp64 slots.
// --------------------------------------------------------------------------------------
constexpr int EV_THREADS = 256;
constexpr int EV_KCH = 128;             // nodes staged per chunk
constexpr int EV_U = 4;                 // directions per thread (one strip of a grid row)

__device__ __forceinline__ double rsq64h(double d2) {
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d2));
    return r;
}

template <int PB, bool TABLE>
__global__ void __launch_bounds__(EV_THREADS)
rbf_eval_kernel(const float *__restrict__ lx, const float *__restrict__ ly, const double *__restrict__ weights,
                int64_t n_pixels, int n, const double *__restrict__ axis, int g, int strips_per_row,
                void *__restrict__ out, int64_t image_stride, int32_t *__restrict__ bad) {
    __shared__ double s_x[PB][EV_KCH], s_y[PB][EV_KCH], s_w[PB][EV_KCH];
    const int64_t p0 = (int64_t)blockIdx.x * PB;
    const int strip = blockIdx.y * EV_THREADS + threadIdx.x;
    const bool live = strip < strips_per_row * g;
    const int v = live ? strip / strips_per_row : 0;
    const int u0 = live ? (strip - v * strips_per_row) * EV_U : 0;
    double gx[EV_U];
#pragma unroll
    for (int j = 0; j < EV_U; ++j) gx[j] = axis[min(u0 + j, g - 1)];
    const double gy = axis[v];
    double acc[PB][EV_U];
#pragma unroll
    for (int q = 0; q < PB; ++q)
#pragma unroll
        for (int j = 0; j < EV_U; ++j) acc[q][j] = 0.0;

    for (int k0 = 0; k0 < n; k0 += EV_KCH) {
        const int kc = min(EV_KCH, n - k0);
        __syncthreads();
        for (int e = threadIdx.x; e < PB * kc; e += EV_THREADS) {
            const int q = e / kc, k = e - q * kc;
            const int64_t p = (p0 + q < n_pixels) ? p0 + q : n_pixels - 1;
            s_x[q][k] = (double)lx[p * n + k0 + k];
            s_y[q][k] = (double)ly[p * n + k0 + k];
            s_w[q][k] = 0.5 * weights[p * n + k0 + k];
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < PB; ++q) {
#pragma unroll 2
            for (int k = 0; k < kc; ++k) {
                const double xk = s_x[q][k], wk = s_w[q][k];
                const double dy = gy - s_y[q][k];
                const double dy2 = fma(dy, dy, 1e-300);
#pragma unroll
                for (int j = 0; j < EV_U; ++j) {
                    const double dx = gx[j] - xk;
                    const double d2 = fma(dx, dx, dy2);
                    const double r = rsq64h(d2);
                    const double t = d2 * r;
                    const double c = fma(-t, r, 3.0);
                    acc[q][j] = fma(t * c, wk, acc[q][j]);
                }
            }
        }
    }
    if (!live) return;
    bool any_bad = false;
#pragma unroll
    for (int j = 0; j < EV_U; ++j) {
        if (u0 + j >= g) break;
        const int64_t d = (int64_t)v * g + u0 + j;
        if (TABLE) {
            int32_t *dst = reinterpret_cast<int32_t *>(out) + d * image_stride + p0;
            if (PB == 8 && p0 + PB <= n_pixels && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                int val[PB];
#pragma unroll
                for (int q = 0; q < PB; ++q) {
                    if (!(fabs(acc[q][j]) < 2147483648.0)) any_bad = true;
                    val[q] = trunc_i32(acc[q][j]);
                }
                __stcs(reinterpret_cast<int4 *>(dst), make_int4(val[0], val[1], val[2], val[3]));
                __stcs(reinterpret_cast<int4 *>(dst) + 1, make_int4(val[4 % PB], val[5 % PB], val[6 % PB], val[7 % PB]));
            } else {
#pragma unroll
                for (int q = 0; q < PB; ++q) {
                    if (p0 + q < n_pixels) {
                        if (!(fabs(acc[q][j]) < 2147483648.0)) any_bad = true;
                        dst[q] = trunc_i32(acc[q][j]);
                    }
                }
            }
        } else {
            double *dst = reinterpret_cast<double *>(out);
#pragma unroll
            for (int q = 0; q < PB; ++q)
                if (p0 + q < n_pixels) dst[(p0 + q) * (int64_t)g * g + d] = acc[q][j];
        }
    }
    if (any_bad && bad != nullptr) atomicAdd(bad, 1);
}

// --------------------------------------------------------------------------------------
// Grid evaluation with float32 distances (rti_rbf_table_fast).  The north star allows 1e-3 on
// RBF grid values; the fp64 kernel above delivers 1e-8 and is bound by the fp64 pipe.  Here the
// distance |g - n_k| is formed in float32 (one FADD, one FFMA, one MUFU.SQRT: relative error
// ~1e-7) and only the product with the weight and the running sum stay in fp64 (one DFMA):
// the weights of a linear RBF are large and cancel (sum |w_k| ~ 1e6 for values ~ 1e2 at
// N = 500), so rounding THEM or the partial sums to float32 costs 1e-3 of the result while a
// 1e-7 relative error on each distance costs ~1e-5.  The float32 -> fp64 widening of r >= 0 is
// done with two shifts and an add on the integer pipe (exponent re-bias 127 -> 1023) because
// F2F.F64.F32 shares the XU pipe with MUFU, which is the unit this kernel is bound by:
// G*G*N square roots per pixel at 16 per clock per SM.
// A CTA owns PB consecutive pixels; a thread owns a small patch of the light grid (below).
// --------------------------------------------------------------------------------------
constexpr int FE_KCH = 128;             // nodes staged per chunk

__device__ __forceinline__ double widen_pos_f32(float r) {
    const uint32_t b = __float_as_uint(r);
    return __hiloint2double((int)((b >> 3) + 0x38000000u), (int)(b << 29));   // r = 0 maps to 2^-127
}

// A thread owns a patch of FE_V consecutive grid rows x FE_U consecutive columns for PB pixels, so dx is shared by
// the rows and dy^2 by the columns of its patch: one node costs 2 LDS + FE_U + 2 FE_V float instructions and
// FE_U * FE_V terms of 5 instructions each (FFMA, MUFU.SQRT, LEA.HI, IMAD.SHL, DFMA).
template <int PB, int FE_U, int FE_V>
__global__ void __launch_bounds__(EV_THREADS)
rbf_eval_f32_kernel(const float *__restrict__ lx, const float *__restrict__ ly, const double *__restrict__ weights,
                          int64_t n_pixels, int n, const double *__restrict__ axis, int g, int strips_per_row,
                          int row_groups, int32_t *__restrict__ out, int64_t image_stride, int32_t *__restrict__ bad) {
    __shared__ float2 s_xy[PB][FE_KCH];
    __shared__ double s_w[PB][FE_KCH];
    const int64_t p0 = (int64_t)blockIdx.x * PB;
    const int patch = blockIdx.y * EV_THREADS + threadIdx.x;
    const bool live = patch < strips_per_row * row_groups;
    const int vg = live ? patch / strips_per_row : 0;
    const int v0 = vg * FE_V;
    const int u0 = live ? (patch - vg * strips_per_row) * FE_U : 0;
    float gx[FE_U], gy[FE_V];
#pragma unroll
    for (int j = 0; j < FE_U; ++j) gx[j] = (float)axis[min(u0 + j, g - 1)];
#pragma unroll
    for (int i = 0; i < FE_V; ++i) gy[i] = (float)axis[min(v0 + i, g - 1)];
    double acc[PB][FE_V][FE_U];
#pragma unroll
    for (int q = 0; q < PB; ++q)
#pragma unroll
        for (int i = 0; i < FE_V; ++i)
#pragma unroll
            for (int j = 0; j < FE_U; ++j) acc[q][i][j] = 0.0;

    for (int k0 = 0; k0 < n; k0 += FE_KCH) {
        const int kc = min(FE_KCH, n - k0);
        __syncthreads();
        for (int e = threadIdx.x; e < PB * kc; e += EV_THREADS) {
            const int q = e / kc, k = e - q * kc;
            const int64_t p = (p0 + q < n_pixels) ? p0 + q : n_pixels - 1;
            s_xy[q][k] = make_float2(lx[p * n + k0 + k], ly[p * n + k0 + k]);
            s_w[q][k] = weights[p * n + k0 + k];
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < PB; ++q) {
#pragma unroll 2
            for (int k = 0; k < kc; ++k) {
                const float2 xy = s_xy[q][k];
                const double wk = s_w[q][k];
                float dx[FE_U], dy2[FE_V];
#pragma unroll
                for (int j = 0; j < FE_U; ++j) dx[j] = gx[j] - xy.x;
#pragma unroll
                for (int i = 0; i < FE_V; ++i) {
                    const float dy = gy[i] - xy.y;
                    dy2[i] = dy * dy;
                }
#pragma unroll
                for (int i = 0; i < FE_V; ++i)
#pragma unroll
                    for (int j = 0; j < FE_U; ++j) {
                        float r;
                        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(dx[j], dx[j], dy2[i])));
                        acc[q][i][j] = fma(widen_pos_f32(r), wk, acc[q][i][j]);
                    }
            }
        }
    }
    if (!live) return;
    bool any_bad = false;
#pragma unroll
    for (int i = 0; i < FE_V; ++i) {
        if (v0 + i >= g) break;
#pragma unroll
        for (int j = 0; j < FE_U; ++j) {
            if (u0 + j >= g) break;
            const int64_t d = (int64_t)(v0 + i) * g + u0 + j;
            int32_t *dst = out + d * image_stride + p0;
            if (PB % 4 == 0 && p0 + PB <= n_pixels && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                for (int q = 0; q < PB; q += 4) {
                    int val[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        if (!(fabs(acc[q + e][i][j]) < 2147483648.0)) any_bad = true;
                        val[e] = trunc_i32(acc[q + e][i][j]);
                    }
                    __stcs(reinterpret_cast<int4 *>(dst + q), make_int4(val[0], val[1], val[2], val[3]));
                }
            } else {
#pragma unroll
                for (int q = 0; q < PB; ++q) {
                    if (p0 + q < n_pixels) {
                        if (!(fabs(acc[q][i][j]) < 2147483648.0)) any_bad = true;
                        dst[q] = trunc_i32(acc[q][i][j]);
                    }
                }
            }
        }
    }
    if (any_bad && bad != nullptr) atomicAdd(bad, 1);
}

}  // namespace rti

using namespace rti;

static size_t rbf_fit_smem(int n) {
    return (size_t)(4 * STG + TS * LDS + TS + 8 + 4 * n) * sizeof(double);
}

extern "C" int64_t rti_rbf_workspace_bytes(int n, int batch) {
    if (n < 2 || batch < 1) return 0;
    const int64_t mpad = rbf_mpad(n);
    const int64_t nT = mpad / TS, nKC = mpad / KC;
    return (int64_t)batch * (nT * nKC * BLK + nT * TS * TS) * (int64_t)sizeof(double);
}

extern "C" int rti_rbf_fit_soa(const float *d_lx, const float *d_ly, const int32_t *d_inten, int64_t n_pixels,
                               int n, double *d_weights, void *d_work, int64_t work_bytes, int batch,
                               int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_inten && d_weights && d_work, "rti_rbf_fit_soa: NULL pointer");
    RTI_REQUIRE(n_pixels > 0 && batch > 0, "rti_rbf_fit_soa: empty input");
    RTI_REQUIRE(n >= 2, "rti_rbf_fit_soa: need >= 2 nodes, got %d", n);
    RTI_REQUIRE(work_bytes >= rti_rbf_workspace_bytes(n, batch), "rti_rbf_fit_soa: workspace too small");
    RTI_REQUIRE((reinterpret_cast<uintptr_t>(d_work) & 15) == 0, "rti_rbf_fit_soa: workspace not 16-byte aligned");
    const size_t smem = rbf_fit_smem(n);
    // the four substitution vectors of length mpad alias the operand stages and sW
    RTI_REQUIRE(4 * rbf_mpad(n) <= 4 * STG + TS * LDS, "rti_rbf_fit_soa: n = %d exceeds the supported %d nodes", n,
                (4 * STG + TS * LDS) / 4 - 1);
    RTI_REQUIRE(smem <= 227 * 1024, "rti_rbf_fit_soa: n = %d needs %zu bytes of shared memory (> 227 KB)", n, smem);
    RTI_CUDA(cudaFuncSetAttribute(rbf_fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t grid = n_pixels < batch ? n_pixels : batch;
    const int64_t stride = rti_rbf_workspace_bytes(n, 1) / (int64_t)sizeof(double);
    rbf_fit_kernel<<<(unsigned)grid, RBF_THREADS, smem, as_stream(stream)>>>(
        d_lx, d_ly, d_inten, n_pixels, n, d_weights, reinterpret_cast<double *>(d_work), stride, d_bad);
    RTI_LAUNCHED("rbf_fit_kernel");
    return RTI_OK;
}

// ---- the same fit for callers that may not allocate (no counterpart in the reference) ----------
// The L workspace lives in a file-scope device array instead of a caller-provided buffer: every
// GPU that loads the library holds RTI_RBF_STATIC_SCRATCH_BYTES until the process exits.  In host
// code the array's name is only a stand-in for the symbol, so its device address is looked up once
// per device (with that device current) and cached.  All calls share the array: calls that overlap
// on the GPU (different streams, or different host threads) race on it.  Nothing in it has to start
// at zero -- rbf_fit_kernel writes every L block and D^-T tile before it reads them, which is also
// what lets rti_rbf_fit_soa run on uninitialised workspaces -- so the contents the previous call
// left behind are simply overwritten.
constexpr int64_t RBF_STATIC_SCRATCH_BYTES = 32ll << 20;
static __device__ __align__(256) double g_rbf_static_scratch[RBF_STATIC_SCRATCH_BYTES / sizeof(double)];
static double *g_rbf_static_ptr[64] = {};
static std::mutex g_rbf_static_mutex;

extern "C" int rti_rbf_fit_soa_static(const float *d_lx, const float *d_ly, const int32_t *d_inten,
                                      int64_t n_pixels, int n, double *d_weights, int32_t *d_bad,
                                      rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_inten && d_weights, "rti_rbf_fit_soa_static: NULL pointer");
    RTI_REQUIRE(n_pixels > 0, "rti_rbf_fit_soa_static: empty input");
    RTI_REQUIRE(n >= 2, "rti_rbf_fit_soa_static: need >= 2 nodes, got %d", n);
    const int64_t per_pixel = rti_rbf_workspace_bytes(n, 1);
    RTI_REQUIRE(per_pixel <= RBF_STATIC_SCRATCH_BYTES,
                "rti_rbf_fit_soa_static: n = %d needs %lld bytes of scratch per pixel, the built-in scratch holds "
                "%lld (n <= 1983); use rti_rbf_fit_soa with a caller-provided workspace", n, (long long)per_pixel,
                (long long)RBF_STATIC_SCRATCH_BYTES);
    const size_t smem = rbf_fit_smem(n);
    RTI_REQUIRE(4 * rbf_mpad(n) <= 4 * STG + TS * LDS && smem <= 227 * 1024,
                "rti_rbf_fit_soa_static: n = %d exceeds what the kernel's shared memory holds", n);
    int dev = 0;
    RTI_CUDA(cudaGetDevice(&dev));
    RTI_REQUIRE(dev >= 0 && dev < 64, "rti_rbf_fit_soa_static: device ordinal %d out of range", dev);
    double *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_rbf_static_mutex);
        if (g_rbf_static_ptr[dev] == nullptr) {
            void *p = nullptr;
            RTI_CUDA(cudaGetSymbolAddress(&p, g_rbf_static_scratch));
            RTI_REQUIRE(p != nullptr, "rti_rbf_fit_soa_static: the scratch symbol has no address on device %d", dev);
            g_rbf_static_ptr[dev] = static_cast<double *>(p);
        }
        scratch = g_rbf_static_ptr[dev];
    }
    // pixels in flight: what the scratch holds, at most two CTAs per SM of a 148-SM part
    int64_t batch = RBF_STATIC_SCRATCH_BYTES / per_pixel;
    if (batch > 2 * 148) batch = 2 * 148;
    RTI_CUDA(cudaFuncSetAttribute(rbf_fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t grid = n_pixels < batch ? n_pixels : batch;
    const int64_t stride = per_pixel / (int64_t)sizeof(double);
    rbf_fit_kernel<<<(unsigned)grid, RBF_THREADS, smem, as_stream(stream)>>>(
        d_lx, d_ly, d_inten, n_pixels, n, d_weights, scratch, stride, d_bad);
    RTI_LAUNCHED("rbf_fit_kernel");
    return RTI_OK;
}

template <bool TABLE>
static int rbf_eval_launch(const float *d_lx, const float *d_ly, const double *d_weights, int64_t n_pixels, int n,
                           const double *d_axis, int g, void *d_out, int64_t image_stride, int32_t *d_bad,
                           rti_stream_t stream, const char *who) {
    RTI_REQUIRE(d_lx && d_ly && d_weights && d_axis && d_out, "%s: NULL pointer", who);
    RTI_REQUIRE(n_pixels > 0 && n > 0 && g > 0 && g <= 2048, "%s: bad sizes", who);
    const int strips_per_row = (g + EV_U - 1) / EV_U;
    const unsigned gy = (unsigned)(((int64_t)strips_per_row * g + EV_THREADS - 1) / EV_THREADS);
    if (TABLE) {
        RTI_REQUIRE(image_stride >= n_pixels, "%s: image_stride < pixels", who);
        const int64_t bx = (n_pixels + 7) / 8;
        RTI_REQUIRE(bx <= 0x7fffffffLL, "%s: too many pixels", who);
        rbf_eval_kernel<8, true><<<dim3((unsigned)bx, gy), EV_THREADS, 0, as_stream(stream)>>>(
            d_lx, d_ly, d_weights, n_pixels, n, d_axis, g, strips_per_row, d_out, image_stride, d_bad);
    } else {
        RTI_REQUIRE(n_pixels <= 0x7fffffffLL, "%s: too many pixels", who);
        rbf_eval_kernel<1, false><<<dim3((unsigned)n_pixels, gy), EV_THREADS, 0, as_stream(stream)>>>(
            d_lx, d_ly, d_weights, n_pixels, n, d_axis, g, strips_per_row, d_out, 0, d_bad);
    }
    RTI_LAUNCHED("rbf_eval_kernel");
    return RTI_OK;
}

extern "C" int rti_rbf_table(const float *d_lx, const float *d_ly, const double *d_weights, int64_t n_pixels, int n,
                             const double *d_axis, int g, int32_t *d_table, int64_t image_stride, int32_t *d_bad,
                             rti_stream_t stream) {
    return rbf_eval_launch<true>(d_lx, d_ly, d_weights, n_pixels, n, d_axis, g, d_table, image_stride, d_bad,
                                 stream, "rti_rbf_table");
}

extern "C" int rti_rbf_grids(const float *d_lx, const float *d_ly, const double *d_weights, int64_t n_pixels, int n,
                             const double *d_axis, int g, double *d_grids, rti_stream_t stream) {
    return rbf_eval_launch<false>(d_lx, d_ly, d_weights, n_pixels, n, d_axis, g, d_grids, 0, nullptr, stream,
                                  "rti_rbf_grids");
}

template <int PB, int U, int V>
static int rbf_fast_launch(const float *d_lx, const float *d_ly, const double *d_weights, int64_t n_pixels, int n,
                                 const double *d_axis, int g, int32_t *d_table, int64_t image_stride, int32_t *d_bad,
                                 rti_stream_t stream, const char *who) {
    const int strips_per_row = (g + U - 1) / U, row_groups = (g + V - 1) / V;
    const unsigned gy = (unsigned)(((int64_t)strips_per_row * row_groups + EV_THREADS - 1) / EV_THREADS);
    const int64_t bx = (n_pixels + PB - 1) / PB;
    RTI_REQUIRE(bx <= 0x7fffffffLL, "%s: too many pixels", who);
    rbf_eval_f32_kernel<PB, U, V><<<dim3((unsigned)bx, gy), EV_THREADS, 0, as_stream(stream)>>>(
        d_lx, d_ly, d_weights, n_pixels, n, d_axis, g, strips_per_row, row_groups, d_table, image_stride, d_bad);
    RTI_LAUNCHED("rbf_eval_f32_kernel");
    return RTI_OK;
}

extern "C" int rti_rbf_table_fast(const float *d_lx, const float *d_ly, const double *d_weights, int64_t n_pixels,
                                  int n, const double *d_axis, int g, int32_t *d_table, int64_t image_stride,
                                  int32_t *d_bad, rti_stream_t stream) {
    const char *who = "rti_rbf_table_fast";
    RTI_REQUIRE(d_lx && d_ly && d_weights && d_axis && d_table, "%s: NULL pointer", who);
    RTI_REQUIRE(n_pixels > 0 && n > 0 && g > 0 && g <= 2048, "%s: bad sizes", who);
    RTI_REQUIRE(image_stride >= n_pixels, "%s: image_stride < pixels", who);
#define RTI_FAST_ARGS d_lx, d_ly, d_weights, n_pixels, n, d_axis, g, d_table, image_stride, d_bad, stream, who
    static const int variant = getenv("RTI_RBF_FAST_VARIANT") ? atoi(getenv("RTI_RBF_FAST_VARIANT")) : 0;
    // development A/B (tools/prof_rbf.py, N = 500, 4736 pixels; <pixels, columns, rows> per thread):
    // <4,4,2> 6.19 ms = 0.82 of the SFU peak, <4,2,4> 6.19, <4,5,2> 6.36, <2,4,2> 6.52, <4,4,1> 6.55 (ncu of that
    // shape: XU pipe 80 %, issue slots 76 %, stall = MIO throttle, i.e. the MUFU queue is full), <4,3,2> 6.69,
    // <8,2,2> 6.95, <2,4,4> 7.66, <4,4,3> 8.33; 1-row strips <8,4> 6.94, <4,10> 7.14, <2,10> 6.62.  Also measured and
    // dropped: MUFU.RSQ + FMUL instead of MUFU.SQRT 7.30 ms; register caps for 4 / 5 CTAs per SM 7.13 / 7.47 ms
    switch (variant) {
    case 1: return rbf_fast_launch<4, 4, 1>(RTI_FAST_ARGS);
    case 2: return rbf_fast_launch<4, 2, 4>(RTI_FAST_ARGS);
    default: return rbf_fast_launch<4, 4, 2>(RTI_FAST_ARGS);
    }
#undef RTI_FAST_ARGS
}
