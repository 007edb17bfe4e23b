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
#include <stdlib.h>
#include <mutex>
#include "tma.cuh"

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


// ---------------------------------------------------------------------------------------------
// Warp-specialised persistent form.  One CTA of 768 threads per SM walks tiles of 256 pixels
// claimed from an atomic counter:
//   warps 0-7  ("fit"):   thread = pixel, produce the coefficients of the next tiles
//   warps 8-23 ("table"): stream the int32 table of the current tile (lane = 4 pixels)
// so the fp64/XU-bound fit and the HBM-bound stores overlap inside every SM all the time instead
// of only when two co-resident CTAs happen to be in different phases.  Under the table's write
// stream a global read takes thousands of cycles, far more than eight fit warps can hide with
// register prefetch, so the fit warps pull their frame bytes through a cp.async ring in shared
// memory ([2 stages][256 frames][256 pixels], the next chunk in flight, one barrier per chunk); the
// camera table (n x 24 B) is loaded once per CTA.  Coefficients pass through a three-deep smem
// ring guarded by mbarriers (full[b] / empty[b], one arrival per warp); named barrier 1 synchronises
// the fit warps alone.
constexpr int WS_PIX = 256;
constexpr int WS_TABLE_WARPS = 16;
constexpr int WS_TABLE_THREADS = 32 * WS_TABLE_WARPS;
constexpr int WS_THREADS = WS_PIX + WS_TABLE_THREADS;          // 768: 8 fit warps + 16 table warps
// register split (setmaxnreg, per warpgroup): 768 x 80 at launch = 256 x 112 (fit) + 512 x 64 (table)
#ifndef WS_CHUNK_FRAMES
#define WS_CHUNK_FRAMES 256       /* measured: 64x4 7.02 ms, 128x3 6.95, 256x2 6.91 at 1024^2 x 500: fewer, larger */
#define WS_RING_STAGES 2         /* read bursts disturb the table's write stream less                            */
#endif
constexpr int WS_CHUNK = WS_CHUNK_FRAMES;          // frames per ring stage
constexpr int WS_STAGES = WS_RING_STAGES;
#ifndef WS_UNROLL
#define WS_UNROLL 4
#endif
constexpr int WS_DEPTH = WS_STAGES - 1;            // chunks in flight
constexpr int WS_STAGE_BYTES = WS_CHUNK * WS_PIX;
constexpr int WS_MAX_LIGHTS = 4096;                // camera table resident in smem (96 KB at the cap)
#ifndef WS_COEF_BUFS
#define WS_COEF_BUFS 3
#endif
constexpr int WS_NBUF = WS_COEF_BUFS;              // coefficient tiles between the fit and the table warps
constexpr int WS_COEF_BYTES = WS_NBUF * 6 * WS_PIX * 8;
#ifndef WS_FIT_REGS
#define WS_FIT_REGS 112
#define WS_TABLE_REGS 64
#endif
#define WS_STR2(x) #x
#define WS_STR(x) WS_STR2(x)

__device__ __forceinline__ void bar_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void cp_async16_zfill(void *smem, const void *gmem, int src_bytes) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(sa), "l"(gmem), "r"(src_bytes) : "memory");
}

// Tile counters for the dynamic scheduler: a ring of 64 per device, one per launch, zeroed on the
// launch's stream (launches on different streams may overlap; 64 in flight at once would not).
constexpr int WS_COUNTERS = 64;
static unsigned int *g_tile_counters[64] = {};
static unsigned g_tile_counter_next[64] = {};
static std::mutex g_tile_counter_mutex;

static int claim_tile_counter(int dev, cudaStream_t stream, unsigned int **out) {
    RTI_REQUIRE(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
    unsigned slot;
    {
        std::lock_guard<std::mutex> lock(g_tile_counter_mutex);
        if (!g_tile_counters[dev]) RTI_CUDA(cudaMalloc(&g_tile_counters[dev], WS_COUNTERS * sizeof(unsigned int)));
        slot = g_tile_counter_next[dev]++ % WS_COUNTERS;
    }
    *out = g_tile_counters[dev] + slot;
    RTI_CUDA(cudaMemsetAsync(*out, 0, sizeof(unsigned int), stream));
    return RTI_OK;
}

void release_tile_counters() {
    std::lock_guard<std::mutex> lock(g_tile_counter_mutex);
    int cur = 0;
    cudaGetDevice(&cur);
    for (int d = 0; d < 64; ++d)
        if (g_tile_counters[d]) {
            cudaSetDevice(d);
            cudaFree(g_tile_counters[d]);
            g_tile_counters[d] = nullptr;
        }
    cudaSetDevice(cur);
}

constexpr size_t WS_SMEM_LIMIT = 227 * 1024 - 1024;      // dynamic smem the kernel may ask for (static part is < 1 KB)
static size_t ws_smem_bytes(int n, int g) {
    return (size_t)WS_STAGES * WS_STAGE_BYTES + WS_COEF_BYTES + ((size_t)3 * n + g) * sizeof(double);
}

template <bool VEC4>
__global__ void __launch_bounds__(WS_THREADS, 1)
ptm_fused_ws_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride, const double *__restrict__ cams,
                    int n, int row0, int cols, int64_t n_pixels, const double *__restrict__ axis, int g,
                    double *__restrict__ coef, int64_t plane_stride, int32_t *__restrict__ table,
                    int64_t image_stride, int32_t *__restrict__ bad, unsigned int *__restrict__ tile_counter) {
    extern __shared__ __align__(16) uint8_t s_raw[];
    uint8_t *s_ring = s_raw;                                                   // [stage][frame][pixel]
    double (*s_coef)[6][WS_PIX] = reinterpret_cast<double (*)[6][WS_PIX]>(s_raw + WS_STAGES * WS_STAGE_BYTES);
    double *s_cam = reinterpret_cast<double *>(s_raw + WS_STAGES * WS_STAGE_BYTES + WS_COEF_BYTES);   // [n][3]: x, y, z^2
    double *s_axis = s_cam + 3 * (size_t)n;
    // full[b]: one arrival per fit warp; empty[b]: one per table warp.  mbarriers instead of bar.sync so
    // that the table warps are not lock-stepped with each other at every tile.
    __shared__ __align__(8) uint64_t s_full[WS_NBUF], s_empty[WS_NBUF];
    // Tiles are claimed dynamically (atomic counter): SMs do not get equal shares of the HBM write
    // bandwidth, and with a static split the slowest SM set the kernel time (18 % idle SM cycles).
    __shared__ int64_t s_tileq[8];             // claimed tile ids, indexed by the CTA's tile count
    __shared__ int64_t s_buf_tile[WS_NBUF];    // tile held by each coefficient buffer, -1 = no more work
    if (threadIdx.x == 0) {
        for (int b = 0; b < WS_NBUF; ++b) {
            mbar_init(smem_u32(&s_full[b]), WS_PIX / 32);
            mbar_init(smem_u32(&s_empty[b]), WS_TABLE_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }

    const int64_t n_tiles = (n_pixels + WS_PIX - 1) / WS_PIX;
    for (int i = threadIdx.x; i < g; i += WS_THREADS) s_axis[i] = axis[i];
    for (int i = threadIdx.x; i < n; i += WS_THREADS) {
        const double cz = cams[(int64_t)i * 3 + 2];
        s_cam[3 * i + 0] = cams[(int64_t)i * 3 + 0];
        s_cam[3 * i + 1] = cams[(int64_t)i * 3 + 1];
        s_cam[3 * i + 2] = cz * cz;
    }
    __syncthreads();

    if (threadIdx.x < WS_PIX) {
        // ------------------------------------------------------------------ fit warps
        asm volatile("setmaxnreg.inc.sync.aligned.u32 " WS_STR(WS_FIT_REGS) ";\n");
        const int t = threadIdx.x;
        const int chunks_per_tile = (n + WS_CHUNK - 1) / WS_CHUNK;
        // the cp.async cursor runs WS_DEPTH chunks ahead, i.e. up to `ahead` tiles: thread 0 keeps that
        // many tile ids claimed beyond the one being fitted (published by the per-chunk barrier)
        const int ahead = (WS_DEPTH + chunks_per_tile - 1) / chunks_per_tile;
        if (t == 0)
            for (int i = 0; i <= ahead; ++i) s_tileq[i] = (int64_t)atomicAdd(tile_counter, 1u);
        bar_sync(1, WS_PIX);

        // producer cursor: chunk (ptile, pj) goes to ring stage pstage
        int pk = 0, pj = 0, pstage = 0;
        int64_t ptile = s_tileq[0];
        auto issue = [&]() {
            if (ptile < n_tiles) {
                const int f0 = pj * WS_CHUNK;
                const int fc = min(WS_CHUNK, n - f0);
                const int64_t px0 = ptile * WS_PIX;
                uint8_t *dst = s_ring + pstage * WS_STAGE_BYTES;
#pragma unroll
                for (int r = 0; r < WS_CHUNK * (WS_PIX / 16) / WS_PIX; ++r) {
                    const int piece = t + r * WS_PIX;
                    const int f = piece >> 4, seg = piece & 15;
                    const int64_t px = px0 + seg * 16;
                    const int64_t left = n_pixels - px;
                    if (f < fc && left > 0)
                        cp_async16_zfill(dst + f * WS_PIX + seg * 16, frames + (int64_t)(f0 + f) * frame_stride + px,
                                         left >= 16 ? 16 : (int)left);
                }
                if (++pj == chunks_per_tile) {
                    pj = 0;
                    ptile = s_tileq[++pk & 7];
                }
            }
            pstage = (pstage + 1 == WS_STAGES) ? 0 : pstage + 1;
            asm volatile("cp.async.commit_group;\n" ::: "memory");
        };
#pragma unroll
        for (int d = 0; d < WS_DEPTH; ++d) issue();

        int cstage = 0, b = 0, round = 0;                     // coefficient buffer b, its use count
        for (int k = 0;; ++k) {
            const int64_t tile = s_tileq[k & 7];
            if (tile >= n_tiles) break;
            if (t == 0) s_tileq[(k + ahead + 1) & 7] = (int64_t)atomicAdd(tile_counter, 1u);
            const int64_t p = tile * WS_PIX + t;
            const bool active = p < n_pixels;
            const int64_t pc = active ? p : n_pixels - 1;
            const int yl = (int)(pc / cols);
            const double xd = (double)(int)(pc - (int64_t)yl * cols);
            const double yd = (double)(row0 + yl);
            NormalEq ne;
            ne.clear();
            for (int j = 0; j < chunks_per_tile; ++j) {
                asm volatile("cp.async.wait_group %0;\n" ::"n"(WS_DEPTH - 1) : "memory");
                bar_sync(1, WS_PIX);          // chunk landed for everyone; stage of the previous chunk is free
                issue();                      // refill that stage, WS_DEPTH chunks ahead
                const int f0 = j * WS_CHUNK;
                const int fc = min(WS_CHUNK, n - f0);
                const uint8_t *src = s_ring + cstage * WS_STAGE_BYTES + t;
                const double *cam = s_cam + 3 * f0;
                int i = 0;
                for (; i + WS_UNROLL <= fc; i += WS_UNROLL) {
#pragma unroll
                    for (int k = 0; k < WS_UNROLL; ++k) {
                        float lu, lv;
                        light_vector(cam[3 * (i + k)], cam[3 * (i + k) + 1], cam[3 * (i + k) + 2], xd, yd, lu, lv);
                        ne.add(lu, lv, u8_to_double(src[(i + k) * WS_PIX]));
                    }
                }
                for (; i < fc; ++i) {
                    float lu, lv;
                    light_vector(cam[3 * i], cam[3 * i + 1], cam[3 * i + 2], xd, yd, lu, lv);
                    ne.add(lu, lv, u8_to_double(src[i * WS_PIX]));
                }
                cstage = (cstage + 1 == WS_STAGES) ? 0 : cstage + 1;
            }
            double a[6];
            const bool ok = solve_normal_eq(ne, (double)n, a);
            if (round > 0) mbar_wait(smem_u32(&s_empty[b]), (round - 1) & 1);   // previous table of buffer b is done
            double bound = 0.0;
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                s_coef[b][k][t] = a[k];
                bound += fabs(a[k]);
                if (active && coef != nullptr) coef[k * plane_stride + p] = a[k];
            }
            if (t == 0) s_buf_tile[b] = tile;
            __syncwarp();
            if ((t & 31) == 0) mbar_arrive(smem_u32(&s_full[b]));
            if (active && bad != nullptr) {
                if (!ok) atomicAdd(bad, 1);
                // |value| <= sum|a_k| on [-1,1]^2; NaN coefficients fail the comparison as well
                else if (!(bound < 2147483648.0)) atomicAdd(bad + 1, 1);
            }
            if (++b == WS_NBUF) {
                b = 0;
                ++round;
            }
        }
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        // tell the table warps there is nothing left
        if (round > 0) mbar_wait(smem_u32(&s_empty[b]), (round - 1) & 1);
        if (t == 0) s_buf_tile[b] = -1;
        __syncwarp();
        if ((t & 31) == 0) mbar_arrive(smem_u32(&s_full[b]));
    } else {
        // ------------------------------------------------------------------ table warps
        // warp w: pixel half (w / 8) of the tile, lane = 4 consecutive pixels (512 B per warp store);
        // the eight warps of a half interleave the grid rows (start rotated per tile so the 100 rows
        // balance).  Only a0 and the two per-row terms live in registers (56 per thread); the other
        // coefficients are re-read from the smem tile once per grid row, so the tile is handed back
        // to the fit warps when its table is finished.
        asm volatile("setmaxnreg.dec.sync.aligned.u32 " WS_STR(WS_TABLE_REGS) ";\n");
        const int t = threadIdx.x - WS_PIX;
        const int lane = t & 31, warp = t >> 5;
        const int q0 = 128 * (warp >> 3) + 4 * lane;
        int it = 0, b = 0, round = 0;
        for (;; ++it) {
            mbar_wait(smem_u32(&s_full[b]), round & 1);
            const int64_t tile = s_buf_tile[b];
            if (tile < 0) break;
            const int64_t p0 = tile * WS_PIX + q0;
            if (p0 < n_pixels) {
                const bool full = p0 + 4 <= n_pixels;
                double a0[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) a0[q] = s_coef[b][0][q0 + q];
                for (int v = (warp + it) & 7; v < g; v += 8) {
                    const double lv = s_axis[v];
                    double bv[4], cv[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        bv[q] = fma(s_coef[b][2][q0 + q], lv, s_coef[b][3][q0 + q]);
                        cv[q] = fma(fma(s_coef[b][1][q0 + q], lv, s_coef[b][4][q0 + q]), lv, s_coef[b][5][q0 + q]);
                    }
                    int32_t *dst = table + ((int64_t)v * g) * image_stride + p0;
#pragma unroll 4
                    for (int u = 0; u < g; ++u) {
                        const double lu = s_axis[u];
                        int val[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const double x = ptm_eval(a0[q], bv[q], cv[q], lu);
                            val[q] = trunc_i32(x);
                        }
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
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&s_empty[b]));      // this warp is done with buffer b
            if (++b == WS_NBUF) {
                b = 0;
                ++round;
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
    const bool vec4 = (image_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_table) & 15) == 0);
    // The warp-specialised kernel needs 16-byte aligned frame rows (cp.async) and the camera table
    // resident in shared memory; anything else takes the phased kernel.
    static const char *variant = getenv("RTI_FUSED_KERNEL");       // "phased" forces the phased kernel (A/B runs)
    const bool ws_ok = n <= WS_MAX_LIGHTS && ws_smem_bytes(n, g) <= WS_SMEM_LIMIT && frame_stride % 16 == 0 &&
                       (reinterpret_cast<uintptr_t>(d_frames) & 15) == 0 && !(variant && variant[0] == 'p');
    if (!ws_ok) {
        const size_t smem = (size_t)g * sizeof(double);
        const int64_t blocks = (n_pixels + FU_THREADS - 1) / FU_THREADS;
        if (vec4)
            ptm_fused_kernel<true><<<(unsigned)blocks, FU_THREADS, smem, as_stream(stream)>>>(
                d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_axis, g, d_coef, plane_stride, d_table,
                image_stride, d_bad);
        else
            ptm_fused_kernel<false><<<(unsigned)blocks, FU_THREADS, smem, as_stream(stream)>>>(
                d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_axis, g, d_coef, plane_stride, d_table,
                image_stride, d_bad);
    } else {
        int dev = 0, sms = 0;
        RTI_CUDA(cudaGetDevice(&dev));
        RTI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        const int64_t n_tiles = (n_pixels + WS_PIX - 1) / WS_PIX;
        const unsigned grid = (unsigned)(n_tiles < sms ? n_tiles : sms);
        const size_t smem = ws_smem_bytes(n, g);
        unsigned int *counter = nullptr;
        int rc = claim_tile_counter(dev, as_stream(stream), &counter);
        if (rc != RTI_OK) return rc;
#define RTI_WS_LAUNCH(V)                                                                                        \
    do {                                                                                                        \
        RTI_CUDA(cudaFuncSetAttribute(ptm_fused_ws_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                      (int)WS_SMEM_LIMIT));                                \
        ptm_fused_ws_kernel<V><<<grid, WS_THREADS, smem, as_stream(stream)>>>(                                  \
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_axis, g, d_coef, plane_stride, d_table,  \
            image_stride, d_bad, counter);                                                                      \
    } while (0)
        if (vec4) RTI_WS_LAUNCH(true);
        else RTI_WS_LAUNCH(false);
#undef RTI_WS_LAUNCH
    }
    RTI_LAUNCHED("ptm_fused_kernel");
    return RTI_OK;
}
