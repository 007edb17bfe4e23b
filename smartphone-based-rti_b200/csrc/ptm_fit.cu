// PTM fit kernels: per-pixel least squares of the 6 biquadratic coefficients
// (reference: _interpolate_PTM, analysis.py:282-298).
#include "common.cuh"

namespace rti {

// --------------------------------------------------------------------------------------
// Fused gather + fit straight from the uint8 frames (no SoA materialisation).
// One thread per pixel; consecutive threads own consecutive pixels so each frame is
// read as one fully used 32-byte sector per warp.  The camera table (cx, cy, cz^2 per
// frame) is staged in shared memory in chunks.  Bound: fp64 pipe (~50 fp64 issue slots
// per pixel.light against 1 byte of HBM traffic).
// --------------------------------------------------------------------------------------
constexpr int FIT_THREADS = 128;
constexpr int FIT_CAM_CHUNK = 512;      // frames of camera data resident in smem at a time
constexpr int FIT_UNROLL = 4;           // independent byte loads in flight per thread

// MASKED: per-pixel validity masks.  A sample takes part in pixel p's fit only when its intensity lies in
// [lo, hi] (shadow / highlight rejection), so every pixel has its own sample set; the oracle is the
// reference's _interpolate_PTM called with the pixel's masked sub-arrays.  Rejected
// samples enter with a zero design row and zero intensity, the sample count becomes per pixel, and a
// pixel with fewer than 6 valid samples is flagged like a singular one (the reference raises, :296-298).
// ASYNC: the CTA's frame bytes ([64 frames][128 pixels] per chunk) come through a three-stage cp.async
// ring two chunks ahead instead of four byte loads per thread (needs 16-byte aligned frame rows).
constexpr int FIT_CHUNK = 64;
constexpr int FIT_STAGES = 3;

template <bool MASKED, bool ASYNC>
__global__ void __launch_bounds__(FIT_THREADS)
ptm_fit_frames_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride,
                      const double *__restrict__ cams, int n, int row0, int cols,
                      int64_t n_pixels, double *__restrict__ coef, int64_t plane_stride,
                      int32_t *__restrict__ bad, int lo, int hi, int32_t *__restrict__ valid_count) {
    __shared__ double s_cam[FIT_CAM_CHUNK][3];
    __shared__ __align__(16) uint8_t s_ring[ASYNC ? FIT_STAGES * FIT_CHUNK * FIT_THREADS : 16];
    const int64_t p = (int64_t)blockIdx.x * FIT_THREADS + threadIdx.x;
    const bool active = p < n_pixels;
    const int64_t pc = active ? p : n_pixels - 1;           // clamp: tail threads redo the last pixel
    const int yl = (int)(pc / cols);
    const double xd = (double)(int)(pc - (int64_t)yl * cols);
    const double yd = (double)(row0 + yl);
    const uint8_t *src = frames + pc;

    NormalEq ne;
    ne.clear();
    int count = MASKED ? 0 : n;
    auto add_sample = [&](float lu, float lv, unsigned v) {
        if (MASKED) {
            const bool valid = (int)v >= lo && (int)v <= hi;
            count += valid ? 1 : 0;
            lu = valid ? lu : 0.0f;
            lv = valid ? lv : 0.0f;
            v = valid ? v : 0u;
        }
        ne.add(lu, lv, u8_to_double(v));
    };
    // ASYNC: chunk c = frames [64 c, 64 c + 64) goes to ring stage c % 3; 4 x 16 bytes per thread
    const int64_t pbase = (int64_t)blockIdx.x * FIT_THREADS;
    auto issue = [&](int c) {
        const int f0 = c * FIT_CHUNK;
#pragma unroll
        for (int r = 0; r < FIT_CHUNK * (FIT_THREADS / 16) / FIT_THREADS; ++r) {
            const int piece = threadIdx.x + r * FIT_THREADS;
            const int f = piece >> 3, seg = piece & 7;
            const int64_t left = n_pixels - (pbase + seg * 16);
            if (f0 + f < n && left > 0) {
                const unsigned dst = (unsigned)__cvta_generic_to_shared(
                    s_ring + ((c % FIT_STAGES) * FIT_CHUNK + f) * FIT_THREADS + seg * 16);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst),
                             "l"(frames + (int64_t)(f0 + f) * frame_stride + pbase + seg * 16),
                             "r"(left >= 16 ? 16 : (int)left) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    if (ASYNC) {
        issue(0);
        issue(1);
    }
    for (int f0 = 0; f0 < n; f0 += FIT_CAM_CHUNK) {
        const int fc = min(FIT_CAM_CHUNK, n - f0);
        __syncthreads();
        for (int i = threadIdx.x; i < fc; i += FIT_THREADS) {
            const double cz = cams[(int64_t)(f0 + i) * 3 + 2];
            s_cam[i][0] = cams[(int64_t)(f0 + i) * 3 + 0];
            s_cam[i][1] = cams[(int64_t)(f0 + i) * 3 + 1];
            s_cam[i][2] = cz * cz;
        }
        __syncthreads();
        if (ASYNC) {
            for (int i0 = 0; i0 < fc; i0 += FIT_CHUNK) {
                const int c = (f0 + i0) / FIT_CHUNK;
                asm volatile("cp.async.wait_group 1;\n" ::: "memory");
                __syncthreads();                  // chunk c has landed for everyone; chunk c-1's stage is free
                issue(c + 2);
                const uint8_t *bytes = s_ring + (c % FIT_STAGES) * FIT_CHUNK * FIT_THREADS + threadIdx.x;
                const int kc = min(FIT_CHUNK, fc - i0);
                int k = 0;
                for (; k + FIT_UNROLL <= kc; k += FIT_UNROLL) {
#pragma unroll
                    for (int j = 0; j < FIT_UNROLL; ++j) {
                        float lu, lv;
                        light_vector(s_cam[i0 + k + j][0], s_cam[i0 + k + j][1], s_cam[i0 + k + j][2], xd, yd, lu, lv);
                        add_sample(lu, lv, bytes[(k + j) * FIT_THREADS]);
                    }
                }
                for (; k < kc; ++k) {
                    float lu, lv;
                    light_vector(s_cam[i0 + k][0], s_cam[i0 + k][1], s_cam[i0 + k][2], xd, yd, lu, lv);
                    add_sample(lu, lv, bytes[k * FIT_THREADS]);
                }
            }
            continue;
        }
        int i = 0;
        for (; i + FIT_UNROLL <= fc; i += FIT_UNROLL) {
            uint8_t v[FIT_UNROLL];
#pragma unroll
            for (int j = 0; j < FIT_UNROLL; ++j) v[j] = __ldg(src + (int64_t)(f0 + i + j) * frame_stride);
#pragma unroll
            for (int j = 0; j < FIT_UNROLL; ++j) {
                float lu, lv;
                light_vector(s_cam[i + j][0], s_cam[i + j][1], s_cam[i + j][2], xd, yd, lu, lv);
                add_sample(lu, lv, v[j]);
            }
        }
        for (; i < fc; ++i) {
            const uint8_t v = __ldg(src + (int64_t)(f0 + i) * frame_stride);
            float lu, lv;
            light_vector(s_cam[i][0], s_cam[i][1], s_cam[i][2], xd, yd, lu, lv);
            add_sample(lu, lv, v);
        }
    }
    if (ASYNC) asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    double a[6];
    bool ok = solve_normal_eq(ne, (double)count, a);
    if (MASKED && count < 6) {
        ok = false;
#pragma unroll
        for (int k = 0; k < 6; ++k) a[k] = __longlong_as_double(0x7FF8000000000000LL);
    }
    if (active) {
#pragma unroll
        for (int k = 0; k < 6; ++k) coef[k * plane_stride + p] = a[k];
        if (MASKED && valid_count != nullptr) valid_count[p] = count;
        if (!ok && bad != nullptr) atomicAdd(bad, 1);
    }
}

// --------------------------------------------------------------------------------------
// Fit from materialised (lx, ly, intensity) arrays [pixel][n], light index fastest
// (the compute_intensities layout, analysis.py:217-219).
// A CTA owns 128 consecutive pixels and runs 256 threads: thread (half, pixel) accumulates lights
// [16 half, 16 half + 16) of every stage of its pixel, the upper half hands its 26 sums to the lower one
// through shared memory at the end (16 warps per SM instead of 8: with 2 warps per scheduler the fp64
// and conversion latencies were exposed, issue slots 49 % busy; 0.381 -> 0.355 ms at 262 144 px x 500).
// The three arrays are streamed through shared memory in stages of 32 lights with cp.async (16-byte
// chunks: 8 consecutive threads fetch one pixel's 128-byte line, so global reads are whole lines),
// double buffered; the source addresses are computed once per thread (12 pointers), a stage costs one
// 64-bit add per copy (the per-copy 64-bit multiplies it replaces were 10 of ~50 issue slots per sample:
// 0.408 -> 0.381 ms).  Chunk c of row t lives at chunk position c ^ (t & 7): the cp.async writes and
// the per-thread LDS.128 reads (row = pixel) are both bank-conflict free.  Every pixel's 26 running
// sums stay in registers: no cross-lane reduction, and the 6x6 solves run on all lanes of the lower half.
// Lights past n and pixels past n_pixels are zero-filled by cp.async (src-size 0), which
// adds exact zeros to the sums.  The 16-byte chunks are aligned in memory for ANY n: when
// n % 4 != 0 a row's first chunk starts with 1-3 elements of the previous row, which the
// owning thread zeroes after the load.
// Bound: 12 B of HBM per pixel.light against 26 fp64 instructions + 6 conversions; tools/peak_pipes.cu
// (profiles/r02_peak_pipes.jsonl) measures 2.17 clocks per DFMA and scheduler, and a conversion that
// involves fp64 costs ~2.8 clocks of the fp64 pipe on top of ~8.5 of the XU pipe: ~73 fp64-pipe clocks
// per warp and sample = 0.31 ms at this size.  The kernel runs at 0.355 ms = 0.69 of the HBM rate
// (0.73 at 1 M pixels); integer-pipe replacements of the conversions were measured slower (0.450 ms).
// --------------------------------------------------------------------------------------
constexpr int SOA_PIX = 128;                 // pixels per CTA
constexpr int SOA_THREADS = 256;             // two threads per pixel, each takes half of a stage's lights
constexpr int SOA_KL = 32;                   // lights per stage (one 128-byte line per array and pixel)
constexpr int SOA_ARRAY_BYTES = SOA_PIX * SOA_KL * 4;            // 16 KB
constexpr int SOA_STAGE_BYTES = 3 * SOA_ARRAY_BYTES;             // 48 KB
constexpr size_t SOA_SMEM = 2 * SOA_STAGE_BYTES;                 // 96 KB -> 2 CTAs per SM

__device__ __forceinline__ void cp_async_zfill(void *smem, const void *gmem, int bytes, int src_bytes) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    if (bytes == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(src_bytes));
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(s), "l"(gmem), "r"(src_bytes));
}

template <bool VEC16>
__global__ void __launch_bounds__(SOA_THREADS, 2)
ptm_fit_soa_kernel(const float *__restrict__ lx, const float *__restrict__ ly,
                   const int32_t *__restrict__ inten, int64_t n_pixels, int n,
                   double *__restrict__ coef, int64_t plane_stride, int32_t *__restrict__ bad) {
    extern __shared__ __align__(16) unsigned char s_soa[];
    const int t = threadIdx.x;
    const int tp = t & (SOA_PIX - 1);            // the thread's pixel (row of the stage)
    const int half = t >> 7;                     // which 16 of a stage's 32 lights it accumulates
    const int64_t p0 = (int64_t)blockIdx.x * SOA_PIX;
    const void *base[3] = {lx, ly, inten};

    // VEC16 addressing, hoisted out of the stage loop: thread t always fetches chunk position c = t & 7 of
    // rows j*32 + (t >> 3), j = 0..3, of all three arrays.  16-byte chunks are aligned in MEMORY, not in
    // the light index: a row starts at element p*n, a multiple of 4 only when n is, so the chunk holding
    // the row start may begin with up to 3 elements of the previous row (the consumer masks them).
    // Per row: the address of this thread's chunk in stage 0 (per array) and the number of elements from
    // there to the row's end (<= 0: nothing to fetch, also for rows past n_pixels).  A stage then costs
    // one 64-bit add per copy; a copy with nothing to fetch has src-size 0 (zero fill, no access).
    constexpr int VJ = SOA_PIX * 8 / SOA_THREADS;                    // rows per thread and array
    const int vc = t & 7, vr = t >> 3;
    const char *v_src[3][VJ];
    int v_len[VJ];
    if (VEC16) {
#pragma unroll
        for (int j = 0; j < VJ; ++j) {
            const int64_t p = p0 + j * (SOA_THREADS / 8) + vr;
            const int64_t e0 = p * n;
            const int lead_j = (int)(e0 & 3);
            const bool in = p < n_pixels;
            v_len[j] = in ? n + lead_j - 4 * vc : 0;
            const int64_t first = in ? e0 - lead_j + 4 * vc : 0;
#pragma unroll
            for (int a = 0; a < 3; ++a) v_src[a][j] = reinterpret_cast<const char *>(base[a]) + first * 4;
        }
    }
    const unsigned v_dst = (unsigned)__cvta_generic_to_shared(s_soa) + vr * 128 + ((vc ^ (vr & 7)) << 4);

    auto issue = [&](int stage, int i0) {
        unsigned char *dst0 = s_soa + stage * SOA_STAGE_BYTES;
        if (VEC16) {
#pragma unroll
            for (int j = 0; j < VJ; ++j) {
                const int left = v_len[j] - i0;                      // elements of the row from this chunk on
                const int nb = 4 * (left < 0 ? 0 : (left > 4 ? 4 : left));
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const unsigned d = v_dst + stage * SOA_STAGE_BYTES + a * SOA_ARRAY_BYTES + j * (SOA_THREADS / 8) * 128;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d),
                                 "l"(v_src[a][j] + (unsigned)i0 * 4u), "r"(nb));
                }
            }
        } else {
            for (int it = 0; it < 3 * SOA_PIX * SOA_KL / SOA_THREADS; ++it) {
                const int idx = it * SOA_THREADS + t;
                const int a = idx / (SOA_PIX * SOA_KL);
                const int rem = idx - a * (SOA_PIX * SOA_KL);
                const int row = rem >> 5, w = rem & 31;
                const int64_t p = p0 + row;
                const int light = i0 + w;
                const int nb = (p < n_pixels && light < n) ? 4 : 0;
                const char *src = reinterpret_cast<const char *>(base[a]);
                if (nb > 0) src += (p * n + light) * 4;
                cp_async_zfill(dst0 + a * SOA_ARRAY_BYTES + row * 128 + (((w >> 2) ^ (row & 7)) << 4) + ((w & 3) << 2),
                               src, 4, nb);
            }
        }
        asm volatile("cp.async.commit_group;\n" ::);
    };

    NormalEq ne;
    ne.clear();
    // VEC16 with n % 4 != 0: rows are misaligned by up to 3 elements, one more stage may be needed
    const int n_stages = (n + ((VEC16 && (n & 3)) ? 3 : 0) + SOA_KL - 1) / SOA_KL;
    const int lead = VEC16 ? (int)(((p0 + tp) * n) & 3) : 0;        // elements of the previous row in chunk 0
    issue(0, 0);
    for (int s = 0; s < n_stages; ++s) {
        if (s + 1 < n_stages) {
            issue((s + 1) & 1, (s + 1) * SOA_KL);
            asm volatile("cp.async.wait_group 1;\n" ::);
        } else {
            asm volatile("cp.async.wait_group 0;\n" ::);
        }
        __syncthreads();
        const unsigned char *row = s_soa + (s & 1) * SOA_STAGE_BYTES + tp * 128;
#pragma unroll 2
        for (int cc = 0; cc < 4; ++cc) {
            const int c = 4 * half + cc;
            const int off = (c ^ (tp & 7)) << 4;
            float4 u = *reinterpret_cast<const float4 *>(row + off);
            float4 v = *reinterpret_cast<const float4 *>(row + SOA_ARRAY_BYTES + off);
            int4 w = *reinterpret_cast<const int4 *>(row + 2 * SOA_ARRAY_BYTES + off);
            if (VEC16 && s == 0 && c == 0 && lead > 0) {               // zero = no contribution (see above)
                u.x = v.x = 0.0f, w.x = 0;
                if (lead > 1) u.y = v.y = 0.0f, w.y = 0;
                if (lead > 2) u.z = v.z = 0.0f, w.z = 0;
            }
            ne.add(u.x, v.x, (double)w.x);
            ne.add(u.y, v.y, (double)w.y);
            ne.add(u.z, v.z, (double)w.z);
            ne.add(u.w, v.w, (double)w.w);
        }
        __syncthreads();                     // buffer s & 1 is refilled by the next iteration's issue
    }
    // the upper half hands its 26 sums over through shared memory ([sum][pixel], conflict free)
    double *s_sum = reinterpret_cast<double *>(s_soa);
    if (half == 1) {
#pragma unroll
        for (int k = 0; k < 15; ++k) s_sum[k * SOA_PIX + tp] = ne.g[k];
#pragma unroll
        for (int k = 0; k < 5; ++k) s_sum[(15 + k) * SOA_PIX + tp] = ne.s[k];
#pragma unroll
        for (int k = 0; k < 6; ++k) s_sum[(20 + k) * SOA_PIX + tp] = ne.r[k];
    }
    __syncthreads();
    if (half == 1) return;
#pragma unroll
    for (int k = 0; k < 15; ++k) ne.g[k] += s_sum[k * SOA_PIX + tp];
#pragma unroll
    for (int k = 0; k < 5; ++k) ne.s[k] += s_sum[(15 + k) * SOA_PIX + tp];
#pragma unroll
    for (int k = 0; k < 6; ++k) ne.r[k] += s_sum[(20 + k) * SOA_PIX + tp];
    const int64_t p = p0 + tp;
    double a[6];
    const bool ok = solve_normal_eq(ne, (double)n, a);
    if (p < n_pixels) {
#pragma unroll
        for (int k = 0; k < 6; ++k) coef[k * plane_stride + p] = a[k];
        if (!ok && bad != nullptr) atomicAdd(bad, 1);
    }
}

// --------------------------------------------------------------------------------------
// Shared-light fast path: coef[6][P] = pinv[6][n] . I[n][P] (fp64 accumulate).
// Each thread owns 4 consecutive pixels (one 32-bit load per frame), the pseudo-inverse
// columns are broadcast from shared memory.
// --------------------------------------------------------------------------------------
constexpr int SH_THREADS = 128;
constexpr int SH_CHUNK = 256;
constexpr int SH_UNROLL = 16;

// Measured variants (262 144 px x 500 lights, tools/prof_shared_fit.py; ncu of this form: fp64 pipe 45 %,
// warps active 18 %, stalled on the frame loads): this form (4 pixels per thread, 16 loads in flight) 0.133 ms =
// 0.32 of the DFMA peak, 0.45 at 1 M pixels; frames split over the 8 warps of a CTA with a shared-memory
// reduction (4x the threads) 0.195 ms; 2 pixels per thread 0.133 / 0.426 ms, 1 pixel 0.150 / 0.512 ms: more
// threads do not help, the pinv LDS per DFMA and the byte unpacking grow instead.  Kept as is.

__global__ void __launch_bounds__(SH_THREADS)
ptm_fit_shared_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride,
                      const double *__restrict__ pinv, int n, int64_t n_pixels,
                      double *__restrict__ coef, int64_t plane_stride) {
    __shared__ double s_p[SH_CHUNK][6];
    const int64_t p0 = ((int64_t)blockIdx.x * SH_THREADS + threadIdx.x) * 4;
    const bool vec = (p0 + 3 < n_pixels) && ((frame_stride & 3) == 0) &&
                     ((reinterpret_cast<uintptr_t>(frames) & 3) == 0);
    double acc[4][6];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int k = 0; k < 6; ++k) acc[q][k] = 0.0;
    for (int f0 = 0; f0 < n; f0 += SH_CHUNK) {
        const int fc = min(SH_CHUNK, n - f0);
        __syncthreads();
        for (int i = threadIdx.x; i < fc * 6; i += SH_THREADS)
            s_p[i / 6][i % 6] = pinv[(int64_t)(i % 6) * n + f0 + i / 6];
        __syncthreads();
        if (p0 >= n_pixels) continue;
        // SH_UNROLL frame words are requested before the first is used: the kernel is a stream of
        // 4-byte loads with 24 DFMAs each, and without the batch it is bound by load latency
        auto load_word = [&](int i) -> uint32_t {
            const uint8_t *src = frames + (int64_t)(f0 + i) * frame_stride + p0;
            if (vec) return __ldg(reinterpret_cast<const uint32_t *>(src));
            uint32_t w = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (p0 + q < n_pixels) w |= (uint32_t)__ldg(src + q) << (8 * q);
            return w;
        };
        auto accumulate = [&](int i, uint32_t w) {
            double pk[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) pk[k] = s_p[i][k];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double v = u8_to_double((w >> (8 * q)) & 0xffu);     // no I2F (XU pipe)
#pragma unroll
                for (int k = 0; k < 6; ++k) acc[q][k] = fma(pk[k], v, acc[q][k]);
            }
        };
        int i = 0;
        for (; i + SH_UNROLL <= fc; i += SH_UNROLL) {
            uint32_t w[SH_UNROLL];
#pragma unroll
            for (int j = 0; j < SH_UNROLL; ++j) w[j] = load_word(i + j);
#pragma unroll
            for (int j = 0; j < SH_UNROLL; ++j) accumulate(i + j, w[j]);
        }
        for (; i < fc; ++i) accumulate(i, load_word(i));
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
        if (p0 + q < n_pixels)
#pragma unroll
            for (int k = 0; k < 6; ++k) coef[k * plane_stride + p0 + q] = acc[q][k];
}

}  // namespace rti

using namespace rti;

extern "C" int rti_ptm_fit_frames(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams,
                                  int n, int row0, int rows, int cols, double *d_coef,
                                  int64_t plane_stride, int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_cams && d_coef, "rti_ptm_fit_frames: NULL pointer");
    RTI_REQUIRE(rows > 0 && cols > 0 && row0 >= 0, "rti_ptm_fit_frames: bad slab %d+%d x %d", row0, rows, cols);
    if (n < 6) {
        set_error("rti_ptm_fit_frames: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    const int64_t n_pixels = (int64_t)rows * cols;
    RTI_REQUIRE(plane_stride >= n_pixels, "rti_ptm_fit_frames: plane_stride < pixels");
    RTI_REQUIRE(frame_stride >= n_pixels, "rti_ptm_fit_frames: frame_stride < slab pixels");
    const int64_t blocks = (n_pixels + FIT_THREADS - 1) / FIT_THREADS;
    if (frame_stride % 16 == 0 && (reinterpret_cast<uintptr_t>(d_frames) & 15) == 0)
        ptm_fit_frames_kernel<false, true><<<(unsigned)blocks, FIT_THREADS, 0, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_coef, plane_stride, d_bad, 0, 255, nullptr);
    else
        ptm_fit_frames_kernel<false, false><<<(unsigned)blocks, FIT_THREADS, 0, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_coef, plane_stride, d_bad, 0, 255, nullptr);
    RTI_LAUNCHED("ptm_fit_frames_kernel");
    return RTI_OK;
}

extern "C" int rti_ptm_fit_frames_masked(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams,
                                         int n, int row0, int rows, int cols, int lo, int hi, double *d_coef,
                                         int64_t plane_stride, int32_t *d_valid_count, int32_t *d_bad,
                                         rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_cams && d_coef, "rti_ptm_fit_frames_masked: NULL pointer");
    RTI_REQUIRE(rows > 0 && cols > 0 && row0 >= 0, "rti_ptm_fit_frames_masked: bad slab %d+%d x %d", row0, rows, cols);
    RTI_REQUIRE(lo >= 0 && hi <= 255 && lo <= hi, "rti_ptm_fit_frames_masked: bad intensity range [%d, %d]", lo, hi);
    if (n < 6) {
        set_error("rti_ptm_fit_frames_masked: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    const int64_t n_pixels = (int64_t)rows * cols;
    RTI_REQUIRE(plane_stride >= n_pixels && frame_stride >= n_pixels, "rti_ptm_fit_frames_masked: stride < slab pixels");
    const int64_t blocks = (n_pixels + FIT_THREADS - 1) / FIT_THREADS;
    if (frame_stride % 16 == 0 && (reinterpret_cast<uintptr_t>(d_frames) & 15) == 0)
        ptm_fit_frames_kernel<true, true><<<(unsigned)blocks, FIT_THREADS, 0, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_coef, plane_stride, d_bad, lo, hi, d_valid_count);
    else
        ptm_fit_frames_kernel<true, false><<<(unsigned)blocks, FIT_THREADS, 0, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_coef, plane_stride, d_bad, lo, hi, d_valid_count);
    RTI_LAUNCHED("ptm_fit_frames_kernel<masked>");
    return RTI_OK;
}

extern "C" int rti_ptm_fit_soa(const float *d_lx, const float *d_ly, const int32_t *d_inten,
                               int64_t n_pixels, int n, double *d_coef, int64_t plane_stride,
                               int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_inten && d_coef, "rti_ptm_fit_soa: NULL pointer");
    RTI_REQUIRE(n_pixels > 0, "rti_ptm_fit_soa: no pixels");
    if (n < 6) {
        set_error("rti_ptm_fit_soa: PTM needs >= 6 light samples, got %d", n);
        return n <= 0 ? RTI_ERR_ARG : RTI_ERR_TOO_FEW;
    }
    RTI_REQUIRE(plane_stride >= n_pixels, "rti_ptm_fit_soa: plane_stride < pixels");
    const int64_t blocks = (n_pixels + SOA_PIX - 1) / SOA_PIX;
    RTI_REQUIRE(blocks <= 0x7fffffffLL, "rti_ptm_fit_soa: too many pixels");
    const bool vec16 = ((reinterpret_cast<uintptr_t>(d_lx) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(d_ly) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(d_inten) & 15) == 0);
    RTI_CUDA(cudaFuncSetAttribute(ptm_fit_soa_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SOA_SMEM));
    RTI_CUDA(cudaFuncSetAttribute(ptm_fit_soa_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SOA_SMEM));
    if (vec16)
        ptm_fit_soa_kernel<true><<<(unsigned)blocks, SOA_THREADS, SOA_SMEM, as_stream(stream)>>>(
            d_lx, d_ly, d_inten, n_pixels, n, d_coef, plane_stride, d_bad);
    else
        ptm_fit_soa_kernel<false><<<(unsigned)blocks, SOA_THREADS, SOA_SMEM, as_stream(stream)>>>(
            d_lx, d_ly, d_inten, n_pixels, n, d_coef, plane_stride, d_bad);
    RTI_LAUNCHED("ptm_fit_soa_kernel");
    return RTI_OK;
}

extern "C" int rti_ptm_fit_shared(const uint8_t *d_frames, int64_t frame_stride, const double *d_pinv,
                                  int n, int64_t n_pixels, double *d_coef, int64_t plane_stride,
                                  rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_pinv && d_coef, "rti_ptm_fit_shared: NULL pointer");
    RTI_REQUIRE(n > 0 && n_pixels > 0, "rti_ptm_fit_shared: empty input");
    RTI_REQUIRE(plane_stride >= n_pixels && frame_stride >= n_pixels, "rti_ptm_fit_shared: stride < pixels");
    const int64_t blocks = (n_pixels + SH_THREADS * 4 - 1) / (SH_THREADS * 4);
    ptm_fit_shared_kernel<<<(unsigned)blocks, SH_THREADS, 0, as_stream(stream)>>>(
        d_frames, frame_stride, d_pinv, n, n_pixels, d_coef, plane_stride);
    RTI_LAUNCHED("ptm_fit_shared_kernel");
    return RTI_OK;
}
