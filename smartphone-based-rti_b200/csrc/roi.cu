// ROI extraction upstream of the fit: crop the ROI out of BGR frames and
// convert it like OpenCV's 8-bit COLOR_BGR2HSV.
//   Utils/utilities.py:302-342  get_ROI(static_img, points, hsv=True): crop + cv2.cvtColor
//   FeatureMatcher.py:183-184   roi_intensities = get_ROI(..)[:, :, 2]   (V channel = one frame
//                               of the stack compute_intensities reads, analysis.py:196-246)
//   interactive_relighting.py:109  the HSV ROI the relighter writes V into
// OpenCV's integer arithmetic (checked against cv2 4.13 for all 2^24 colours):
//   v = max(b,g,r), diff = v - min(b,g,r)
//   s = (diff * sdiv[v] + 2^11) >> 12,              sdiv[i] = rint(255 * 4096 / i)
//   h = (v==r ? g-b : v==g ? b-r+2 diff : r-g+4 diff),
//   h = (h * hdiv[diff] + 2^11) >> 12 (arithmetic shift), h += 180 if h < 0,
//                                                   hdiv[i] = rint(180 * 4096 / (6 i))
// The op is a byte stream: 3 B read + 1 B (V) or 3 B (HSV) written per ROI pixel, HBM-bound.
// The unit of work is a run of 512 pixels of one ROI row, owned by ONE WARP: lane 0 asks the
// TMA unit for the run's 16-byte granules (one cp.async.bulk per run, rows start at arbitrary
// byte offsets inside the frame), completion is signalled on a warp-private mbarrier, and each
// lane converts 16 consecutive pixels and stores them as 16 (V) / 48 (HSV) bytes.  Every warp keeps
// the next three runs in flight in its own four-stage ring; there is no CTA-wide barrier in the
// loop.  (ncu history: the first version - one CTA per 4 rows x 1024 px, per-thread 16-byte loads -
// was issue-bound at 35 instructions per pixel, most of them address arithmetic for the loads.)
#include "common.cuh"
#include "tma.cuh"
#include <algorithm>

namespace rti {

constexpr int ROI_WARPS = 8;
constexpr int ROI_THREADS = ROI_WARPS * 32;
constexpr int ROI_PIX_PER_LANE = 16;                             // 48 B read, 16 / 48 B written per lane and run
constexpr int ROI_UNIT = 32 * ROI_PIX_PER_LANE;                  // pixels per run
constexpr int ROI_UNIT_BYTES = ROI_UNIT * 3 + 32;                // + up to 15 B of skew, whole granules, word over-read
constexpr int ROI_STAGES = 4;                                    // runs per warp: one converted, three in flight
constexpr int ROI_RING_BYTES = ROI_WARPS * ROI_STAGES * ROI_UNIT_BYTES;    // dynamic shared memory per CTA
constexpr int ROI_CTAS_PER_SM = 4;                               // resident at most; the host launches 2-4 per SM
static_assert(ROI_UNIT_BYTES % 16 == 0, "stages hold whole granules");

struct RoiShape {
    int64_t frame_stride, out_stride;
    uintptr_t safe_end;                                          // first byte after the last frame
    int row_pitch, x_min, y_min, rows, cols, segs_x, n_frames;
    int sx, sy, sf;                                              // the grid's stride in warps as (run, row, frame)
};

// Position of a run as (run in the row, ROI row, frame).  A warp advances it by the grid's stride,
// decomposed once by the host, with carries: no division in the loop.
struct RoiCursor {
    int xs, y, f;
    __device__ __forceinline__ void advance(const RoiShape &g) {
        xs += g.sx;
        if (xs >= g.segs_x) { xs -= g.segs_x; ++y; }
        y += g.sy;
        if (y >= g.rows) { y -= g.rows; ++f; }
        f += g.sf;
    }
};

__device__ __forceinline__ const uint8_t *roi_run_src(const uint8_t *frames, const RoiShape &g, const RoiCursor &c) {
    return frames + c.f * g.frame_stride + (int64_t)(g.y_min + c.y) * g.row_pitch + (g.x_min + c.xs * ROI_UNIT) * 3;
}

__device__ __forceinline__ void roi_bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// Lane 0: request the granules that cover the run.  The first granule may start up to 15 bytes
// before the run (same frame row or the previous one: inside the buffer); the last one may end up
// to 15 bytes after it, which is read too unless the frame buffer itself ends there (then the
// consumer copies the last < 16 bytes itself, see roi_wait).
__device__ __forceinline__ void roi_request(uint32_t stage, uint32_t bar, const uint8_t *__restrict__ frames,
                                            const RoiShape &g, const RoiCursor &c) {
    if (c.f >= g.n_frames) return;
    const uint8_t *src = roi_run_src(frames, g, c);
    const uint32_t skew = (uint32_t)(uintptr_t)src & 15u;
    const uint32_t span = skew + (uint32_t)min(ROI_UNIT, g.cols - c.xs * ROI_UNIT) * 3u;
    uint32_t bytes = (span + 15u) & ~15u;
    if ((uintptr_t)(src - skew) + bytes > g.safe_end) bytes = span & ~15u;
    if (bytes) {
        mbar_expect_tx(bar, bytes);
        roi_bulk_load(stage, src - skew, bytes, bar);
    } else {
        mbar_arrive(bar);
    }
}

// All lanes: wait for the run's bytes; returns the skew of the run inside its stage.
__device__ __forceinline__ uint32_t roi_wait(uint8_t *stage, uint32_t bar, uint32_t parity,
                                             const uint8_t *__restrict__ frames, const RoiShape &g,
                                             const RoiCursor &c) {
    const uint8_t *src = roi_run_src(frames, g, c);
    const uint32_t skew = (uint32_t)(uintptr_t)src & 15u;
    const uint32_t span = skew + (uint32_t)min(ROI_UNIT, g.cols - c.xs * ROI_UNIT) * 3u;
    mbar_wait(bar, parity);
    if ((uintptr_t)(src - skew) + ((span + 15u) & ~15u) > g.safe_end) {      // the buffer's very last granule
        for (uint32_t k = (span & ~15u) + (threadIdx.x & 31); k < span; k += 32) stage[k] = (src - skew)[k];
        __syncwarp();
    }
    return skew;
}

// max(b, g, r) of four pixels held in three words b0 g0 r0 b1 | g1 r1 b2 g2 | r2 b3 g3 r3: each
// channel's bytes go to the HIGH byte of a 16-bit lane (the low byte is whatever PRMT leaves
// there; it only breaks ties), one three-input u16x2 max per pixel pair, and the four high bytes
// are packed into one word.
__device__ __forceinline__ uint32_t roi_max4(uint32_t i0, uint32_t i1, uint32_t i2) {
    const uint32_t bA = __byte_perm(i0, i1, 0x3000), gA = __byte_perm(i0, i1, 0x4010), rA = __byte_perm(i0, i1, 0x5020);
    const uint32_t bB = __byte_perm(i1, i2, 0x5020), gB = __byte_perm(i1, i2, 0x6030), rB = __byte_perm(i1, i2, 0x7040);
    return __byte_perm(__vimax3_u16x2(bA, gA, rA), __vimax3_u16x2(bB, gB, rB), 0x7531);
}

// OpenCV's 8-bit BGR -> HSV of pixel k (0..3) of three words, written into the same layout.
// (Measured and not adopted: the hue as selects of its operands instead of the nested ternary, which
// compiles to two divergent branches per pixel: 0.277 -> 0.282 ms; the lanes' 48-byte outputs exchanged
// through the consumed stage and stored as coalesced 512-byte instructions: 0.307 ms.  The HSV
// path is bound by its ~40 integer instructions per pixel, not by branches or the store pattern.)
__device__ __forceinline__ void roi_hsv1(const uint32_t *in, uint32_t *res, int k, const int *sdiv, const int *hdiv) {
    const int b = (in[(3 * k) >> 2] >> (8 * ((3 * k) & 3))) & 0xff;
    const int gg = (in[(3 * k + 1) >> 2] >> (8 * ((3 * k + 1) & 3))) & 0xff;
    const int rr = (in[(3 * k + 2) >> 2] >> (8 * ((3 * k + 2) & 3))) & 0xff;
    const int v = max(max(b, gg), rr);
    const int diff = v - min(min(b, gg), rr);
    const int sat = (diff * sdiv[v] + (1 << 11)) >> 12;
    int h = (v == rr) ? (gg - b) : (v == gg) ? (b - rr + 2 * diff) : (rr - gg + 4 * diff);
    h = (h * hdiv[diff] + (1 << 11)) >> 12;
    h += h < 0 ? 180 : 0;
    res[(3 * k) >> 2] |= (uint32_t)h << (8 * ((3 * k) & 3));
    res[(3 * k + 1) >> 2] |= (uint32_t)sat << (8 * ((3 * k + 1) & 3));
    res[(3 * k + 2) >> 2] |= (uint32_t)v << (8 * ((3 * k + 2) & 3));
}

template <bool HSV>
__global__ void __launch_bounds__(ROI_THREADS, ROI_CTAS_PER_SM)
roi_extract_kernel(const uint8_t *__restrict__ frames, uint8_t *__restrict__ out, const RoiShape g) {
    extern __shared__ __align__(16) uint8_t roi_ring[];          // [ROI_WARPS][ROI_STAGES][ROI_UNIT_BYTES]
    __shared__ __align__(8) uint64_t bars[ROI_WARPS][ROI_STAGES];
    __shared__ int sdiv[HSV ? 256 : 1];
    __shared__ int hdiv[HSV ? 256 : 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (HSV) {
        // tables as OpenCV builds them: saturate_cast<int>(x) = rint(x) for doubles
        const int i = threadIdx.x;
        sdiv[i] = i ? __double2int_rn((double)(255 << 12) / (double)i) : 0;
        hdiv[i] = i ? __double2int_rn((double)(180 << 12) / (6.0 * (double)i)) : 0;
        __syncthreads();
    }
    const uint32_t ring0 = smem_u32(roi_ring + warp * ROI_STAGES * ROI_UNIT_BYTES), bar0 = smem_u32(&bars[warp][0]);
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < ROI_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    // run blockIdx.x * ROI_WARPS + warp, then every (sx, sy, sf)
    const unsigned gw = blockIdx.x * ROI_WARPS + warp;
    RoiCursor cur;                                               // the run being converted
    cur.xs = (int)(gw % g.segs_x);
    cur.y = (int)((gw / g.segs_x) % g.rows);
    cur.f = (int)(gw / g.segs_x / g.rows);
    RoiCursor ahead = cur;                                       // the run being requested
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < ROI_STAGES - 1; ++s) {
            roi_request(ring0 + s * ROI_UNIT_BYTES, bar0 + 8 * s, frames, g, ahead);
            ahead.advance(g);
        }
    }
    constexpr int BPP = HSV ? 3 : 1;
    for (unsigned it = 0; cur.f < g.n_frames; ++it, cur.advance(g)) {
        const unsigned s = it % ROI_STAGES;
        __syncwarp();                                            // every lane has read the stage of run it - 1
        if (lane == 0) {
            const unsigned sn = (it + ROI_STAGES - 1) % ROI_STAGES;
            roi_request(ring0 + sn * ROI_UNIT_BYTES, bar0 + 8 * sn, frames, g, ahead);
            ahead.advance(g);
        }
        uint8_t *stage = roi_ring + (warp * ROI_STAGES + s) * ROI_UNIT_BYTES;
        const uint32_t skew = roi_wait(stage, bar0 + 8 * s, (it / ROI_STAGES) & 1u, frames, g, cur);

        // the lane's 48 bytes start at skew + 48 lane, i.e. (skew & 3) bytes into a 32-bit word for
        // every lane of the run: 13 word loads + funnel shifts
        constexpr int NW = ROI_PIX_PER_LANE * 3 / 4, NQ = ROI_PIX_PER_LANE / 4;
        const uint32_t o = skew + lane * (ROI_PIX_PER_LANE * 3);
        const uint32_t *wsrc = reinterpret_cast<const uint32_t *>(stage + (o & ~3u));
        const unsigned sh = (o & 3u) * 8u;
        uint32_t w[NW + 1], in[NW];
#pragma unroll
        for (int j = 0; j < NW + 1; ++j) w[j] = wsrc[j];
#pragma unroll
        for (int j = 0; j < NW; ++j) in[j] = __funnelshift_r(w[j], w[j + 1], sh);

        const int p0 = cur.xs * ROI_UNIT + lane * ROI_PIX_PER_LANE;
        const int cnt = min(ROI_PIX_PER_LANE, g.cols - p0);
        if (cnt <= 0) continue;
        uint8_t *dst = out + cur.f * g.out_stride + ((int64_t)cur.y * g.cols + p0) * BPP;
        uint32_t res[NQ * BPP];
        if constexpr (!HSV) {
#pragma unroll
            for (int q = 0; q < NQ; ++q) res[q] = roi_max4(in[3 * q], in[3 * q + 1], in[3 * q + 2]);
        } else {
#pragma unroll
            for (int j = 0; j < NW; ++j) res[j] = 0u;
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int q = 0; q < NQ; ++q) roi_hsv1(in + 3 * q, res + 3 * q, k, sdiv, hdiv);
        }
        if (cnt == ROI_PIX_PER_LANE && ((uintptr_t)dst & 15) == 0 && (NQ * BPP) % 4 == 0) {
#pragma unroll
            for (int j = 0; j < NQ * BPP / 4; ++j)
                reinterpret_cast<uint4 *>(dst)[j] = make_uint4(res[4 * j], res[4 * j + 1], res[4 * j + 2], res[4 * j + 3]);
        } else if (cnt == ROI_PIX_PER_LANE && ((uintptr_t)dst & 7) == 0) {
#pragma unroll
            for (int j = 0; j < NQ * BPP / 2; ++j)
                reinterpret_cast<uint2 *>(dst)[j] = make_uint2(res[2 * j], res[2 * j + 1]);
        } else if (cnt == ROI_PIX_PER_LANE && ((uintptr_t)dst & 3) == 0) {
#pragma unroll
            for (int j = 0; j < NQ * BPP; ++j) reinterpret_cast<uint32_t *>(dst)[j] = res[j];
        } else {
#pragma unroll
            for (int j = 0; j < NQ * BPP; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (4 * j + q < cnt * BPP) dst[4 * j + q] = (uint8_t)(res[j] >> (8 * q));
        }
    }
}

}  // namespace rti

using namespace rti;

extern "C" int rti_roi_extract(const uint8_t *d_bgr, int64_t frame_stride, int n_frames, int height, int width,
                               int x_min, int y_min, int rows, int cols, int hsv, uint8_t *d_out,
                               int64_t out_stride, rti_stream_t stream) {
    RTI_REQUIRE(d_bgr && d_out, "rti_roi_extract: NULL buffer");
    RTI_REQUIRE(n_frames > 0 && height > 0 && width > 0, "rti_roi_extract: bad frame shape %d x %d x %d",
                n_frames, height, width);
    RTI_REQUIRE(rows > 0 && cols > 0 && x_min >= 0 && y_min >= 0 && x_min + cols <= width && y_min + rows <= height,
                "rti_roi_extract: ROI [%d,%d)x[%d,%d) outside the %d x %d frame", y_min, y_min + rows, x_min,
                x_min + cols, height, width);
    const int64_t bpp = hsv ? 3 : 1;
    RTI_REQUIRE(frame_stride >= (int64_t)height * width * 3 || n_frames == 1, "rti_roi_extract: frame stride too small");
    RTI_REQUIRE(out_stride >= (int64_t)rows * cols * bpp || n_frames == 1, "rti_roi_extract: output stride too small");
    RTI_REQUIRE((int64_t)width * 3 < (int64_t)1 << 30, "rti_roi_extract: frame rows too wide");
    RoiShape g;
    g.frame_stride = frame_stride; g.out_stride = out_stride; g.row_pitch = width * 3;
    g.safe_end = (uintptr_t)d_bgr + (uintptr_t)((int64_t)(n_frames - 1) * frame_stride + (int64_t)height * width * 3);
    g.x_min = x_min; g.y_min = y_min; g.rows = rows; g.cols = cols; g.n_frames = n_frames;
    g.segs_x = (cols + ROI_UNIT - 1) / ROI_UNIT;
    const int64_t n_units = (int64_t)g.segs_x * rows * n_frames;
    int dev = 0, sms = 148;
    RTI_CUDA(cudaGetDevice(&dev));
    RTI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    // CTAs per SM, measured (tools/prof_roi.py; 40 x 2048^2 out of 3840 x 2160 | 300 x 400^2 out of 1920 x 1080):
    //   V    1: 0.166 | 0.064 ms   2: 0.113 | 0.045   3: 0.123 | 0.041   4: 0.186 | 0.044   5 (3 stages): 0.213 | 0.049
    //   HSV  1: 0.436 | 0.172 ms   2: 0.316 | 0.120   3: 0.284 | 0.109   4: 0.277 | 0.105
    // HSV is bound by its integer arithmetic and wants every warp it can get; the V stream is fastest with
    // 16-24 warps per SM (more TMA requests in flight than that slow it down).
    const int per_sm = hsv ? ROI_CTAS_PER_SM : (cols >= 4 * ROI_UNIT ? 2 : 3);
    const int grid = (int)std::min<int64_t>((n_units + ROI_WARPS - 1) / ROI_WARPS, (int64_t)sms * per_sm);
    const int stride = grid * ROI_WARPS;                         // in runs, as (run, row, frame)
    g.sx = stride % g.segs_x; g.sy = (stride / g.segs_x) % rows; g.sf = stride / g.segs_x / rows;
    // per launch (microseconds): the attribute belongs to the current device's copy of the kernel
    if (hsv)
        RTI_CUDA(cudaFuncSetAttribute(roi_extract_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ROI_RING_BYTES));
    else
        RTI_CUDA(cudaFuncSetAttribute(roi_extract_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ROI_RING_BYTES));
    if (hsv)
        roi_extract_kernel<true><<<grid, ROI_THREADS, ROI_RING_BYTES, as_stream(stream)>>>(d_bgr, d_out, g);
    else
        roi_extract_kernel<false><<<grid, ROI_THREADS, ROI_RING_BYTES, as_stream(stream)>>>(d_bgr, d_out, g);
    RTI_LAUNCHED("roi_extract_kernel");
    return RTI_OK;
}
