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
// One CTA converts runs of 1024 pixels of four ROI rows: the runs' bytes are staged in shared
// memory with 16-byte loads (rows start at arbitrary byte offsets inside the frame), each
// thread then converts 4 consecutive pixels per row and stores them as one / three 4-byte words.
#include "common.cuh"

namespace rti {

constexpr int ROI_THREADS = 256;
constexpr int ROI_PIX_PER_THREAD = 4;
constexpr int ROI_TILE = ROI_THREADS * ROI_PIX_PER_THREAD;       // pixels per CTA
constexpr int ROI_STAGE_BYTES = ROI_TILE * 3 + 48;               // + alignment slack both ends, word over-read

__device__ __forceinline__ uint4 ld_stream_v4(const void *p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

constexpr int ROI_ROWS = 4;                                      // ROI rows per CTA (bytes in flight)

template <bool HSV>
__global__ void __launch_bounds__(ROI_THREADS)
roi_extract_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride, int64_t row_pitch,
                   int x_min, int y_min, int rows, int cols, uint8_t *__restrict__ out,
                   int64_t out_stride) {
    __shared__ __align__(16) uint8_t stage[ROI_ROWS][ROI_STAGE_BYTES];
    __shared__ int sdiv[HSV ? 256 : 1];
    __shared__ int hdiv[HSV ? 256 : 1];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * ROI_TILE;                        // first ROI column of the run
    const int y0 = blockIdx.y * ROI_ROWS;
    const int64_t f = blockIdx.z;
    const int npix = min(ROI_TILE, cols - x0);
    const int nrows = min(ROI_ROWS, rows - y0);

    if (HSV) {
        // tables as OpenCV builds them: saturate_cast<int>(x) = rint(x) for doubles
        const int i = tid;
        sdiv[i] = i ? __double2int_rn((double)(255 << 12) / (double)i) : 0;
        hdiv[i] = i ? __double2int_rn((double)(180 << 12) / (6.0 * (double)i)) : 0;
    }

    // ---- stage [a0, a1) of every row through 16-byte granules; all rows' loads are issued
    // before the barrier so one CTA keeps up to 12 KB in flight
    const uint8_t *src0 = frames + f * frame_stride + (int64_t)(y_min + y0) * row_pitch + (int64_t)(x_min + x0) * 3;
#pragma unroll
    for (int r = 0; r < ROI_ROWS; ++r) {
        if (r >= nrows) continue;
        const uintptr_t a0 = (uintptr_t)(src0 + r * row_pitch);
        const uintptr_t a1 = a0 + (uintptr_t)npix * 3;
        const uintptr_t g0 = a0 & ~(uintptr_t)15;                // stage[r][a0 - g0] = first byte of the run
        const int ngran = (int)((a1 - g0 + 15) >> 4);
        for (int gi = tid; gi < ngran; gi += ROI_THREADS) {
            const uintptr_t ga = g0 + ((uintptr_t)gi << 4);
            if (ga >= a0 && ga + 16 <= a1) {
                *reinterpret_cast<uint4 *>(&stage[r][gi << 4]) = ld_stream_v4((const void *)ga);
            } else {                                             // first / last partial granule
                for (int k = 0; k < 16; ++k) {
                    const uintptr_t a = ga + k;
                    if (a >= a0 && a < a1) stage[r][(gi << 4) + k] = *(const uint8_t *)a;
                }
            }
        }
    }
    __syncthreads();

    // ---- convert 4 pixels per thread and row: the thread's 12 bytes start at skew + 12 tid, i.e.
    // (skew & 3) bytes into a 32-bit word for every thread of the row: 4 word loads + funnel shifts
    const int p0 = tid * ROI_PIX_PER_THREAD;
    if (p0 >= npix) return;
    const int cnt = min(ROI_PIX_PER_THREAD, npix - p0);
    constexpr int BPP = HSV ? 3 : 1;
#pragma unroll
    for (int r = 0; r < ROI_ROWS; ++r) {
        if (r >= nrows) continue;
        const int o = (int)((uintptr_t)(src0 + r * row_pitch) & 15) + p0 * 3;
        const uint32_t *wsrc = reinterpret_cast<const uint32_t *>(&stage[r][o & ~3]);
        const unsigned sh = (unsigned)(o & 3) * 8u;
        const uint32_t w0 = wsrc[0], w1 = wsrc[1], w2 = wsrc[2], w3 = wsrc[3];
        uint32_t in[3];
        in[0] = __funnelshift_r(w0, w1, sh);
        in[1] = __funnelshift_r(w1, w2, sh);
        in[2] = __funnelshift_r(w2, w3, sh);
        uint8_t res[ROI_PIX_PER_THREAD * BPP];
#pragma unroll
        for (int k = 0; k < ROI_PIX_PER_THREAD; ++k) {
            const int b = (in[(3 * k) >> 2] >> (8 * ((3 * k) & 3))) & 0xff;
            const int g = (in[(3 * k + 1) >> 2] >> (8 * ((3 * k + 1) & 3))) & 0xff;
            const int rr = (in[(3 * k + 2) >> 2] >> (8 * ((3 * k + 2) & 3))) & 0xff;
            const int v = max(max(b, g), rr);
            if (HSV) {
                const int diff = v - min(min(b, g), rr);
                const int sat = (diff * sdiv[v] + (1 << 11)) >> 12;
                int h = (v == rr) ? (g - b) : (v == g) ? (b - rr + 2 * diff) : (rr - g + 4 * diff);
                h = (h * hdiv[diff] + (1 << 11)) >> 12;
                h += h < 0 ? 180 : 0;
                res[3 * k + 0] = (uint8_t)h;
                res[3 * k + 1] = (uint8_t)sat;
                res[3 * k + 2] = (uint8_t)v;
            } else {
                res[k] = (uint8_t)v;
            }
        }
        uint8_t *dst = out + f * out_stride + ((int64_t)(y0 + r) * cols + x0 + p0) * BPP;
        if (cnt == ROI_PIX_PER_THREAD && ((uintptr_t)dst & 3) == 0) {
#pragma unroll
            for (int w = 0; w < BPP; ++w) {
                const uint32_t word = (uint32_t)res[4 * w] | ((uint32_t)res[4 * w + 1] << 8) |
                                      ((uint32_t)res[4 * w + 2] << 16) | ((uint32_t)res[4 * w + 3] << 24);
                reinterpret_cast<uint32_t *>(dst)[w] = word;
            }
        } else {
            for (int k = 0; k < cnt * BPP; ++k) dst[k] = res[k];
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
    RTI_REQUIRE(rows <= 65535 && n_frames <= 65535, "rti_roi_extract: more than 65535 rows or frames per call");
    const int64_t bpp = hsv ? 3 : 1;
    RTI_REQUIRE(frame_stride >= (int64_t)height * width * 3 || n_frames == 1, "rti_roi_extract: frame stride too small");
    RTI_REQUIRE(out_stride >= (int64_t)rows * cols * bpp || n_frames == 1, "rti_roi_extract: output stride too small");
    dim3 grid((cols + ROI_TILE - 1) / ROI_TILE, (rows + ROI_ROWS - 1) / ROI_ROWS, n_frames);
    const int64_t pitch = (int64_t)width * 3;
    if (hsv)
        roi_extract_kernel<true><<<grid, ROI_THREADS, 0, as_stream(stream)>>>(d_bgr, frame_stride, pitch, x_min, y_min,
                                                                              rows, cols, d_out, out_stride);
    else
        roi_extract_kernel<false><<<grid, ROI_THREADS, 0, as_stream(stream)>>>(d_bgr, frame_stride, pitch, x_min, y_min,
                                                                               rows, cols, d_out, out_stride);
    RTI_LAUNCHED("roi_extract_kernel");
    return RTI_OK;
}
