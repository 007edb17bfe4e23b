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
// A work item is a run of 1024 pixels of four ROI rows: the runs' bytes are staged in shared
// memory in 16-byte granules (rows start at arbitrary byte offsets inside the frame), each
// thread then converts 4 consecutive pixels per row and stores them as one / three 4-byte words.
// CTAs are persistent and keep the next two items' bytes in flight with cp.async.
#include "common.cuh"
#include <algorithm>

namespace rti {

constexpr int ROI_THREADS = 256;
constexpr int ROI_PIX_PER_THREAD = 4;
constexpr int ROI_TILE = ROI_THREADS * ROI_PIX_PER_THREAD;       // pixels per work item and row
constexpr int ROI_STAGE_BYTES = ROI_TILE * 3 + 48;               // + alignment slack both ends, word over-read
constexpr int ROI_ROWS = 4;                                      // ROI rows per work item
constexpr int ROI_STAGES = 3;                                    // work items staged ahead per CTA
constexpr int ROI_CTAS_PER_SM = 5;
static_assert(ROI_STAGE_BYTES % 16 == 0 && ROI_STAGE_BYTES / 16 <= ROI_THREADS, "one 16-byte granule per thread and row");

// Position of a work item as (column run, row group, frame).  A CTA advances it by its stride,
// decomposed once by the host, with carries: no division in the item loop.
struct RoiCursor {
    int xt, yg, f;
    __device__ __forceinline__ void advance(int sx, int sy, int sf, int tiles_x, int groups_y) {
        xt += sx;
        if (xt >= tiles_x) { xt -= tiles_x; ++yg; }
        yg += sy;
        if (yg >= groups_y) { yg -= groups_y; ++f; }
        f += sf;
    }
};

struct RoiShape {
    int64_t frame_stride, out_stride;
    int row_pitch, x_min, y_min, rows, cols, tiles_x, groups_y, n_frames;
};

__device__ __forceinline__ void roi_cp_async16(uint32_t smem, const void *gmem, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem), "l"(gmem), "r"(src_bytes) : "memory");
}

__device__ __forceinline__ const uint8_t *roi_item_src(const uint8_t *frames, const RoiShape &g, const RoiCursor &c) {
    return frames + c.f * g.frame_stride + (int64_t)(g.y_min + c.yg * ROI_ROWS) * g.row_pitch +
           (g.x_min + c.xt * ROI_TILE) * 3;
}

// ---- request the bytes [a0, a1) of every row of item c in 16-byte granules, one per thread and row.
// The first granule may start up to 15 bytes before a0 (same frame row or the previous one: inside
// the allocation), the last one is cut at a1 and zero-filled.  Only the low address bits vary with
// the row, so the per-row arithmetic is 32-bit.
__device__ __forceinline__ void roi_issue(uint32_t stage_smem, const uint8_t *__restrict__ frames, const RoiShape &g,
                                          const RoiCursor &c) {
    if (c.f < g.n_frames) {
        const int tid = threadIdx.x;
        const int nbytes = min(ROI_TILE, g.cols - c.xt * ROI_TILE) * 3;
        const int nrows = min(ROI_ROWS, g.rows - c.yg * ROI_ROWS);
        const uint8_t *src0 = roi_item_src(frames, g, c);
        const uint32_t lo = (uint32_t)(uintptr_t)src0;
#pragma unroll
        for (int r = 0; r < ROI_ROWS; ++r) {
            if (r >= nrows) continue;
            const int skew = (int)((lo + (uint32_t)(r * g.row_pitch)) & 15u);   // stage[r][skew] = first byte of the run
            const int left = skew + nbytes - (tid << 4);                        // bytes of the run from this granule on
            if (left > 0)
                roi_cp_async16(stage_smem + r * ROI_STAGE_BYTES + (tid << 4),
                               src0 + ((int64_t)r * g.row_pitch - skew + (tid << 4)), min(16, left));
        }
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
}

// ---- convert 4 pixels per thread and row: the thread's 12 bytes start at skew + 12 tid, i.e.
// (skew & 3) bytes into a 32-bit word for every thread of the row: 4 word loads + funnel shifts.
template <bool HSV>
__device__ __forceinline__ void roi_convert(const uint8_t *stage, const uint8_t *__restrict__ frames,
                                            uint8_t *__restrict__ out, const RoiShape &g, const RoiCursor &c,
                                            const int *sdiv, const int *hdiv) {
    constexpr int BPP = HSV ? 3 : 1;
    const int tid = threadIdx.x;
    const int x0 = c.xt * ROI_TILE, y0 = c.yg * ROI_ROWS;
    const int npix = min(ROI_TILE, g.cols - x0);
    const int nrows = min(ROI_ROWS, g.rows - y0);
    const int p0 = tid * ROI_PIX_PER_THREAD;
    if (p0 >= npix) return;
    const int cnt = min(ROI_PIX_PER_THREAD, npix - p0);
    const uint32_t lo = (uint32_t)(uintptr_t)roi_item_src(frames, g, c);
    uint8_t *dst0 = out + c.f * g.out_stride + ((int64_t)y0 * g.cols + x0 + p0) * BPP;
#pragma unroll
    for (int r = 0; r < ROI_ROWS; ++r) {
        if (r >= nrows) continue;
        const int o = (int)((lo + (uint32_t)(r * g.row_pitch)) & 15u) + p0 * 3;
        const uint32_t *wsrc = reinterpret_cast<const uint32_t *>(stage + r * ROI_STAGE_BYTES + (o & ~3));
        const unsigned sh = (unsigned)(o & 3) * 8u;
        const uint32_t w0 = wsrc[0], w1 = wsrc[1], w2 = wsrc[2], w3 = wsrc[3];
        uint32_t in[3];                                          // b0 g0 r0 b1 | g1 r1 b2 g2 | r2 b3 g3 r3
        in[0] = __funnelshift_r(w0, w1, sh);
        in[1] = __funnelshift_r(w1, w2, sh);
        in[2] = __funnelshift_r(w2, w3, sh);
        uint8_t *dst = dst0 + (int64_t)r * g.cols * BPP;
        if (!HSV) {
            // V = max(b, g, r) for two pixels per instruction: each channel's bytes go to the HIGH
            // byte of a 16-bit lane (the low byte is whatever PRMT leaves there; it only breaks ties),
            // one three-input u16x2 max, and the four high bytes are packed into the output word.
            const uint32_t bA = __byte_perm(in[0], in[1], 0x3000), gA = __byte_perm(in[0], in[1], 0x4010),
                           rA = __byte_perm(in[0], in[1], 0x5020);
            const uint32_t bB = __byte_perm(in[1], in[2], 0x5020), gB = __byte_perm(in[1], in[2], 0x6030),
                           rB = __byte_perm(in[1], in[2], 0x7040);
            const uint32_t vA = __vimax3_u16x2(bA, gA, rA), vB = __vimax3_u16x2(bB, gB, rB);
            const uint32_t word = __byte_perm(vA, vB, 0x7531);
            if (cnt == ROI_PIX_PER_THREAD && ((uintptr_t)dst & 3) == 0) {
                *reinterpret_cast<uint32_t *>(dst) = word;
            } else {
                for (int k = 0; k < cnt; ++k) dst[k] = (uint8_t)(word >> (8 * k));
            }
            continue;
        }
        uint32_t res[3] = {0u, 0u, 0u};
#pragma unroll
        for (int k = 0; k < ROI_PIX_PER_THREAD; ++k) {
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
        if (cnt == ROI_PIX_PER_THREAD && ((uintptr_t)dst & 3) == 0) {
#pragma unroll
            for (int w = 0; w < 3; ++w) reinterpret_cast<uint32_t *>(dst)[w] = res[w];
        } else {
            for (int k = 0; k < cnt * 3; ++k) dst[k] = (uint8_t)(res[k >> 2] >> (8 * (k & 3)));
        }
    }
}

// Persistent: a CTA walks work items (frame, group of ROI_ROWS rows, run of ROI_TILE columns) with a
// stride of gridDim.x = (sx, sy, sf); the bytes of the next two items are in flight (cp.async into a
// three-stage ring) while the current one is converted, so a CTA never waits on a cold load.
template <bool HSV>
__global__ void __launch_bounds__(ROI_THREADS, ROI_CTAS_PER_SM)
roi_extract_kernel(const uint8_t *__restrict__ frames, uint8_t *__restrict__ out, const RoiShape g, int sx, int sy,
                   int sf) {
    __shared__ __align__(16) uint8_t ring[ROI_STAGES][ROI_ROWS * ROI_STAGE_BYTES];
    __shared__ int sdiv[HSV ? 256 : 1];
    __shared__ int hdiv[HSV ? 256 : 1];
    if (HSV) {
        // tables as OpenCV builds them: saturate_cast<int>(x) = rint(x) for doubles
        const int i = threadIdx.x;
        sdiv[i] = i ? __double2int_rn((double)(255 << 12) / (double)i) : 0;
        hdiv[i] = i ? __double2int_rn((double)(180 << 12) / (6.0 * (double)i)) : 0;
    }
    const uint32_t ring_smem = (uint32_t)__cvta_generic_to_shared(&ring[0][0]);
    constexpr uint32_t STAGE_BYTES = ROI_ROWS * ROI_STAGE_BYTES;

    // item blockIdx.x (< tiles_x * groups_y * n_frames, the host sizes the grid so)
    RoiCursor cur;                                               // the item being converted
    cur.xt = (int)(blockIdx.x % g.tiles_x);
    cur.yg = (int)((blockIdx.x / g.tiles_x) % g.groups_y);
    cur.f = (int)(blockIdx.x / g.tiles_x / g.groups_y);
    RoiCursor ahead = cur;                                       // the item being requested
    roi_issue(ring_smem, frames, g, ahead);
    ahead.advance(sx, sy, sf, g.tiles_x, g.groups_y);
    roi_issue(ring_smem + STAGE_BYTES, frames, g, ahead);
    int stage = 0;
    for (; cur.f < g.n_frames; cur.advance(sx, sy, sf, g.tiles_x, g.groups_y), stage = stage + 1 == ROI_STAGES ? 0 : stage + 1) {
        asm volatile("cp.async.wait_group 1;\n" ::: "memory");   // this item has landed (the next may be in flight)
        __syncthreads();                                         // ... for every thread; the previous item is converted
        ahead.advance(sx, sy, sf, g.tiles_x, g.groups_y);
        roi_issue(ring_smem + (stage == 0 ? ROI_STAGES - 1 : stage - 1) * STAGE_BYTES, frames, g, ahead);
        roi_convert<HSV>(ring[stage], frames, out, g, cur, sdiv, hdiv);
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
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
    const int tiles_x = (cols + ROI_TILE - 1) / ROI_TILE, groups_y = (rows + ROI_ROWS - 1) / ROI_ROWS;
    const int64_t n_items = (int64_t)tiles_x * groups_y * n_frames;
    int dev = 0, sms = 148;
    RTI_CUDA(cudaGetDevice(&dev));
    RTI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = (int)std::min<int64_t>(n_items, (int64_t)sms * ROI_CTAS_PER_SM);
    const int sx = grid % tiles_x, sy = (grid / tiles_x) % groups_y, sf = grid / tiles_x / groups_y;   // the stride as (run, group, frame)
    RTI_REQUIRE((int64_t)width * 3 * ROI_ROWS < (int64_t)1 << 31, "rti_roi_extract: frame rows too wide");
    RoiShape g;
    g.frame_stride = frame_stride; g.out_stride = out_stride; g.row_pitch = width * 3;
    g.x_min = x_min; g.y_min = y_min; g.rows = rows; g.cols = cols;
    g.tiles_x = tiles_x; g.groups_y = groups_y; g.n_frames = n_frames;
    if (hsv)
        roi_extract_kernel<true><<<grid, ROI_THREADS, 0, as_stream(stream)>>>(d_bgr, d_out, g, sx, sy, sf);
    else
        roi_extract_kernel<false><<<grid, ROI_THREADS, 0, as_stream(stream)>>>(d_bgr, d_out, g, sx, sy, sf);
    RTI_LAUNCHED("roi_extract_kernel");
    return RTI_OK;
}
