// compute_intensities (analysis.py:196-246): frame-major uint8 ROIs + camera
// positions -> pixel-major SoA buffers (lx f32, ly f32, intensity i32), light index
// fastest.  This is a transpose with 12 bytes written per byte read, bound by HBM
// (13 B per pixel.light).
#include "common.cuh"

namespace rti {

// Tile = 32 consecutive pixels x 128 consecutive frames.
//  phase 1: thread (pixel lane, frame quad) reads 4 bytes from 4 frames (each warp
//           load is one fully used 32-byte sector), evaluates the light vectors in
//           fp64 and parks a float4/int4 per array in shared memory;
//  phase 2: each warp owns pixels and streams their 128 values per array out as one
//           512-byte run of 16-byte stores.
// Shared memory is addressed in 16-byte chunks, chunk (p, q) stored at p*32 + (q ^ p):
// phase 1 (fixed q, lanes = p) and phase 2 (fixed p, lanes = q) both touch 32 distinct
// chunks per warp access -> conflict free.
constexpr int G_PIX = 32;
constexpr int G_FRAMES = 128;
constexpr int G_THREADS = 256;
constexpr size_t G_SMEM = 3 * G_PIX * 32 * 16 + G_FRAMES * 3 * sizeof(double);

template <bool VEC>
__global__ void __launch_bounds__(G_THREADS)
gather_soa_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride,
                  const double *__restrict__ cams, int n, int row0, int cols, int64_t n_pixels,
                  float *__restrict__ lx, float *__restrict__ ly, int32_t *__restrict__ inten) {
    extern __shared__ __align__(16) unsigned char s_raw[];     // 51 KB: above the static limit
    float4 *s_lx = reinterpret_cast<float4 *>(s_raw);
    float4 *s_ly = s_lx + G_PIX * 32;
    int4 *s_in = reinterpret_cast<int4 *>(s_ly + G_PIX * 32);
    double (*s_cam)[3] = reinterpret_cast<double (*)[3]>(s_in + G_PIX * 32);

    const int64_t pix0 = (int64_t)blockIdx.x * G_PIX;
    const int f0 = blockIdx.y * G_FRAMES;
    const int fc = min(G_FRAMES, n - f0);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int QPW = 32 / (G_THREADS / 32);          // frame quads per warp (4)

    // ---- phase 1a: all 16 byte loads of this thread are issued before anything waits on
    // them (each warp load is one fully used 32-byte sector; 16 sectors in flight per warp)
    const int64_t p = pix0 + lane;
    const int64_t pc = p < n_pixels ? p : n_pixels - 1;
    const int yl = (int)(pc / cols);
    const double xd = (double)(int)(pc - (int64_t)yl * cols);
    const double yd = (double)(row0 + yl);
    const uint8_t *src = frames + pc;
    uint8_t raw[QPW][4];
#pragma unroll
    for (int i = 0; i < QPW; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int f = (warp + i * (G_THREADS / 32)) * 4 + j;
            raw[i][j] = f < fc ? __ldg(src + (int64_t)(f0 + f) * frame_stride) : (uint8_t)0;
        }
    for (int i = threadIdx.x; i < fc; i += G_THREADS) {
        const double cz = cams[(int64_t)(f0 + i) * 3 + 2];
        s_cam[i][0] = cams[(int64_t)(f0 + i) * 3 + 0];
        s_cam[i][1] = cams[(int64_t)(f0 + i) * 3 + 1];
        s_cam[i][2] = cz * cz;
    }
    __syncthreads();

    // ---- phase 1b: light vectors in fp64, parked as float4/int4 chunks
#pragma unroll
    for (int i = 0; i < QPW; ++i) {
        const int q = warp + i * (G_THREADS / 32);
        float ux[4], uy[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int f = min(q * 4 + j, fc - 1);
            light_vector(s_cam[f][0], s_cam[f][1], s_cam[f][2], xd, yd, ux[j], uy[j]);
        }
        const int slot = lane * 32 + (q ^ lane);
        s_lx[slot] = make_float4(ux[0], ux[1], ux[2], ux[3]);
        s_ly[slot] = make_float4(uy[0], uy[1], uy[2], uy[3]);
        s_in[slot] = make_int4(raw[i][0], raw[i][1], raw[i][2], raw[i][3]);
    }
    __syncthreads();
    {   // ---- phase 2: lane = frame quad
        const int q = lane;
        const int f = f0 + q * 4;
        for (int pl = warp; pl < G_PIX; pl += G_THREADS / 32) {
            const int64_t po = pix0 + pl;
            if (po >= n_pixels) break;
            const int slot = pl * 32 + (q ^ pl);
            const float4 a = s_lx[slot];
            const float4 b = s_ly[slot];
            const int4 c = s_in[slot];
            const int64_t o = po * n + f;
            if (VEC && f + 3 < n) {
                __stcs(reinterpret_cast<float4 *>(lx + o), a);
                __stcs(reinterpret_cast<float4 *>(ly + o), b);
                __stcs(reinterpret_cast<int4 *>(inten + o), c);
            } else {
                const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
                const int cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (f + j < n && q * 4 + j < fc) {
                        lx[o + j] = av[j];
                        ly[o + j] = bv[j];
                        inten[o + j] = cv[j];
                    }
            }
        }
    }
}

}  // namespace rti

using namespace rti;

extern "C" int rti_gather_soa(const uint8_t *d_frames, int64_t frame_stride, const double *d_cams, int n,
                              int row0, int rows, int cols, float *d_lx, float *d_ly, int32_t *d_inten,
                              rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_cams && d_lx && d_ly && d_inten, "rti_gather_soa: NULL pointer");
    RTI_REQUIRE(n > 0 && rows > 0 && cols > 0 && row0 >= 0, "rti_gather_soa: bad sizes");
    RTI_REQUIRE(frame_stride >= (int64_t)rows * cols, "rti_gather_soa: frame_stride < slab pixels");
    const int64_t n_pixels = (int64_t)rows * cols;
    const int64_t bx = (n_pixels + G_PIX - 1) / G_PIX;
    const unsigned by = (unsigned)((n + G_FRAMES - 1) / G_FRAMES);
    RTI_REQUIRE(bx <= 0x7fffffffLL && by <= 65535u, "rti_gather_soa: problem too large for one launch");
    const bool vec = (n % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_lx) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(d_ly) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(d_inten) & 15) == 0);
    // per device and idempotent; cheap enough to repeat on every call
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM));
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM));
    if (vec)
        gather_soa_kernel<true><<<dim3((unsigned)bx, by), G_THREADS, G_SMEM, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    else
        gather_soa_kernel<false><<<dim3((unsigned)bx, by), G_THREADS, G_SMEM, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    RTI_LAUNCHED("gather_soa_kernel");
    return RTI_OK;
}
