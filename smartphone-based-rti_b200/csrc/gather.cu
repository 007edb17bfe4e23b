// compute_intensities (analysis.py:196-246): frame-major uint8 ROIs + camera
// positions -> pixel-major SoA buffers (lx f32, ly f32, intensity i32), light index
// fastest.  This is a transpose with 12 bytes written per byte read, bound by HBM
// (13 B per pixel.light).
#include "common.cuh"

namespace rti {

// A CTA owns 32 consecutive pixels and walks the light axis in chunks of 128 frames, so
// everything it writes is one contiguous 32*n*4-byte region per array: partial sectors at
// the seam between two chunks are completed by the same CTA a few microseconds later and
// merge in L2 instead of going to DRAM as masked writes.
//  phase 1: thread (pixel lane, frame quad) reads 4 bytes from 4 frames (each warp
//           load is one fully used 32-byte sector), evaluates the light vectors in
//           fp64 and parks a float4/int4 per array in shared memory; the bytes and the
//           camera positions of chunk c+1 are fetched into registers before chunk c is
//           evaluated;
//  phase 2: each warp owns pixels and streams their 128 values per array out as one
//           512-byte run of 16-byte stores.
// Shared memory is addressed in 16-byte chunks, chunk (p, q) stored at p*32 + (q ^ p):
// phase 1 (fixed q, lanes = p) and phase 2 (fixed p, lanes = q) both touch 32 distinct
// chunks per warp access -> conflict free.
// (A warp-private variant without block barriers, 16 frames per step, was measured 2.5x
// slower: its 64-byte runs per pixel and store cost more than the barriers save.)
constexpr int G_PIX = 32;
constexpr int G_FRAMES = 128;
constexpr int G_THREADS = 256;
constexpr int G_WARPS = G_THREADS / 32;
constexpr int G_QPW = 32 / G_WARPS;                  // frame quads per warp and chunk (4)
constexpr size_t G_SMEM = 3 * G_PIX * 32 * 16 + G_FRAMES * 3 * sizeof(double);

template <bool VEC>
__global__ void __launch_bounds__(G_THREADS, 4)
gather_soa_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride,
                  const double *__restrict__ cams, int n, int row0, int cols, int64_t n_pixels,
                  float *__restrict__ lx, float *__restrict__ ly, int32_t *__restrict__ inten) {
    extern __shared__ __align__(16) unsigned char s_raw[];     // 51 KB: above the static limit
    float4 *s_lx = reinterpret_cast<float4 *>(s_raw);
    float4 *s_ly = s_lx + G_PIX * 32;
    int4 *s_in = reinterpret_cast<int4 *>(s_ly + G_PIX * 32);
    double (*s_cam)[3] = reinterpret_cast<double (*)[3]>(s_in + G_PIX * 32);
    double *s_camflat = &s_cam[0][0];

    const int64_t pix0 = (int64_t)blockIdx.x * G_PIX;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t p = pix0 + lane;
    const int64_t pc = p < n_pixels ? p : n_pixels - 1;
    const int yl = (int)(pc / cols);
    const double xd = (double)(int)(pc - (int64_t)yl * cols);
    const double yd = (double)(row0 + yl);
    const uint8_t *src = frames + pc;

    // 16 byte loads per thread and chunk, all issued before anything waits on them
    // (kept as 16 separate registers: packing them would wait on the loads right here)
    auto fetch = [&](int f0, uint8_t (&raw)[G_QPW][4]) {
#pragma unroll
        for (int i = 0; i < G_QPW; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int f = f0 + (warp + i * G_WARPS) * 4 + j;
                raw[i][j] = f < n ? __ldg(src + (int64_t)f * frame_stride) : (uint8_t)0;
            }
    };
    // camera table of a chunk: 3 * 128 doubles = 1.5 per thread, prefetched the same way
    auto cam_fetch = [&](int f0, double (&c)[2]) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int idx = threadIdx.x + i * G_THREADS;
            c[i] = (idx < 3 * G_FRAMES && f0 * 3 + idx < 3 * n) ? __ldg(cams + (int64_t)f0 * 3 + idx) : 1.0;
        }
    };
    auto cam_park = [&](const double (&c)[2]) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int idx = threadIdx.x + i * G_THREADS;
            if (idx < 3 * G_FRAMES) s_camflat[idx] = (idx % 3 == 2) ? c[i] * c[i] : c[i];   // z is used squared
        }
    };

    uint8_t cur[G_QPW][4], nxt[G_QPW][4];
    double ccam[2];
    cam_fetch(0, ccam);
    fetch(0, cur);
    cam_park(ccam);
    for (int f0 = 0; f0 < n; f0 += G_FRAMES) {
        const int fc = min(G_FRAMES, n - f0);
        const bool has_next = f0 + G_FRAMES < n;
        if (has_next) {
            fetch(f0 + G_FRAMES, nxt);
            cam_fetch(f0 + G_FRAMES, ccam);
        }
        __syncthreads();        // this chunk's camera table is parked; the previous phase 2 has drained
        // ---- phase 1: light vectors in fp64, parked as float4/int4 chunks
#pragma unroll
        for (int i = 0; i < G_QPW; ++i) {
            const int q = warp + i * G_WARPS;
            float ux[4], uy[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int f = min(q * 4 + j, fc - 1);
                light_vector(s_cam[f][0], s_cam[f][1], s_cam[f][2], xd, yd, ux[j], uy[j]);
            }
            const int slot = lane * 32 + (q ^ lane);
            s_lx[slot] = make_float4(ux[0], ux[1], ux[2], ux[3]);
            s_ly[slot] = make_float4(uy[0], uy[1], uy[2], uy[3]);
            s_in[slot] = make_int4(cur[i][0], cur[i][1], cur[i][2], cur[i][3]);
        }
        __syncthreads();
        if (has_next) cam_park(ccam);        // every phase-1 read of s_cam is behind the barrier
        {   // ---- phase 2: lane = frame quad
            const int q = lane;
            const int f = f0 + q * 4;
            for (int pl = warp; pl < G_PIX; pl += G_WARPS) {
                const int64_t po = pix0 + pl;
                if (po >= n_pixels) break;
                const int slot = pl * 32 + (q ^ pl);
                const float4 a = s_lx[slot];
                const float4 b = s_ly[slot];
                const int4 c = s_in[slot];
                const int64_t o = po * n + f;
                if (VEC) {
                    if (f < n) {             // n % 4 == 0: a quad is entirely inside or outside
                        __stcs(reinterpret_cast<float4 *>(lx + o), a);
                        __stcs(reinterpret_cast<float4 *>(ly + o), b);
                        __stcs(reinterpret_cast<int4 *>(inten + o), c);
                    }
                } else {
                    const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
                    const int cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (f + j < n) {
                            lx[o + j] = av[j];
                            ly[o + j] = bv[j];
                            inten[o + j] = cv[j];
                        }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < G_QPW; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) cur[i][j] = nxt[i][j];
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
    RTI_REQUIRE(bx <= 0x7fffffffLL, "rti_gather_soa: problem too large for one launch");
    const bool vec = (n % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_lx) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(d_ly) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(d_inten) & 15) == 0);
    // per device and idempotent; cheap enough to repeat on every call
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM));
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM));
    if (vec)
        gather_soa_kernel<true><<<(unsigned)bx, G_THREADS, G_SMEM, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    else
        gather_soa_kernel<false><<<(unsigned)bx, G_THREADS, G_SMEM, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    RTI_LAUNCHED("gather_soa_kernel");
    return RTI_OK;
}
