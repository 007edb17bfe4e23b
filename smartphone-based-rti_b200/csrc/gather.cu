// compute_intensities (analysis.py:196-246): frame-major uint8 ROIs + camera
// positions -> pixel-major SoA buffers (lx f32, ly f32, intensity i32), light index
// fastest.  This is a transpose with 12 bytes written per byte read, bound by HBM
// (13 B per pixel.light).
#include <stdlib.h>
#include "tma.cuh"

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
// ASYNC: the frame bytes of a chunk ([128 frames][32 pixels] = 4 KB) come through a three-stage cp.async
// ring, two chunks ahead, instead of one chunk ahead in registers: under this kernel's own write
// stream a global read takes thousands of cycles and ncu showed half of the warps waiting on the
// register prefetch (long scoreboard 4.1 warps per issue).  Needs 16-byte aligned 32-pixel runs.
constexpr int G_STAGES = 3;
constexpr size_t G_RING = (size_t)G_STAGES * G_FRAMES * G_PIX;
constexpr size_t G_SMEM_ASYNC = G_SMEM + G_RING;

template <bool VEC, bool ASYNC>
__global__ void __launch_bounds__(G_THREADS, ASYNC ? 3 : 4)
gather_soa_kernel(const uint8_t *__restrict__ frames, int64_t frame_stride,
                  const double *__restrict__ cams, int n, int row0, int cols, int64_t n_pixels,
                  float *__restrict__ lx, float *__restrict__ ly, int32_t *__restrict__ inten) {
    extern __shared__ __align__(16) unsigned char s_raw[];     // 51 KB: above the static limit
    float4 *s_lx = reinterpret_cast<float4 *>(s_raw);
    float4 *s_ly = s_lx + G_PIX * 32;
    int4 *s_in = reinterpret_cast<int4 *>(s_ly + G_PIX * 32);
    double (*s_cam)[3] = reinterpret_cast<double (*)[3]>(s_in + G_PIX * 32);
    double *s_camflat = &s_cam[0][0];
    uint8_t *s_ring = reinterpret_cast<uint8_t *>(s_camflat + 3 * G_FRAMES);   // ASYNC only: [stage][frame][pixel]

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

    // ASYNC: thread t copies half h = t & 1 of frame f0 + (t >> 1)'s 32-byte pixel run
    const int valid_bytes = (int)min((int64_t)G_PIX, n_pixels - pix0);
    auto issue = [&](int f0, int stage) {
        const int f = f0 + (threadIdx.x >> 1), h = threadIdx.x & 1;
        const int nb = min(16, valid_bytes - 16 * h);
        if (f < n && nb > 0) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(s_ring + ((size_t)stage * G_FRAMES + (threadIdx.x >> 1)) * G_PIX + 16 * h);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst),
                         "l"(frames + (int64_t)f * frame_stride + pix0 + 16 * h), "r"(nb) : "memory");
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };

    uint8_t cur[G_QPW][4], nxt[G_QPW][4];
    double ccam[2];
    cam_fetch(0, ccam);
    if (ASYNC) {
        issue(0, 0);
        issue(G_FRAMES, 1);
    } else {
        fetch(0, cur);
    }
    cam_park(ccam);
    int stage = 0;
    for (int f0 = 0; f0 < n; f0 += G_FRAMES) {
        const int fc = min(G_FRAMES, n - f0);
        const bool has_next = f0 + G_FRAMES < n;
        if (has_next) {
            if (!ASYNC) fetch(f0 + G_FRAMES, nxt);
            cam_fetch(f0 + G_FRAMES, ccam);
        }
        if (ASYNC) asm volatile("cp.async.wait_group 1;\n" ::: "memory");     // this chunk's bytes have landed
        __syncthreads();        // this chunk's camera table is parked; the previous phase 2 has drained
        if (ASYNC) {
            issue(f0 + 2 * G_FRAMES, stage == 0 ? 2 : stage - 1);              // the stage of the previous chunk
            const uint8_t *bytes = s_ring + (size_t)stage * G_FRAMES * G_PIX + lane;
#pragma unroll
            for (int i = 0; i < G_QPW; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int fl = (warp + i * G_WARPS) * 4 + j;
                    cur[i][j] = (fl < fc && lane < valid_bytes) ? bytes[fl * G_PIX] : (uint8_t)0;
                }
            stage = stage == 2 ? 0 : stage + 1;
        }
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
        if (VEC) {   // ---- phase 2: lane = frame quad, 16-byte stores (n % 4 == 0)
            const int q = lane;
            const int f = f0 + q * 4;
            for (int pl = warp; pl < G_PIX; pl += G_WARPS) {
                const int64_t po = pix0 + pl;
                if (po >= n_pixels) break;
                const int slot = pl * 32 + (q ^ pl);
                const int64_t o = po * n + f;
                if (f < n) {                 // a quad is entirely inside or outside
                    __stcs(reinterpret_cast<float4 *>(lx + o), s_lx[slot]);
                    __stcs(reinterpret_cast<float4 *>(ly + o), s_ly[slot]);
                    __stcs(reinterpret_cast<int4 *>(inten + o), s_in[slot]);
                }
            }
        } else {     // ---- phase 2, any n: lane = frame, each warp store is one 128-byte run of a pixel's row
            const float *f_lx = reinterpret_cast<const float *>(s_lx);
            const float *f_ly = reinterpret_cast<const float *>(s_ly);
            const int *f_in = reinterpret_cast<const int *>(s_in);
            for (int pl = warp; pl < G_PIX; pl += G_WARPS) {
                const int64_t po = pix0 + pl;
                if (po >= n_pixels) break;
#pragma unroll
                for (int it = 0; it < G_FRAMES / 32; ++it) {
                    const int fl = it * 32 + lane;                       // frame within the chunk
                    const int idx = (pl * 32 + ((fl >> 2) ^ pl)) * 4 + (fl & 3);
                    const int64_t o = po * n + f0 + fl;
                    if (f0 + fl < n) {
                        __stcs(lx + o, f_lx[idx]);
                        __stcs(ly + o, f_ly[idx]);
                        __stcs(inten + o, f_in[idx]);
                    }
                }
            }
        }
        if (!ASYNC) {
#pragma unroll
            for (int i = 0; i < G_QPW; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) cur[i][j] = nxt[i][j];
        }
    }
    if (ASYNC) asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

// ---- TMA-store variant (n % 4 == 0, 16-byte aligned outputs) ---------------------------------
// Same CTA = 32 pixels walking the light axis, but the transposed tile never goes back through
// the register file: phase 1 parks float4/int4 chunks in 128B-swizzled [32 pixels][32 frames]
// boxes and six elected threads hand each box to the TMA unit (cp.async.bulk.tensor store into
// the [P][n] arrays, clipped at n and P by the tensor map).  Two smem buffers alternate, so the
// drain of chunk c overlaps the light vectors of chunk c+1, with one block barrier per chunk.
constexpr int T_PIX = 32;
constexpr int T_FRAMES = 64;                      // frames per chunk = 2 boxes of 32 per array
constexpr int T_THREADS = 256;
constexpr int T_BOX = 32 * 128;                   // 4 KB
constexpr int T_ARRAY = (T_FRAMES / 32) * T_BOX;  // 8 KB per array and buffer
constexpr int T_BUF = 3 * T_ARRAY;                // 24 KB
constexpr size_t T_SMEM = 2 * T_BUF + 1024 + 2 * T_FRAMES * 3 * sizeof(double);

__global__ void __launch_bounds__(T_THREADS, 4)
gather_soa_tma_kernel(const __grid_constant__ CUtensorMap map_lx, const __grid_constant__ CUtensorMap map_ly,
                      const __grid_constant__ CUtensorMap map_in, const uint8_t *__restrict__ frames,
                      int64_t frame_stride, const double *__restrict__ cams, int n, int row0, int cols,
                      int64_t n_pixels) {
    extern __shared__ unsigned char t_raw[];
    const uint32_t base = (smem_u32(t_raw) + 1023u) & ~1023u;
    double *s_cam = reinterpret_cast<double *>(t_raw + (base - smem_u32(t_raw)) + 2 * T_BUF);   // [2][T_FRAMES*3]

    const int64_t pix0 = (int64_t)blockIdx.x * T_PIX;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t p = pix0 + lane;
    const int64_t pc = p < n_pixels ? p : n_pixels - 1;
    const int yl = (int)(pc / cols);
    const double xd = (double)(int)(pc - (int64_t)yl * cols);
    const double yd = (double)(row0 + yl);
    const uint8_t *src = frames + pc;

    // quad q = warp + 8*i covers frames f0 + 4q .. 4q+3: box i, 16-byte chunk `warp` of row `lane`
    auto fetch = [&](int f0, uint8_t (&raw)[2][4]) {
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int f = f0 + (warp + i * 8) * 4 + j;
                raw[i][j] = f < n ? __ldg(src + (int64_t)f * frame_stride) : (uint8_t)0;
            }
    };
    auto cam_fetch = [&](int f0) -> double {
        const int idx = threadIdx.x;
        return (idx < 3 * T_FRAMES && f0 * 3 + idx < 3 * n) ? __ldg(cams + (int64_t)f0 * 3 + idx) : 1.0;
    };
    auto cam_park = [&](int b, double c) {
        const int idx = threadIdx.x;
        if (idx < 3 * T_FRAMES) s_cam[b * (3 * T_FRAMES) + idx] = (idx % 3 == 2) ? c * c : c;   // z is used squared
    };
    const bool issuer = lane == 0 && warp < 6;
    const CUtensorMap *my_map = warp < 2 ? &map_lx : (warp < 4 ? &map_ly : &map_in);
    if (issuer) prefetch_tmap(my_map);

    // One chunk: prefetch the next chunk's bytes into `nxt`, evaluate this chunk from `cur`.  The
    // loop below alternates two register sets instead of copying nxt -> cur: a copy would wait
    // for the prefetch right away and serialise the loads with the arithmetic.
    int b = 0;
    double ccam;
    auto chunk = [&](int f0, const uint8_t (&cur)[2][4], uint8_t (&nxt)[2][4]) {
        const int fc = min(T_FRAMES, n - f0);
        const bool has_next = f0 + T_FRAMES < n;
        if (has_next) {
            fetch(f0 + T_FRAMES, nxt);
            // the camera chunk of the NEXT iteration was requested one iteration ago (its slot was last
            // read before the previous barrier); the request for the one after goes out now, so neither
            // ever waits a full L2 round trip inside one chunk
            cam_park(b ^ 1, ccam);
            ccam = cam_fetch(f0 + 2 * T_FRAMES);
        }
        const double *cam = s_cam + b * (3 * T_FRAMES);
        const uint32_t buf = base + b * T_BUF;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            float ux[4], uy[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int f = min((warp + i * 8) * 4 + j, fc - 1);
                light_vector(cam[3 * f], cam[3 * f + 1], cam[3 * f + 2], xd, yd, ux[j], uy[j]);
            }
            const uint32_t dst = buf + i * T_BOX + lane * 128 + ((warp ^ (lane & 7)) << 4);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "f"(ux[0]), "f"(ux[1]), "f"(ux[2]),
                         "f"(ux[3]) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + T_ARRAY), "f"(uy[0]), "f"(uy[1]),
                         "f"(uy[2]), "f"(uy[3]) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 2 * T_ARRAY), "r"((int)cur[i][0]),
                         "r"((int)cur[i][1]), "r"((int)cur[i][2]), "r"((int)cur[i][3]) : "memory");
        }
        fence_async_smem();
        if (issuer) tma_store_wait_read<0>();     // the boxes of the other buffer have been read out
        __syncthreads();
        if (issuer) {
            const int arr = warp >> 1, sub = warp & 1;
            if (f0 + sub * 32 < n) {
                tma_store_2d(my_map, buf + arr * T_ARRAY + sub * T_BOX, f0 + sub * 32, (int)pix0);
                tma_store_commit();
            }
        }
        b ^= 1;
    };

    uint8_t ra[2][4], rb[2][4];
    cam_park(0, cam_fetch(0));
    ccam = cam_fetch(T_FRAMES);
    fetch(0, ra);
    __syncthreads();
    for (int f0 = 0; f0 < n; f0 += 2 * T_FRAMES) {
        chunk(f0, ra, rb);
        if (f0 + T_FRAMES < n) chunk(f0 + T_FRAMES, rb, ra);
    }
    if (issuer) tma_store_wait_read<0>();
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
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM));
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM));
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM_ASYNC));
    RTI_CUDA(cudaFuncSetAttribute(gather_soa_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM_ASYNC));
    const char *env_async = getenv("RTI_GATHER_ASYNC");
    const bool async = (env_async == nullptr || env_async[0] != '0') && frame_stride % 16 == 0 &&
                       (reinterpret_cast<uintptr_t>(d_frames) & 15) == 0;
    const char *env_tma = getenv("RTI_GATHER_TMA");
    const bool want_tma = env_tma == nullptr || env_tma[0] != '0';
    // TMA-store kernel when every pixel row starts on a 32-byte sector (n % 8 == 0): 5.0 TB/s against 4.2 for
    // the register path; with n % 8 == 4 every other row is half a sector off, the boxes straddle sectors and
    // the register path is the faster one (measured, profiles/).
    if (want_tma && vec && n % 8 == 0 && n_pixels <= 0x7fffffffLL && encode_tiled_fn() != nullptr) {
        CUtensorMap map_lx, map_ly, map_in;
        int rc;
        if ((rc = make_map_2d(&map_lx, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, d_lx, n, n_pixels, (uint64_t)n * 4, 32, 32, "lx")))
            return rc;
        if ((rc = make_map_2d(&map_ly, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, d_ly, n, n_pixels, (uint64_t)n * 4, 32, 32, "ly")))
            return rc;
        if ((rc = make_map_2d(&map_in, CU_TENSOR_MAP_DATA_TYPE_INT32, d_inten, n, n_pixels, (uint64_t)n * 4, 32, 32,
                              "intensity")))
            return rc;
        RTI_CUDA(cudaFuncSetAttribute(gather_soa_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_SMEM));
        gather_soa_tma_kernel<<<(unsigned)bx, T_THREADS, T_SMEM, as_stream(stream)>>>(
            map_lx, map_ly, map_in, d_frames, frame_stride, d_cams, n, row0, cols, n_pixels);
    } else if (vec && async)
        gather_soa_kernel<true, true><<<(unsigned)bx, G_THREADS, G_SMEM_ASYNC, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    else if (vec)
        gather_soa_kernel<true, false><<<(unsigned)bx, G_THREADS, G_SMEM, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    else if (async)
        gather_soa_kernel<false, true><<<(unsigned)bx, G_THREADS, G_SMEM_ASYNC, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    else
        gather_soa_kernel<false, false><<<(unsigned)bx, G_THREADS, G_SMEM, as_stream(stream)>>>(
            d_frames, frame_stride, d_cams, n, row0, cols, n_pixels, d_lx, d_ly, d_inten);
    RTI_LAUNCHED("gather_soa_kernel");
    return RTI_OK;
}
