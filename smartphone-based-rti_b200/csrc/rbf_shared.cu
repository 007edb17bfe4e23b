// Linear RBF, shared-light case: every pixel sees the same N nodes (lx_i, ly_i), so scipy's
// per-pixel  w = solve(A, d);  out = cdist(grid, nodes) @ w   (analysis.py:249-260 -> scipy
// _rbf.py:270-292) collapses to ONE dense contraction over all pixels
//
//     table[G*G][P] = M[G*G][N] . D[N][P],      M = Phi_grid . A^-1   (fp64, built once)
//
// (and weights[N][P] = A^-1 . D with the same kernel).  That contraction is the only
// GEMM-shaped piece of the whole path, so it is the only one on the tcgen05 tensor cores:
//
//   * D = the uint8 frame intensities, exact in fp16;
//   * every row of M is scaled by a power of two and split into fp16 hi + lo (22 mantissa
//     bits, absolute error 2^-22 of the row maximum), both accumulated into the same fp32
//     TMEM accumulator: 2 MMAs per k-step instead of the 3 of a bf16x3 split;
//   * CTA tile 128 operator rows x 256 pixels, K stepped 64 lights per stage; TMA
//     (SWIZZLE_128B) feeds a 3-stage smem ring: A_hi, A_lo K-major, the frame tile MN-major
//     (pixels contiguous, exactly how the frames lie in HBM - no transpose anywhere);
//   * warp 0 = TMA producer, warp 1 = MMA issuer (one thread), warps 2-5 = epilogue:
//     tcgen05.ld 32 columns at a time, unscale (a power of two: exact in fp32), truncate toward zero like the
//     reference's int32 element store (analysis.py:404-407), stage through swizzled smem
//     and TMA-store 32x32 int32 boxes straight into the [ly][lx] images of the table;
//   * two 256-column TMEM accumulators, so the epilogue of tile t overlaps the MMAs of t+1;
//   * persistent grid, operator-row blocks fastest so the frame tile of a pixel block stays
//     in L2 across its 79 row blocks and the 20 MB operator never leaves L2.
#include <cuda_fp16.h>
#include "tma.cuh"

namespace rti {
namespace {

constexpr int BM = 128;                 // operator rows per tile   (UMMA M)
constexpr int BN = 256;                 // pixels per tile          (UMMA N)
constexpr int BK = 64;                  // lights per stage (64 fp16 = one 128-byte swizzle row)
constexpr int UK = 16;                  // lights per tcgen05.mma (fp16)
constexpr int STAGES = 3;
constexpr int A_TILE = BM * BK * 2;     // 16 KB (hi and lo each)
constexpr int B_BOX = 64 * BK * 2;      // 8 KB: 64 pixels x 64 lights
constexpr int B_TILE = (BN / 64) * B_BOX;          // 32 KB
constexpr int STAGE_BYTES = 2 * A_TILE + B_TILE;   // 64 KB
constexpr int EPI_BUF = 32 * 128;       // one 32x32 int32 box = 4 KB
constexpr int EPI_BYTES = 4 * 2 * EPI_BUF;         // 4 warps x double buffer
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int THREADS = 192;
constexpr int TMEM_COLS = 512;          // two fp32 accumulators of 256 columns

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem], fp16 operands, fp32 accumulator
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets lane (warp%4)*32 + t
__device__ __forceinline__ void tc_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- descriptors ------------------------------------------------------------------------
// Shared-memory matrix descriptor (sm_100 "version 1"), SWIZZLE_128B, tile base 1024-aligned:
//   [0,14) start >> 4, [16,30) leading byte offset >> 4, [32,46) stride byte offset >> 4,
//   [46,48) version = 1, [61,64) layout = 2 (128-byte swizzle).
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) |
           ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor, kind::f16: D fp32 (bit 4), A/B fp16 (format 0), A K-major,
// B MN-major (bit 16), N >> 3 at [17,23), M >> 4 at [24,29).
constexpr uint32_t IDESC = (1u << 4) | (1u << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
constexpr uint32_t IDESC_BK = IDESC & ~(1u << 16);      // B K-major as well (batched form: vectors stored [column][k])


// ---- CTA-pair form (tcgen05 cta_group::2) --------------------------------------------------
// Two CTAs of a cluster (the two SMs of a TPC) work on one 256-row x 256-pixel tile: each CTA stages ITS 128
// operator rows (hi and lo) and ITS half of the intensity tile (128 pixels), the leader's single thread issues
// M = 256 MMAs that read both halves, every CTA's TMEM receives its 128 rows of the accumulator.  Per CTA the
// intensity operand is filled and read at half the rate of the single-CTA kernel, which is what that kernel is
// short of (shared-memory bandwidth: TMA fill + operand reads + epilogue staging against 128 B/clk), and a stage
// shrinks from 64 KB to 48 KB: four stages instead of three.
constexpr int STAGES2 = 4;
constexpr int B_HALF = (BN / 2 / 64) * B_BOX;              // 16 KB
constexpr int STAGE2_BYTES = 2 * A_TILE + B_HALF;          // 48 KB
constexpr int SMEM2_BYTES = STAGES2 * STAGE2_BYTES + EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t IDESC2 = (1u << 4) | (1u << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {     // shared::cluster address
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA loads whose completion bytes are counted on a barrier of the pair's leader CTA
__device__ __forceinline__ void tma2_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma2_load_3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tc2_commit(uint32_t bar) {        // arrives on `bar` of BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ void tc2_mma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

// ---- the contraction --------------------------------------------------------------------
struct GemmArgs {
    int op_rows;               // R: rows of the operator (G*G for the table, N for weights)
    int n;                     // lights (K)
    long long n_pixels;        // P
    const float *inv_scale;    // [R] power-of-two unscale factors
    int32_t *table;            // MODE 1: int32 [R][image_stride]
    long long image_stride;
    double *out_f64;           // MODE 2: float64 [P][R] (pixel-major)
    int32_t *bad;              // pixels x rows whose value does not fit int32 (MODE 0/1)
    int num_m_blks, num_n_blks;
    // batched form (rbf_pcg.cu: one preconditioner per pixel tile): pixel block n_blk uses operator
    // n_blk / blks_per_op, whose rows start at that index times op_row_stride in the operator arrays and
    // in inv_scale; the operator's K extent wraps after a_num_k stages (B = [r_hi; r_lo] stacked along K
    // against one copy of the operator); tile_active[op] == 0 skips the operator's pixel blocks.
    int a_num_k, blks_per_op, op_row_stride;
    const int32_t *tile_active;
};

enum { STORE_TMA = 0, STORE_DIRECT = 1, STORE_F64_PIXEL_MAJOR = 2 };

template <int MODE, bool B_KMAJOR = false>
__global__ void __launch_bounds__(THREADS, 1)
rbf_shared_gemm_kernel(const __grid_constant__ CUtensorMap map_hi, const __grid_constant__ CUtensorMap map_lo,
                       const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_out,
                       const GemmArgs args) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t epi_base = base + STAGES * STAGE_BYTES;
    const uint32_t bar_base = epi_base + EPI_BYTES;
    const uint32_t bar_full = bar_base, bar_empty = bar_base + 8 * STAGES;
    const uint32_t bar_tfull = bar_base + 16 * STAGES, bar_tempty = bar_tfull + 16;
    const uint32_t tmem_slot = bar_tempty + 16;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_k = (args.n + BK - 1) / BK;
    const int num_tiles = args.num_m_blks * args.num_n_blks;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&map_hi);
        prefetch_tmap(&map_lo);
        prefetch_tmap(&map_b);
        if (MODE == STORE_TMA) prefetch_tmap(&map_out);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, 4);      // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                     "r"((uint32_t)TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m_blk = tile % args.num_m_blks, n_blk = tile / args.num_m_blks;
                const int op = n_blk / args.blks_per_op;
                if (args.tile_active != nullptr && args.tile_active[op] == 0) continue;
                const int a_row = op * args.op_row_stride + m_blk * BM;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    const uint32_t full = bar_full + 8 * stage;
                    const uint32_t sa = base + stage * STAGE_BYTES;
                    const int a_col = (kb % args.a_num_k) * BK;
                    mbar_expect_tx(full, STAGE_BYTES);
                    tma_load_2d(sa, &map_hi, a_col, a_row, full);
                    tma_load_2d(sa + A_TILE, &map_lo, a_col, a_row, full);
                    if (B_KMAJOR) {
                        // one box of 256 columns x 64 k (each column's k run is contiguous in HBM)
                        tma_load_2d(sa + 2 * A_TILE, &map_b, kb * BK, n_blk * BN, full);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_3d(sa + 2 * A_TILE + j * B_BOX, &map_b, j * 64, kb * BK, n_blk, full);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one thread drives the tensor core for the whole CTA =====
        if (lane == 0) {
            int stage = 0, acc = 0;
            uint32_t phase = 0, aphase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                if (args.tile_active != nullptr &&
                    args.tile_active[(tile / args.num_m_blks) / args.blks_per_op] == 0) continue;
                mbar_wait(bar_tempty + 8 * acc, aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after();
                    const uint32_t sa = base + stage * STAGE_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UK; ++k) {
                        // A (K-major): 128-byte rows, 8-row groups 1024 B apart; a k-step is 32 B along the row
                        const uint64_t da_hi = smem_desc(sa + k * (UK * 2), 16, 1024);
                        const uint64_t da_lo = smem_desc(sa + A_TILE + k * (UK * 2), 16, 1024);
                        // B (MN-major): 64-pixel atoms 8 KB apart (LBO), 8-light groups 1024 B apart (SBO);
                        // a k-step is two 8-light groups
                        // B K-major (batched form): the same layout as A, 256 rows
                        const uint64_t db = B_KMAJOR ? smem_desc(sa + 2 * A_TILE + k * (UK * 2), 16, 1024)
                                                     : smem_desc(sa + 2 * A_TILE + k * (UK * 128), B_BOX, 1024);
                        constexpr uint32_t idesc = B_KMAJOR ? IDESC_BK : IDESC;
                        tc_mma_f16(d_tmem, da_hi, db, idesc, (kb | k) != 0);
                        tc_mma_f16(d_tmem, da_lo, db, idesc, 1u);
                    }
                    tc_commit(bar_empty + 8 * stage);              // smem slot free once these MMAs retire
                    if (kb == num_k - 1) tc_commit(bar_tfull + 8 * acc);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                acc ^= 1;
                if (acc == 0) aphase ^= 1;
            }
        }
    } else {
        // ===== epilogue: TMEM -> registers -> (unscale, truncate) -> global =====
        const int q = warp & 3;                                    // TMEM lane quarter this warp may read
        int acc = 0;
        uint32_t aphase = 0;
        int my_bad = 0;
        int chunk_parity = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m_blk = tile % args.num_m_blks, n_blk = tile / args.num_m_blks;
            const int op = n_blk / args.blks_per_op;
            if (args.tile_active != nullptr && args.tile_active[op] == 0) continue;
            const int row = m_blk * BM + q * 32 + lane;
            const bool row_ok = row < args.op_rows;
            const float sc = row_ok ? __ldg(args.inv_scale + (long long)op * args.op_row_stride + row) : 0.0f;
            mbar_wait(bar_tfull + 8 * acc, aphase);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tc_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + c * 32, v);
                tc_wait_ld();
                if (c == BN / 32 - 1) {                            // accumulator drained: hand it back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
                }
                const long long p0 = (long long)n_blk * BN + c * 32;
                if (MODE == STORE_F64_PIXEL_MAJOR) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const long long p = p0 + j;
                        if (row_ok && p < args.n_pixels)
                            args.out_f64[p * args.op_rows + row] = (double)(__uint_as_float(v[j]) * sc);
                    }
                } else {
                    int iv[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float f = __uint_as_float(v[j]) * sc;        // exact: sc is a power of two
                        my_bad |= !(fabsf(f) < 2147483648.0f);
                        iv[j] = __float2int_rz(f);
                    }
                    if (MODE == STORE_TMA) {
                        const uint32_t buf = epi_base + (q * 2 + chunk_parity) * EPI_BUF;
                        if (lane == 0) tma_store_wait_read<1>();           // the store that last used buf is done reading
                        __syncwarp();
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const uint32_t dst = buf + lane * 128 + ((j ^ (lane & 7)) << 4);
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(iv[4 * j]),
                                         "r"(iv[4 * j + 1]), "r"(iv[4 * j + 2]), "r"(iv[4 * j + 3])
                                         : "memory");
                        }
                        fence_async_smem();
                        __syncwarp();
                        // Boxes wholly outside the table are skipped (partial ones are clipped by TMA).  The
                        // buffers alternate per ISSUED store: wait_group.read 1 above then guarantees that
                        // the store which last read `buf` has finished.
                        if (p0 < args.n_pixels && m_blk * BM + q * 32 < args.op_rows) {
                            if (lane == 0) {
                                tma_store_2d(&map_out, buf, (int)p0, m_blk * BM + q * 32);
                                tma_store_commit();
                            }
                            chunk_parity ^= 1;
                        }
                    } else {
                        int32_t *dst = args.table + (long long)row * args.image_stride + p0;
                        if (row_ok && p0 + 32 <= args.n_pixels && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                            // a lane owns one row: 32 consecutive pixels = one 128-byte line, as 8 streaming 16-byte stores
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                __stcs(reinterpret_cast<int4 *>(dst) + j,
                                       make_int4(iv[4 * j], iv[4 * j + 1], iv[4 * j + 2], iv[4 * j + 3]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const long long p = p0 + j;
                                if (row_ok && p < args.n_pixels) dst[j] = iv[j];
                            }
                        }
                    }
                }
            }
            acc ^= 1;
            if (acc == 0) aphase ^= 1;
        }
        if (MODE == STORE_TMA && lane == 0) tma_store_wait_all();
        if (MODE != STORE_F64_PIXEL_MAJOR && args.bad != nullptr && my_bad) atomicAdd(args.bad, 1);
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS)
                     : "memory");
    }
}


// CTA-pair variant of rbf_shared_gemm_kernel (shared-light tables; B MN-major).  Barrier protocol:
//   full[s]   leader only: one arrival (the leader's expect_tx of both CTAs' bytes); both CTAs' TMA loads
//             complete on it;          empty[s]  per CTA: the leader's commit arrives on both (multicast);
//   tfull[a]  per CTA: commit multicast; tempty[a] leader only: 8 arrivals, one per epilogue warp of both CTAs.
template <int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
rbf_shared_gemm2_kernel(const __grid_constant__ CUtensorMap map_hi, const __grid_constant__ CUtensorMap map_lo,
                        const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_out,
                        const GemmArgs args) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t epi_base = base + STAGES2 * STAGE2_BYTES;
    const uint32_t bar_base = epi_base + EPI_BYTES;
    const uint32_t bar_full = bar_base, bar_empty = bar_base + 8 * STAGES2;
    const uint32_t bar_tfull = bar_base + 16 * STAGES2, bar_tempty = bar_tfull + 16;
    const uint32_t tmem_slot = bar_tempty + 16;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();
    const bool leader = rank == 0;
    const int num_k = (args.n + BK - 1) / BK;
    const int num_m2 = (args.num_m_blks + 1) / 2;                    // 256-row blocks
    const int num_tiles = num_m2 * args.num_n_blks;
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&map_hi);
        prefetch_tmap(&map_lo);
        prefetch_tmap(&map_b);
        if (MODE == STORE_TMA) prefetch_tmap(&map_out);
        for (int s = 0; s < STAGES2; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, 8);      // four epilogue warps of each CTA
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                     "r"((uint32_t)TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();                            // both CTAs' barriers initialised, both allocations done
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    if (warp == 0) {
        // ===== TMA producer (both CTAs): own 128 operator rows, own 128 pixels =====
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                const int m2 = tile % num_m2, n_blk = tile / num_m2;
                const int a_row = m2 * 2 * BM + (int)rank * BM;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    const uint32_t full = map_to_cta(bar_full + 8 * stage, 0);       // the leader's barrier
                    const uint32_t sa = base + stage * STAGE2_BYTES;
                    if (leader) mbar_expect_tx(bar_full + 8 * stage, 2 * STAGE2_BYTES);
                    tma2_load_2d(sa, &map_hi, kb * BK, a_row, full);
                    tma2_load_2d(sa + A_TILE, &map_lo, kb * BK, a_row, full);
#pragma unroll
                    for (int j = 0; j < BN / 2 / 64; ++j)
                        tma2_load_3d(sa + 2 * A_TILE + j * B_BOX, &map_b, ((int)rank * (BN / 2 / 64) + j) * 64, kb * BK,
                                     n_blk, full);
                    if (++stage == STAGES2) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one thread of the LEADER drives both SMs' tensor cores =====
        if (lane == 0 && leader) {
            int stage = 0, acc = 0;
            uint32_t phase = 0, aphase = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                mbar_wait(bar_tempty + 8 * acc, aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(bar_full + 8 * stage, phase);
                    tc_fence_after();
                    const uint32_t sa = base + stage * STAGE2_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UK; ++k) {
                        const uint64_t da_hi = smem_desc(sa + k * (UK * 2), 16, 1024);
                        const uint64_t da_lo = smem_desc(sa + A_TILE + k * (UK * 2), 16, 1024);
                        const uint64_t db = smem_desc(sa + 2 * A_TILE + k * (UK * 128), B_BOX, 1024);
                        tc2_mma_f16(d_tmem, da_hi, db, IDESC2, (kb | k) != 0);
                        tc2_mma_f16(d_tmem, da_lo, db, IDESC2, 1u);
                    }
                    tc2_commit(bar_empty + 8 * stage);             // both CTAs' slots free once these MMAs retire
                    if (kb == num_k - 1) tc2_commit(bar_tfull + 8 * acc);
                    if (++stage == STAGES2) { stage = 0; phase ^= 1; }
                }
                acc ^= 1;
                if (acc == 0) aphase ^= 1;
            }
        }
    } else {
        // ===== epilogue (both CTAs): own 128 rows of the accumulator =====
        const int q = warp & 3;
        int acc = 0;
        uint32_t aphase = 0;
        int my_bad = 0;
        int chunk_parity = 0;
        for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
            const int m2 = tile % num_m2, n_blk = tile / num_m2;
            const int row0 = m2 * 2 * BM + (int)rank * BM + q * 32;
            const int row = row0 + lane;
            const bool row_ok = row < args.op_rows;
            const float sc = row_ok ? __ldg(args.inv_scale + row) : 0.0f;
            mbar_wait(bar_tfull + 8 * acc, aphase);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tc_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + c * 32, v);
                tc_wait_ld();
                if (c == BN / 32 - 1) {                            // accumulator drained: tell the leader
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(map_to_cta(bar_tempty + 8 * acc, 0));
                }
                const long long p0 = (long long)n_blk * BN + c * 32;
                if (MODE == STORE_F64_PIXEL_MAJOR) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const long long p = p0 + j;
                        if (row_ok && p < args.n_pixels)
                            args.out_f64[p * args.op_rows + row] = (double)(__uint_as_float(v[j]) * sc);
                    }
                } else {
                    int iv[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float f = __uint_as_float(v[j]) * sc;
                        my_bad |= !(fabsf(f) < 2147483648.0f);
                        iv[j] = __float2int_rz(f);
                    }
                    if (MODE == STORE_TMA) {
                        const uint32_t buf = epi_base + (q * 2 + chunk_parity) * EPI_BUF;
                        if (lane == 0) tma_store_wait_read<1>();
                        __syncwarp();
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const uint32_t dst = buf + lane * 128 + ((j ^ (lane & 7)) << 4);
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(iv[4 * j]),
                                         "r"(iv[4 * j + 1]), "r"(iv[4 * j + 2]), "r"(iv[4 * j + 3])
                                         : "memory");
                        }
                        fence_async_smem();
                        __syncwarp();
                        if (p0 < args.n_pixels && row0 < args.op_rows) {
                            if (lane == 0) {
                                tma_store_2d(&map_out, buf, (int)p0, row0);
                                tma_store_commit();
                            }
                            chunk_parity ^= 1;
                        }
                    } else {
                        int32_t *dst = args.table + (long long)row * args.image_stride + p0;
                        if (row_ok && p0 + 32 <= args.n_pixels && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                            // a lane owns one row: 32 consecutive pixels = one 128-byte line, as 8 streaming 16-byte stores
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                __stcs(reinterpret_cast<int4 *>(dst) + j,
                                       make_int4(iv[4 * j], iv[4 * j + 1], iv[4 * j + 2], iv[4 * j + 3]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const long long p = p0 + j;
                                if (row_ok && p < args.n_pixels) dst[j] = iv[j];
                            }
                        }
                    }
                }
            }
            acc ^= 1;
            if (acc == 0) aphase ^= 1;
        }
        if (MODE == STORE_TMA && lane == 0) tma_store_wait_all();
        if (MODE != STORE_F64_PIXEL_MAJOR && args.bad != nullptr && my_bad) atomicAdd(args.bad, 1);
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();                            // neither CTA frees TMEM while the other may still be written
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS)
                     : "memory");
    }
}


// ---- operand packing --------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t *>(&h);
}

// Packed intensity operand: pixel blocks of BN = 256, each a contiguous [n lights][256 pixels]
// fp16 slab, so that the 64 light rows of one pipeline stage lie in one 32 KB run (one page for
// the TMA unit) instead of 64 rows a whole frame pitch apart.
__device__ __forceinline__ long long packed_index(long long p, int i, int n) {
    return ((p / BN) * n + i) * BN + (p % BN);
}

// uint8 frames [n][frame_stride] -> packed fp16 (0..255 are exact), pad columns zero.
// One thread = 8 consecutive pixels of one frame: a warp reads 256 contiguous bytes and writes 512.
__global__ void pack_frames_kernel(const uint8_t *__restrict__ frames, long long frame_stride, int n,
                                   long long n_pixels, long long p_pad, uint16_t *__restrict__ out) {
    const long long p = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    const int i = blockIdx.y;
    if (p >= p_pad) return;
    const uint8_t *src = frames + (long long)i * frame_stride + p;
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = (p + j < n_pixels) ? (float)__ldg(src + j) : 0.0f;
    uint4 o;
    o.x = pack_half2(f[0], f[1]);
    o.y = pack_half2(f[2], f[3]);
    o.z = pack_half2(f[4], f[5]);
    o.w = pack_half2(f[6], f[7]);
    *reinterpret_cast<uint4 *>(out + packed_index(p, i, n)) = o;
}

// int32 intensities [P][n] (compute_intensities layout, analysis.py:219) -> packed fp16.
// fp16 holds integers exactly up to 2048 in magnitude; anything else is counted in *bad.
__global__ void pack_soa_kernel(const int32_t *__restrict__ inten, long long n_pixels, int n, long long p_pad,
                                uint16_t *__restrict__ out, int32_t *__restrict__ bad) {
    __shared__ float tile[32][33];
    const long long p0 = (long long)blockIdx.x * 32;
    const int i0 = blockIdx.y * 32;
    int my_bad = 0;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const long long p = p0 + r;
        const int i = i0 + threadIdx.x;
        int v = 0;
        if (p < n_pixels && i < n) v = __ldg(inten + p * n + i);
        my_bad |= (v > 2048) | (v < -2048);
        tile[r][threadIdx.x] = (float)v;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int i = i0 + r;
        const long long p = p0 + threadIdx.x;
        if (i < n && p < p_pad) out[packed_index(p, i, n)] = __half_as_ushort(__float2half_rn(tile[threadIdx.x][r]));
    }
    if (my_bad && bad != nullptr) atomicAdd(bad, 1);
}

// ---- operator construction ----------------------------------------------------------------
// n pseudo-pixels, all with the shared nodes; pseudo-pixel j carries the unit intensity e_j, so
// the per-pixel RBF fit returns column j of A^-1 and its grid evaluation column j of Phi.A^-1.
__global__ void replicate_nodes_kernel(const float *__restrict__ lx, const float *__restrict__ ly, int n,
                                       float *__restrict__ lx_rep, float *__restrict__ ly_rep,
                                       int32_t *__restrict__ ident) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (k >= n) return;
    lx_rep[(long long)j * n + k] = lx[k];
    ly_rep[(long long)j * n + k] = ly[k];
    ident[(long long)j * n + k] = (j == k);
}

// Row scale of the operator: src[j * src_stride + r], j < n.  scale_r = 2^(13 - ilogb(max_j |.|)),
// so the largest scaled entry lies in [2^13, 2^14): far from fp16 overflow, and the products
// with intensities <= 255 stay far below fp32 overflow.
__global__ void operator_scale_kernel(const double *__restrict__ src, long long src_stride, int n, int rows,
                                      double *__restrict__ scale, float *__restrict__ inv_scale,
                                      int32_t *__restrict__ bad) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    double m = 0.0;
    bool finite = true;
    for (int j = 0; j < n; ++j) {
        const double v = fabs(src[(long long)j * src_stride + r]);
        finite = finite && (v < 1.0e300);
        m = fmax(m, v);
    }
    int e = (m > 0.0 && finite) ? ilogb(m) : 13;
    e = max(-100, min(100, e));
    scale[r] = scalbn(1.0, 13 - e);
    inv_scale[r] = (float)scalbn(1.0, e - 13);
    if (!finite && bad != nullptr) atomicAdd(bad, 1);
}

// hi = fp16(v * scale_r), lo = fp16(v * scale_r - hi), written K-major [row][n_pad] (lights fastest)
__global__ void operator_split_kernel(const double *__restrict__ src, long long src_stride, int n, int rows,
                                      const double *__restrict__ scale, long long n_pad,
                                      uint16_t *__restrict__ hi, uint16_t *__restrict__ lo) {
    __shared__ double tile[32][33];
    const int r0 = blockIdx.x * 32, j0 = blockIdx.y * 32;
    for (int t = threadIdx.y; t < 32; t += blockDim.y) {
        const int j = j0 + t, r = r0 + threadIdx.x;
        tile[t][threadIdx.x] = (j < n && r < rows) ? src[(long long)j * src_stride + r] * scale[r] : 0.0;
    }
    __syncthreads();
    for (int t = threadIdx.y; t < 32; t += blockDim.y) {
        const int r = r0 + t, j = j0 + threadIdx.x;
        if (r < rows && j < n) {
            const double v = tile[threadIdx.x][t];
            const __half h = __double2half(v);
            const __half l = __double2half(v - (double)__half2float(h));
            hi[(long long)r * n_pad + j] = __half_as_ushort(h);
            lo[(long long)r * n_pad + j] = __half_as_ushort(l);
        }
    }
}

// ---- host side -----------------------------------------------------------------------------
int make_map_packed(CUtensorMap *map, const void *ptr, uint64_t n, uint64_t n_blks) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (fn == nullptr) {
        set_error("cuTensorMapEncodeTiled is not available from this driver");
        return RTI_ERR_CUDA;
    }
    const cuuint64_t dims[3] = {(cuuint64_t)BN, n, n_blks};
    const cuuint64_t strides[2] = {(cuuint64_t)BN * 2, n * BN * 2};
    const cuuint32_t box[3] = {64, (cuuint32_t)BK, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult rc = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void *>(ptr), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(packed intensities) failed with CUresult %d (n %llu, blocks %llu)", (int)rc,
                  (unsigned long long)n, (unsigned long long)n_blks);
        return RTI_ERR_CUDA;
    }
    return RTI_OK;
}

inline long long round_up(long long v, long long m) { return (v + m - 1) / m * m; }

int gemm_launch(const uint16_t *d_packed, int n, long long n_pixels, const uint16_t *d_op_hi, const uint16_t *d_op_lo,
                const float *d_inv_scale, int op_rows, int32_t *d_table, long long image_stride, double *d_out_f64,
                int32_t *d_bad, cudaStream_t stream) {
    const long long n_pad = round_up(n, BK), p_pad = round_up(n_pixels, BN), r_pad = round_up(op_rows, BM);
    CUtensorMap map_hi, map_lo, map_b, map_out;
    int rc;
    if ((rc = make_map_2d(&map_hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, d_op_hi, n_pad, r_pad, n_pad * 2, BK, BM, "op_hi")))
        return rc;
    if ((rc = make_map_2d(&map_lo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, d_op_lo, n_pad, r_pad, n_pad * 2, BK, BM, "op_lo")))
        return rc;
    if ((rc = make_map_packed(&map_b, d_packed, n, p_pad / BN))) return rc;
    int mode = STORE_F64_PIXEL_MAJOR;
    map_out = map_b;
    if (d_table != nullptr) {
        // (register stores straight to global, one 128-byte line per lane, were measured instead of the
        // swizzled staging + TMA store to relieve shared memory: 77 ms against 16 ms -- 32 lines per store
        // instruction; they remain the path of unaligned tables only)
        const bool tma_ok = (image_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_table) & 15) == 0);
        mode = tma_ok ? STORE_TMA : STORE_DIRECT;
        if (tma_ok && (rc = make_map_2d(&map_out, CU_TENSOR_MAP_DATA_TYPE_INT32, d_table, n_pixels, op_rows,
                                        image_stride * 4, 32, 32, "table")))
            return rc;
    }
    GemmArgs args;
    args.op_rows = op_rows;
    args.n = n;
    args.n_pixels = n_pixels;
    args.inv_scale = d_inv_scale;
    args.table = d_table;
    args.image_stride = image_stride;
    args.out_f64 = d_out_f64;
    args.bad = d_bad;
    args.num_m_blks = (int)(r_pad / BM);
    args.num_n_blks = (int)((n_pixels + BN - 1) / BN);
    args.a_num_k = (int)(n_pad / BK);
    args.blks_per_op = 0x7fffffff;
    args.op_row_stride = 0;
    args.tile_active = nullptr;
    int dev = 0, sms = 148;
    RTI_CUDA(cudaGetDevice(&dev));
    RTI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const long long tiles = (long long)args.num_m_blks * args.num_n_blks;
    RTI_REQUIRE(tiles < (1ll << 31), "rbf shared: too many tiles (%lld); split the pixel range", tiles);
    // CTA pairs (cta_group::2) unless RTI_GEMM_2CTA=0
    static const bool two_cta = !(getenv("RTI_GEMM_2CTA") && getenv("RTI_GEMM_2CTA")[0] == '0');
    if (two_cta) {
        const long long tiles2 = (long long)((args.num_m_blks + 1) / 2) * args.num_n_blks;
        // persistent pairs with a static tile stride: launch exactly as many clusters as can be resident at once
        // (a GPC with an odd number of usable SMs cannot pair its last one)
#define RTI_GEMM2_LAUNCH(M)                                                                                      \
    do {                                                                                                         \
        RTI_CUDA(cudaFuncSetAttribute(rbf_shared_gemm2_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                      SMEM2_BYTES));                                                             \
        cudaLaunchConfig_t cfg = {};                                                                             \
        cfg.gridDim = dim3(sms / 2 * 2);                                                                         \
        cfg.blockDim = dim3(THREADS);                                                                            \
        cfg.dynamicSmemBytes = SMEM2_BYTES;                                                                      \
        int max_clusters = 0;                                                                                    \
        RTI_CUDA(cudaOccupancyMaxActiveClusters(&max_clusters, rbf_shared_gemm2_kernel<M>, &cfg));               \
        if (max_clusters < 1) max_clusters = 1;                                                                  \
        const int clusters = (int)(tiles2 < max_clusters ? tiles2 : max_clusters);                               \
        rbf_shared_gemm2_kernel<M><<<2 * clusters, THREADS, SMEM2_BYTES, stream>>>(map_hi, map_lo, map_b, map_out, \
                                                                                  args);                         \
    } while (0)
        if (mode == STORE_TMA) RTI_GEMM2_LAUNCH(STORE_TMA);
        else if (mode == STORE_DIRECT) RTI_GEMM2_LAUNCH(STORE_DIRECT);
        else RTI_GEMM2_LAUNCH(STORE_F64_PIXEL_MAJOR);
#undef RTI_GEMM2_LAUNCH
        RTI_LAUNCHED("rbf_shared_gemm2_kernel");
        return RTI_OK;
    }
    const int grid = (int)(tiles < sms ? tiles : sms);
#define RTI_GEMM_LAUNCH(M)                                                                                       \
    do {                                                                                                         \
        RTI_CUDA(cudaFuncSetAttribute(rbf_shared_gemm_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize,    \
                                      SMEM_BYTES));                                                              \
        rbf_shared_gemm_kernel<M><<<grid, THREADS, SMEM_BYTES, stream>>>(map_hi, map_lo, map_b, map_out, args);  \
    } while (0)
    if (mode == STORE_TMA) RTI_GEMM_LAUNCH(STORE_TMA);
    else if (mode == STORE_DIRECT) RTI_GEMM_LAUNCH(STORE_DIRECT);
    else RTI_GEMM_LAUNCH(STORE_F64_PIXEL_MAJOR);
#undef RTI_GEMM_LAUNCH
    RTI_LAUNCHED("rbf_shared_gemm_kernel");
    return RTI_OK;
}

}  // namespace

// Batched contraction for the tile-preconditioned conjugate gradients of rbf_pcg.cu:
//   out[col][0:op_rows_pad] = (op_hi + op_lo)[col / (blks_per_op*256)] . ([B_hi; B_lo])[:, col]
// d_packed: fp16 [n_cols][k_rows] with k_rows = 2 * a_k_pad (each column: the hi halves then the lo halves
// of its vector, contiguous -- K-major B operand); operators: fp16 [n_ops][op_rows_pad][a_k_pad]; out float64 [n_cols][op_rows_pad].
int rbf_gemm_batched(const uint16_t *d_packed, int k_rows, long long n_cols, const uint16_t *d_op_hi,
                     const uint16_t *d_op_lo, const float *d_inv_scale, int n_ops, int op_rows_pad, int a_k_pad,
                     int blks_per_op, double *d_out, const int32_t *d_tile_active, cudaStream_t stream) {
    RTI_REQUIRE(op_rows_pad % BM == 0 && a_k_pad % BK == 0 && k_rows % BK == 0 && n_cols % BN == 0,
                "rbf_gemm_batched: unpadded shape");
    CUtensorMap map_hi, map_lo, map_b;
    int rc;
    const uint64_t rows_all = (uint64_t)n_ops * op_rows_pad;
    if ((rc = make_map_2d(&map_hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, d_op_hi, a_k_pad, rows_all, (uint64_t)a_k_pad * 2, BK, BM,
                          "pcg op_hi")))
        return rc;
    if ((rc = make_map_2d(&map_lo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, d_op_lo, a_k_pad, rows_all, (uint64_t)a_k_pad * 2, BK, BM,
                          "pcg op_lo")))
        return rc;
    if ((rc = make_map_2d(&map_b, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, d_packed, k_rows, n_cols, (uint64_t)k_rows * 2, BK, BN,
                          "pcg vectors")))
        return rc;
    GemmArgs args;
    args.op_rows = op_rows_pad;
    args.n = k_rows;
    args.n_pixels = n_cols;
    args.inv_scale = d_inv_scale;
    args.table = nullptr;
    args.image_stride = 0;
    args.out_f64 = d_out;
    args.bad = nullptr;
    args.num_m_blks = op_rows_pad / BM;
    args.num_n_blks = (int)(n_cols / BN);
    args.a_num_k = a_k_pad / BK;
    args.blks_per_op = blks_per_op;
    args.op_row_stride = op_rows_pad;
    args.tile_active = d_tile_active;
    int dev = 0, sms = 148;
    RTI_CUDA(cudaGetDevice(&dev));
    RTI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const long long tiles = (long long)args.num_m_blks * args.num_n_blks;
    RTI_REQUIRE(tiles < (1ll << 31), "rbf_gemm_batched: too many tiles");
    const int grid = (int)(tiles < sms ? tiles : sms);
    RTI_CUDA(cudaFuncSetAttribute(rbf_shared_gemm_kernel<STORE_F64_PIXEL_MAJOR, true>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    rbf_shared_gemm_kernel<STORE_F64_PIXEL_MAJOR, true><<<grid, THREADS, SMEM_BYTES, stream>>>(map_hi, map_lo, map_b, map_b,
                                                                                                 args);
    RTI_LAUNCHED("rbf_shared_gemm_kernel");
    return RTI_OK;
}

}  // namespace rti

// ---- C ABI ------------------------------------------------------------------------------------
using namespace rti;

extern "C" int64_t rti_rbf_shared_operator_elems(int op_rows, int n) {
    if (op_rows < 1 || n < 1) return 0;
    return round_up(op_rows, BM) * round_up(n, BK);
}

extern "C" int64_t rti_rbf_shared_packed_elems(int n, int64_t n_pixels) {
    if (n < 1 || n_pixels < 1) return 0;
    return (int64_t)n * round_up(n_pixels, BN);
}

extern "C" int rti_rbf_shared_operator(const float *d_lx, const float *d_ly, int n, const double *d_axis, int g,
                                       int kind, uint16_t *d_op_hi, uint16_t *d_op_lo, float *d_inv_scale,
                                       int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_lx && d_ly && d_op_hi && d_op_lo && d_inv_scale, "rti_rbf_shared_operator: NULL pointer");
    RTI_REQUIRE(n >= 2, "rti_rbf_shared_operator: need >= 2 nodes, got %d", n);
    RTI_REQUIRE(kind == RTI_RBF_OP_GRID || kind == RTI_RBF_OP_WEIGHTS, "rti_rbf_shared_operator: unknown kind %d", kind);
    RTI_REQUIRE(kind == RTI_RBF_OP_WEIGHTS || (d_axis && g >= 1), "rti_rbf_shared_operator: grid axis missing");
    cudaStream_t s = as_stream(stream);
    const int rows = kind == RTI_RBF_OP_GRID ? g * g : n;
    const long long n_pad = round_up(n, BK), nn = (long long)n * n;
    const int batch = n < 148 * 2 ? n : 148 * 2;
    const int64_t fit_bytes = rti_rbf_workspace_bytes(n, batch);
    RTI_REQUIRE(fit_bytes > 0, "rti_rbf_shared_operator: n = %d is not supported by the RBF factorisation", n);

    // one allocation, carved: lx_rep, ly_rep, ident (4 B each), A^-1 (8 B), scale (8 B per row),
    // fit workspace, and for the grid operator the fp64 Phi.A^-1 columns [n][g*g]
    const long long grid_elems = kind == RTI_RBF_OP_GRID ? (long long)n * rows : 0;
    const size_t off_ly = round_up(nn * 4, 256), off_id = 2 * off_ly, off_w = 3 * off_ly;
    const size_t off_scale = off_w + round_up(nn * 8, 256);
    const size_t off_fit = off_scale + round_up((long long)rows * 8, 256);
    const size_t off_grid = off_fit + round_up(fit_bytes, 256);
    const size_t total = off_grid + grid_elems * 8 + 256;
    uint8_t *tmp = nullptr;
    RTI_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&tmp), total, s));
    float *lx_rep = reinterpret_cast<float *>(tmp), *ly_rep = reinterpret_cast<float *>(tmp + off_ly);
    int32_t *ident = reinterpret_cast<int32_t *>(tmp + off_id);
    double *w = reinterpret_cast<double *>(tmp + off_w), *scale = reinterpret_cast<double *>(tmp + off_scale);
    double *grids = reinterpret_cast<double *>(tmp + off_grid);

    int rc = RTI_OK;
    do {
        replicate_nodes_kernel<<<dim3((n + 255) / 256, n), 256, 0, s>>>(d_lx, d_ly, n, lx_rep, ly_rep, ident);
        count_launch();
        if ((rc = check_cuda(cudaGetLastError(), "replicate_nodes_kernel"))) break;
        if ((rc = rti_rbf_fit_soa(lx_rep, ly_rep, ident, n, n, w, tmp + off_fit, fit_bytes, batch, d_bad, stream))) break;
        const double *src = w;
        long long src_stride = n;
        if (kind == RTI_RBF_OP_GRID) {
            if ((rc = rti_rbf_grids(lx_rep, ly_rep, w, n, n, d_axis, g, grids, stream))) break;
            src = grids;
            src_stride = rows;
        }
        const long long op_elems = rti_rbf_shared_operator_elems(rows, n);
        if ((rc = check_cuda(cudaMemsetAsync(d_op_hi, 0, op_elems * 2, s), "memset op_hi"))) break;
        if ((rc = check_cuda(cudaMemsetAsync(d_op_lo, 0, op_elems * 2, s), "memset op_lo"))) break;
        operator_scale_kernel<<<(rows + 127) / 128, 128, 0, s>>>(src, src_stride, n, rows, scale, d_inv_scale, d_bad);
        count_launch();
        operator_split_kernel<<<dim3((rows + 31) / 32, (n + 31) / 32), dim3(32, 8), 0, s>>>(src, src_stride, n, rows,
                                                                                          scale, n_pad, d_op_hi, d_op_lo);
        count_launch();
        rc = check_cuda(cudaGetLastError(), "operator_split_kernel");
    } while (0);
    cudaFreeAsync(tmp, s);
    return rc;
}

extern "C" int rti_rbf_shared_pack_frames(const uint8_t *d_frames, int64_t frame_stride, int n, int64_t n_pixels,
                                          uint16_t *d_packed, rti_stream_t stream) {
    RTI_REQUIRE(d_frames && d_packed, "rti_rbf_shared_pack_frames: NULL pointer");
    RTI_REQUIRE(n >= 1 && n <= 65535 && n_pixels >= 1, "rti_rbf_shared_pack_frames: bad shape n=%d pixels=%lld", n,
                (long long)n_pixels);
    RTI_REQUIRE((reinterpret_cast<uintptr_t>(d_packed) & 15) == 0, "rti_rbf_shared_pack_frames: d_packed not 16-byte aligned");
    const long long p_pad = round_up(n_pixels, BN);
    const long long threads = p_pad / 8;
    pack_frames_kernel<<<dim3((unsigned)((threads + 255) / 256), n), 256, 0, as_stream(stream)>>>(
        d_frames, frame_stride, n, n_pixels, p_pad, d_packed);
    RTI_LAUNCHED("pack_frames_kernel");
    return RTI_OK;
}

extern "C" int rti_rbf_shared_pack_soa(const int32_t *d_inten, int64_t n_pixels, int n, uint16_t *d_packed,
                                       int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_inten && d_packed, "rti_rbf_shared_pack_soa: NULL pointer");
    RTI_REQUIRE(n >= 1 && n <= 65535 * 32 && n_pixels >= 1, "rti_rbf_shared_pack_soa: bad shape");
    const long long p_pad = round_up(n_pixels, BN);
    pack_soa_kernel<<<dim3((unsigned)(p_pad / 32), (n + 31) / 32), dim3(32, 8), 0, as_stream(stream)>>>(
        d_inten, n_pixels, n, p_pad, d_packed, d_bad);
    RTI_LAUNCHED("pack_soa_kernel");
    return RTI_OK;
}

extern "C" int rti_rbf_shared_table(const uint16_t *d_packed, int n, int64_t n_pixels, const uint16_t *d_op_hi,
                                    const uint16_t *d_op_lo, const float *d_inv_scale, int op_rows, int32_t *d_table,
                                    int64_t image_stride, int32_t *d_bad, rti_stream_t stream) {
    RTI_REQUIRE(d_packed && d_op_hi && d_op_lo && d_inv_scale && d_table, "rti_rbf_shared_table: NULL pointer");
    RTI_REQUIRE(n >= 1 && n_pixels >= 1 && op_rows >= 1 && image_stride >= n_pixels, "rti_rbf_shared_table: bad shape");
    return gemm_launch(d_packed, n, n_pixels, d_op_hi, d_op_lo, d_inv_scale, op_rows, d_table, image_stride, nullptr,
                       d_bad, as_stream(stream));
}

extern "C" int rti_rbf_shared_apply_f64(const uint16_t *d_packed, int n, int64_t n_pixels, const uint16_t *d_op_hi,
                                        const uint16_t *d_op_lo, const float *d_inv_scale, int op_rows, double *d_out,
                                        rti_stream_t stream) {
    RTI_REQUIRE(d_packed && d_op_hi && d_op_lo && d_inv_scale && d_out, "rti_rbf_shared_apply_f64: NULL pointer");
    RTI_REQUIRE(n >= 1 && n_pixels >= 1 && op_rows >= 1, "rti_rbf_shared_apply_f64: bad shape");
    return gemm_launch(d_packed, n, n_pixels, d_op_hi, d_op_lo, d_inv_scale, op_rows, nullptr, 0, d_out, nullptr,
                       as_stream(stream));
}
