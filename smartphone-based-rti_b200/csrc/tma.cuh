// TMA (cp.async.bulk.tensor) and mbarrier PTX wrappers plus the host-side tensor-map encoder,
// shared by the kernels that stage tiles through shared memory with the TMA unit.
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace rti {

// ---- PTX wrappers ---------------------------------------------------------------------
static __device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

static __device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
static __device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
static __device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Spin on a phase parity.  Release builds wait for as long as it takes (a debugger, MPS time
// slicing or a long co-scheduled kernel must give a slow run, not a sticky context error).
// -DRTI_MBAR_TIMEOUT (the `debug` target of csrc/Makefile) turns a wait that outlives ~2 s of SM
// clocks into a trap, so that a protocol bug under development fails the launch instead of
// hanging the device.
static __device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
#ifdef RTI_MBAR_TIMEOUT
    const long long t0 = clock64();
#endif
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
#ifdef RTI_MBAR_TIMEOUT
        if (clock64() - t0 > 4000000000LL) __trap();
#endif
    }
}

static __device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
static __device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
static __device__ __forceinline__ void tma_store_2d(const CUtensorMap *map, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1)
                 : "memory");
}
static __device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
static __device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
static __device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
static __device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
static __device__ __forceinline__ void prefetch_tmap(const CUtensorMap *map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// ---- host side -----------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

// 2-D row-major tensor [dim1][dim0] with a row pitch in bytes, 128-byte swizzled boxes
inline int make_map_2d(CUtensorMap *map, CUtensorMapDataType dtype, const void *ptr, uint64_t dim0, uint64_t dim1,
                uint64_t pitch_bytes, uint32_t box0, uint32_t box1, const char *what) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (fn == nullptr) {
        set_error("cuTensorMapEncodeTiled is not available from this driver");
        return RTI_ERR_CUDA;
    }
    const cuuint64_t dims[2] = {dim0, dim1};
    const cuuint64_t strides[1] = {pitch_bytes};
    const cuuint32_t box[2] = {box0, box1};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult rc = fn(map, dtype, 2, const_cast<void *>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(%s) failed with CUresult %d (dims %llu x %llu, pitch %llu)", what, (int)rc,
                  (unsigned long long)dim0, (unsigned long long)dim1, (unsigned long long)pitch_bytes);
        return RTI_ERR_CUDA;
    }
    return RTI_OK;
}

// packed intensities as a 3-D tensor {256 pixels, n lights, pixel blocks}; out-of-range lights read as zero
}  // namespace rti
