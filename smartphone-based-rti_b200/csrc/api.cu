// Library plumbing of librti_b200: error strings, device info, launch counter.
#include <stdarg.h>
#include <string.h>
#include <atomic>
#include "common.cuh"

namespace rti {

static thread_local char g_error[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return RTI_OK;
    set_error("CUDA error in %s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
    return RTI_ERR_CUDA;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace rti

extern "C" int rti_version(void) { return 100; }   // 0.1.0

extern "C" const char *rti_last_error_string(void) { return rti::g_error; }

extern "C" int rti_device_info(int device, int *sm_count, int64_t *hbm_bytes) {
    cudaDeviceProp prop;
    RTI_CUDA(cudaGetDeviceProperties(&prop, device));
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (hbm_bytes) *hbm_bytes = (int64_t)prop.totalGlobalMem;
    return RTI_OK;
}

extern "C" int64_t rti_launch_count(int reset) {
    if (reset) return rti::g_launches.exchange(0);
    return rti::g_launches.load();
}
