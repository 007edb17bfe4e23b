// Linear RBF path (placeholder until the batched Cholesky kernels land).
#include "common.cuh"
using namespace rti;

extern "C" int64_t rti_rbf_workspace_bytes(int n, int batch) { (void)n; (void)batch; return 0; }
extern "C" int rti_rbf_fit_soa(const float *, const float *, const int32_t *, int64_t, int, double *, void *,
                               int64_t, int, int32_t *, rti_stream_t) {
    set_error("rti_rbf_fit_soa: not built yet");
    return RTI_ERR_ARG;
}
extern "C" int rti_rbf_table(const float *, const float *, const double *, int64_t, int, const double *, int,
                             int32_t *, int64_t, int32_t *, rti_stream_t) {
    set_error("rti_rbf_table: not built yet");
    return RTI_ERR_ARG;
}
extern "C" int rti_rbf_grids(const float *, const float *, const double *, int64_t, int, const double *, int,
                             double *, rti_stream_t) {
    set_error("rti_rbf_grids: not built yet");
    return RTI_ERR_ARG;
}
