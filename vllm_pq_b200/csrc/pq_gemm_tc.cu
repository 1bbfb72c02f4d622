// placeholder - filled in below in this round
#include "pq_common.cuh"
namespace pq {
int gemm_tc_dispatch(const void*, int64_t, const void*, const void*, const void*, const void*,
                     void*, int64_t, int64_t, int64_t, int64_t, int, int, int, const int64_t*,
                     int, int, void*, int64_t, cudaStream_t) { return PQ_ERR_UNSUPPORTED; }
int64_t gemm_tc_workspace(int64_t, int64_t, int64_t, int, int, int, int) { return 0; }
bool gemm_tc_supported(int64_t, int64_t, int64_t, int, int, int) { return false; }
}
