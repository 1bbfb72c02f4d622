// bf16 instantiations of the direct GEMV (separate TU so the two dtypes compile in parallel)
#define PQ_GV_DTYPE_BF16 1
#include "pq_gemv_direct.cu"
