// SiLU(gate) * up as its own launch: the fallback of pq_b200_gemm_silu_mul
// for the kernels that do not take the activation on their x load, and the drop-in for the
// reference's silu_and_mul (csrc/activation_kernels.cu via layers/activation.py:65-84).
// HBM-bound elementwise work: 16-byte loads of gate and up, one 16-byte store, grid-stride.
#include "pq_common.cuh"

namespace pq {

extern bool g_use_pdl;

template <typename T>
__global__ void __launch_bounds__(256) silu_and_mul_kernel(const T* __restrict__ x, int64_t x_stride,
                                                           T* __restrict__ out, int64_t out_stride,
                                                           int t, int k, int vec) {
  pdl_launch_dependents();
  pdl_wait();
  const int per_row = vec ? k / 8 : k / 2;             // work items per token: 8 or 2 elements
  const int64_t total = (int64_t)t * per_row;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int row = (int)(i / per_row), c = (int)(i - (int64_t)row * per_row);
    const T* g = x + row * x_stride;
    T* o = out + row * out_stride;
    if (vec) {
      const uint4 gv = ld_stream_16(g + c * 8), uv = ld_stream_16(g + k + c * 8);
      uint4 r;
      r.x = silu_mul_pair<T>(gv.x, uv.x); r.y = silu_mul_pair<T>(gv.y, uv.y);
      r.z = silu_mul_pair<T>(gv.z, uv.z); r.w = silu_mul_pair<T>(gv.w, uv.w);
      *reinterpret_cast<uint4*>(o + c * 8) = r;
    } else {
      const uint32_t gv = *reinterpret_cast<const uint32_t*>(g + c * 2);
      const uint32_t uv = *reinterpret_cast<const uint32_t*>(g + k + c * 2);
      *reinterpret_cast<uint32_t*>(o + c * 2) = silu_mul_pair<T>(gv, uv);
    }
  }
}

int silu_and_mul_dispatch(const void* x, int64_t x_stride, void* out, int64_t out_stride, int64_t t,
                          int64_t k, int dtype, cudaStream_t stream) {
  if (k % 2 || (x_stride * 2) % 4 || (out_stride * 2) % 4 || ((uintptr_t)x | (uintptr_t)out) % 4) return PQ_ERR_ALIGN;
  const int vec = (k % 8 == 0 && x_stride % 8 == 0 && out_stride % 8 == 0 &&
                   (((uintptr_t)x | (uintptr_t)out) & 15) == 0) ? 1 : 0;
  const int64_t total = t * (vec ? k / 8 : k / 2);
  int blocks = (int)((total + 255) / 256);
  const int cap = sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(256); cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 1 : 0;
  cudaError_t e;
  if (dtype == PQ_F16)
    e = cudaLaunchKernelEx(&cfg, silu_and_mul_kernel<__half>, (const __half*)x, x_stride, (__half*)out, out_stride,
                           (int)t, (int)k, vec);
  else
    e = cudaLaunchKernelEx(&cfg, silu_and_mul_kernel<__nv_bfloat16>, (const __nv_bfloat16*)x, x_stride,
                           (__nv_bfloat16*)out, out_stride, (int)t, (int)k, vec);
  return e == cudaSuccess ? PQ_OK : PQ_ERR_DEVICE;
}

}  // namespace pq
