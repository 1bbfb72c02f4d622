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

// ---- (residual add +) RMSNorm as its own launch: the unfused form of what the table GEMV does in its
// epilogue / on its x load, the fallback for the kernels without that fusion, and
// the drop-in for fused_add_rms_norm / rms_norm (csrc/layernorm_kernels.cu:17-98).  Same roundings:
// z = T(x + r); variance over float(z); out = T(T(z * rstd) * w).  One CTA per token; the sum of
// squares is reduced in a fixed order (lane tree, then warps in order).
constexpr int kNormThreads = 512;

template <typename T>
__global__ void __launch_bounds__(kNormThreads) add_rms_norm_kernel(T* __restrict__ input, int64_t in_stride,
                                                                   T* __restrict__ residual, int64_t res_stride,
                                                                   const T* __restrict__ weight, float eps, int hidden) {
  __shared__ float warp_part[kNormThreads / 32];
  __shared__ float s_rstd;
  pdl_launch_dependents();
  pdl_wait();
  T* x = input + blockIdx.x * in_stride;
  T* r = residual ? residual + blockIdx.x * res_stride : nullptr;
  float ss = 0.f;
  for (int i = threadIdx.x; i < hidden; i += kNormThreads) {
    T z = x[i];
    if (r) { z = Traits<T>::from_f(Traits<T>::to_f(z) + Traits<T>::to_f(r[i])); r[i] = z; }
    const float zf = Traits<T>::to_f(z);
    ss = fmaf(zf, zf, ss);
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int i = 0; i < kNormThreads / 32; ++i) tot += warp_part[i];
    s_rstd = rsqrtf(tot / (float)hidden + eps);
  }
  __syncthreads();
  const float rstd = s_rstd;
  const T* src = r ? r : x;
  for (int i = threadIdx.x; i < hidden; i += kNormThreads) {
    const T scaled = Traits<T>::from_f(Traits<T>::to_f(src[i]) * rstd);
    x[i] = Traits<T>::from_f(Traits<T>::to_f(scaled) * Traits<T>::to_f(weight[i]));   // T * T, one rounding
  }
}

int add_rms_norm_dispatch(void* input, int64_t in_stride, void* residual, int64_t res_stride, const void* weight,
                          float eps, int64_t t, int64_t hidden, int dtype, cudaStream_t stream) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)t); cfg.blockDim = dim3(kNormThreads); cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 1 : 0;
  cudaError_t e;
  if (dtype == PQ_F16)
    e = cudaLaunchKernelEx(&cfg, add_rms_norm_kernel<__half>, (__half*)input, in_stride, (__half*)residual, res_stride,
                           (const __half*)weight, eps, (int)hidden);
  else
    e = cudaLaunchKernelEx(&cfg, add_rms_norm_kernel<__nv_bfloat16>, (__nv_bfloat16*)input, in_stride,
                           (__nv_bfloat16*)residual, res_stride, (const __nv_bfloat16*)weight, eps, (int)hidden);
  return (int)e;
}

}  // namespace pq
