// Shared device/host helpers for the PQ kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pq_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libpq_b200 is written for sm_100a only"
#endif

namespace pq {

constexpr int kWarp = 32;

// Output partitions of a fused layer (QKV: 3, gate_up: 2).  Row-blocks handed
// to CTAs never straddle a partition, so a CTA needs one partition's codebooks.
// Reference: the cumulative-size walk of gemm_kernels.cu:43-51,499-514 (<= 4
// partitions there; 8 here).
struct Parts {
  int n_parts;
  int row_end[PQ_MAX_PARTS];   // cumulative row ends
  int cta_end[PQ_MAX_PARTS];   // cumulative CTA ends (rows_per_cta granularity)
};

struct RowBlock {
  int part;
  int row0;
  int row1;
};

__device__ __forceinline__ RowBlock locate_row_block(const Parts& ps, int block, int rows_per_cta) {
  int p = 0;
#pragma unroll
  for (int i = 0; i < PQ_MAX_PARTS - 1; ++i)
    if (i < ps.n_parts - 1 && block >= ps.cta_end[i]) p = i + 1;
  const int cta0 = p ? ps.cta_end[p - 1] : 0;
  const int rs = p ? ps.row_end[p - 1] : 0;
  RowBlock rb;
  rb.part = p;
  rb.row0 = rs + (block - cta0) * rows_per_cta;
  rb.row1 = min(rb.row0 + rows_per_cta, ps.row_end[p]);
  return rb;
}

inline int make_parts(Parts& ps, const int64_t* part_sizes, int n_parts, int64_t n, int rows_per_cta) {
  if (n_parts < 1 || n_parts > PQ_MAX_PARTS) return PQ_ERR_SHAPE;
  int64_t row = 0, cta = 0;
  ps.n_parts = n_parts;
  for (int i = 0; i < PQ_MAX_PARTS; ++i) {
    if (i < n_parts) {
      if (part_sizes[i] <= 0) return PQ_ERR_SHAPE;
      row += part_sizes[i];
      cta += (part_sizes[i] + rows_per_cta - 1) / rows_per_cta;
    }
    ps.row_end[i] = (int)row;
    ps.cta_end[i] = (int)cta;
  }
  return row == n ? PQ_OK : PQ_ERR_SHAPE;
}

// ---- dtype traits -----------------------------------------------------------
template <typename T> struct Traits;
template <> struct Traits<__half> {
  using T2 = __half2;
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
  static __device__ __forceinline__ float2 to_f2(uint32_t v) {
    return __half22float2(*reinterpret_cast<const __half2*>(&v));
  }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};
template <> struct Traits<__nv_bfloat16> {
  using T2 = __nv_bfloat162;
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  static __device__ __forceinline__ float2 to_f2(uint32_t v) {
    // bf16 -> fp32 is a 16-bit shift: two ALU ops, no conversion pipe
    return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
  }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

// ---- memory helpers ----------------------------------------------------------
__device__ __forceinline__ uint4 ld_stream_16(const void* p) {   // read-once data: keep out of L1
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_cg_16(const void* p) {       // L2-resident gathers
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_cg_8(const void* p) {
  uint2 r;
  asm volatile("ld.global.cg.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// SiLU(gate) * up on a pair of packed elements, rounded the way the reference's silu_and_mul is
// (csrc/activation_kernels.cu: (T)(x / (1 + expf(-x))) in fp32, then the product in T).
template <typename T>
__device__ __forceinline__ uint32_t silu_mul_pair(uint32_t gate, uint32_t up) {
  using T2 = typename Traits<T>::T2;
  const float2 g = Traits<T>::to_f2(gate);
  uint32_t s = Traits<T>::pack(g.x / (1.0f + expf(-g.x)), g.y / (1.0f + expf(-g.y)));
  const T2 r = __hmul2(*reinterpret_cast<const T2*>(&s), *reinterpret_cast<const T2*>(&up));
  return *reinterpret_cast<const uint32_t*>(&r);
}

// RMSNorm of a pair of packed elements, rounded the way the reference's fused_add_rms_norm is
// (csrc/layernorm_kernels.cu:91-97: temp *= s_variance rounds to T, temp *= weight rounds to T).
template <typename T>
__device__ __forceinline__ uint32_t rms_norm_pair(uint32_t z, float rstd, uint32_t w) {
  using T2 = typename Traits<T>::T2;
  const float2 f = Traits<T>::to_f2(z);
  uint32_t s = Traits<T>::pack(f.x * rstd, f.y * rstd);
  const T2 r = __hmul2(*reinterpret_cast<const T2*>(&s), *reinterpret_cast<const T2*>(&w));
  return *reinterpret_cast<const uint32_t*>(&r);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Programmatic dependent launch: everything before pdl_wait() may overlap the
// tail of the previous kernel in the stream (weights do not depend on it).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

inline int sm_count() {
  static int cached[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 16 && cached[dev]) return cached[dev];
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (dev < 16) cached[dev] = sms;
  return sms;
}

}  // namespace pq
