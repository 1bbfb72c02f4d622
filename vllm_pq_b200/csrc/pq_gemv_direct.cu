// Decode-path PQ GEMV, "direct" form: gather codewords, FMA against x.
//
// Replaces Code1x16MatVec / Code2x8MatVec and their per-token host loop
// (reference csrc/quantization/aqlm/gemm_kernels.cu:31-183, 424-496) for every
// (nbits, d, m); it is the decode kernel for nbits > 8 (a 2^nbits-entry lookup
// table per group would be larger than the weights) and the small-problem
// fallback for nbits <= 8, where pq_gemv_lut.cu is the bandwidth path.
//
// Differences from the reference kernels, by design:
//   * all T <= 4 tokens of a tile in ONE launch, codes read once per tile
//     (reference: one launch per token, gemm_kernels.cu:437-442)
//   * scale and bias fused, applied in fp32 before the only rounding of y
//     (reference: two extra torch elementwise launches, :443-447)
//   * four rows in flight per warp -> 64 B of codes outstanding per lane
//     (reference: one 16 B load per lane per row)
//   * x staged once per CTA for the whole K, padded so the 128-bit reads of
//     lanes that own consecutive code vectors are bank-conflict free
//   * 256-entry codebooks replicated to one 128-byte line per entry so every
//     lane of a quarter-warp reads its own 16-byte slot (conflict free);
//     4096-entry codebooks unreplicated in smem; 65536-entry ones gathered
//     from L2 with ld.global.cg
//   * no K % 64 or <= 4-partition assumption; rows
//     whose code bytes are not a multiple of 16 take the scalar kernel
//   * programmatic dependent launch: codebook staging overlaps the tail of
//     the previous kernel; x is only touched after griddepcontrol.wait
#include <type_traits>

#include "pq_common.cuh"

namespace pq {

constexpr int kGvWarps = 4;
constexpr int kGvThreads = kGvWarps * kWarp;
constexpr int kGvRB = 4;                 // rows in flight per warp

enum CbMode { CB_SMEM_REPL = 0, CB_SMEM_FLAT = 1, CB_GLOBAL = 2 };

template <int D> struct GV {
  static constexpr int kEntryBytes = D * 2;
  static constexpr int kUnit = kEntryBytes < 16 ? kEntryBytes : 16;   // staging / pad unit
  static constexpr int kRepl = 128 / kEntryBytes;                    // copies per entry line
};

template <int D> struct CW { uint32_t w[D / 2]; };

template <int D>
__device__ __forceinline__ CW<D> ld_cw_smem(const unsigned char* p) {
  CW<D> v;
  if constexpr (D >= 8) {
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      uint4 q = reinterpret_cast<const uint4*>(p)[i];
      v.w[4 * i] = q.x; v.w[4 * i + 1] = q.y; v.w[4 * i + 2] = q.z; v.w[4 * i + 3] = q.w;
    }
  } else {
    uint2 q = *reinterpret_cast<const uint2*>(p);
    v.w[0] = q.x; v.w[1] = q.y;
  }
  return v;
}
template <int D>
__device__ __forceinline__ CW<D> ld_cw_gmem(const unsigned char* p) {
  CW<D> v;
  if constexpr (D >= 8) {
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      uint4 q = ld_cg_16(reinterpret_cast<const uint4*>(p) + i);
      v.w[4 * i] = q.x; v.w[4 * i + 1] = q.y; v.w[4 * i + 2] = q.z; v.w[4 * i + 3] = q.w;
    }
  } else {
    uint2 q = ld_cg_8(p);
    v.w[0] = q.x; v.w[1] = q.y;
  }
  return v;
}

// dot(codeword sum, x group) -> fp32.
//  fp16: codewords summed with HADD2 (M=2: identical to the dequantised weight), then a
//        D/2-long HFMA2 chain from zero and an fp32 add of the two halves - the numerics of
//        the reference GEMV (gemm_kernels.cu:84-87,166-170).
//  bf16: everything in fp32 (bf16 FMA chains would lose the tolerance gate).
template <typename T, int D, int M>
__device__ __forceinline__ float dot_group(const CW<D> (&cw)[M], const CW<D>& xv) {
  if constexpr (sizeof(T) == 2 && std::is_same<T, __half>::value) {
    __half2 acc = __float2half2_rn(0.f);
#pragma unroll
    for (int j = 0; j < D / 2; ++j) {
      __half2 a = *reinterpret_cast<const __half2*>(&cw[0].w[j]);
#pragma unroll
      for (int m = 1; m < M; ++m) a = __hadd2(a, *reinterpret_cast<const __half2*>(&cw[m].w[j]));
      acc = __hfma2(a, *reinterpret_cast<const __half2*>(&xv.w[j]), acc);
    }
    return __low2float(acc) + __high2float(acc);
  } else {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < D / 2; ++j) {
      float2 a = Traits<T>::to_f2(cw[0].w[j]);
#pragma unroll
      for (int m = 1; m < M; ++m) {
        float2 b = Traits<T>::to_f2(cw[m].w[j]);
        a.x += b.x; a.y += b.y;
      }
      float2 x = Traits<T>::to_f2(xv.w[j]);
      s = fmaf(a.x, x.x, s);
      s = fmaf(a.y, x.y, s);
    }
    return s;
  }
}

template <typename CodeT>
__device__ __forceinline__ uint32_t code_at(const uint4& v, int idx) {   // idx is a constant after unrolling
  constexpr int per_word = 4 / (int)sizeof(CodeT);
  const int word = idx / per_word, sub = idx % per_word;
  const uint32_t w = word == 0 ? v.x : word == 1 ? v.y : word == 2 ? v.z : v.w;
  if constexpr (sizeof(CodeT) == 1) return (w >> (8 * sub)) & 0xffu;
  else return (w >> (16 * sub)) & 0xffffu;
}

struct GvArgs {
  const void* x; const void* codes; const void* codebooks; const void* scales; const void* bias;
  void* y;
  long long x_stride, y_stride;
  int T, N, K, G, E;
  int rows_per_cta;
  int xs_stride;       // bytes per token row of the padded x tile
  Parts parts;
};

// Fast path: row code bytes % 16 == 0; one uint4 per lane = GPL groups x M codes.
template <typename T, int D, typename CodeT, int M, int TT, int CBM>
__global__ void __launch_bounds__(kGvThreads)
pq_gemv_direct_kernel(const GvArgs a) {
  constexpr int GPL = 16 / ((int)sizeof(CodeT) * M);
  constexpr int EB = GV<D>::kEntryBytes;
  constexpr int UNIT = GV<D>::kUnit;
  constexpr int REPL = GV<D>::kRepl;
  extern __shared__ __align__(16) unsigned char gv_smem[];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int t0 = blockIdx.y * TT;
  const RowBlock rb = locate_row_block(a.parts, blockIdx.x, a.rows_per_cta);
  const unsigned char* cb_g = reinterpret_cast<const unsigned char*>(a.codebooks) +
                              (size_t)rb.part * M * a.E * EB;

  // ---- stage this partition's codebooks (independent of the previous kernel) ----
  unsigned char* xs = gv_smem;
  const unsigned char* cb = cb_g;
  if constexpr (CBM == CB_SMEM_REPL) {
    unsigned char* cbs = gv_smem;
    xs = gv_smem + (size_t)M * a.E * 128;
    const int units = M * a.E * (EB / UNIT);
    for (int u = threadIdx.x; u < units; u += kGvThreads) {
      const int ent = u / (EB / UNIT), sub = u % (EB / UNIT);
      if constexpr (UNIT == 16) {
        const uint4 v = reinterpret_cast<const uint4*>(cb_g)[u];
#pragma unroll
        for (int r = 0; r < REPL; ++r)
          *reinterpret_cast<uint4*>(cbs + (size_t)ent * 128 + r * EB + sub * 16) = v;
      } else {
        const uint2 v = reinterpret_cast<const uint2*>(cb_g)[u];
#pragma unroll
        for (int r = 0; r < REPL; ++r)
          *reinterpret_cast<uint2*>(cbs + (size_t)ent * 128 + r * EB) = v;
      }
    }
    cb = cbs + (lane & (REPL - 1)) * EB;
  } else if constexpr (CBM == CB_SMEM_FLAT) {
    unsigned char* cbs = gv_smem;
    const int n16 = M * a.E * EB / 16;
    xs = gv_smem + (size_t)n16 * 16;
    for (int u = threadIdx.x; u < n16; u += kGvThreads)
      reinterpret_cast<uint4*>(cbs)[u] = reinterpret_cast<const uint4*>(cb_g)[u];
    cb = cbs;
  }
  constexpr int CB_SHIFT_BYTES = (CBM == CB_SMEM_REPL) ? 128 : EB;   // bytes per entry slot

  // ---- stage x (depends on the previous kernel) ----
  pdl_wait();
  pdl_launch_dependents();
  {
    const int units_per_row = a.K * 2 / UNIT;
    const int upg = EB / UNIT;                      // units per group
    for (int tt = 0; tt < TT; ++tt) {
      const int t = t0 + tt;
      const unsigned char* xrow = reinterpret_cast<const unsigned char*>(a.x) + (size_t)t * a.x_stride * 2;
      for (int u = threadIdx.x; u < units_per_row; u += kGvThreads) {
        const int g = u / upg;
        unsigned char* dst = xs + (size_t)tt * a.xs_stride + (size_t)u * UNIT + (size_t)(g / GPL) * UNIT;
        if constexpr (UNIT == 16) {
          uint4 v = make_uint4(0, 0, 0, 0);
          if (t < a.T) v = reinterpret_cast<const uint4*>(xrow)[u];
          *reinterpret_cast<uint4*>(dst) = v;
        } else {
          uint2 v = make_uint2(0, 0);
          if (t < a.T) v = reinterpret_cast<const uint2*>(xrow)[u];
          *reinterpret_cast<uint2*>(dst) = v;
        }
      }
    }
  }
  __syncthreads();

  const int vpr = a.G / GPL;                         // 16-byte code vectors per row
  const uint4* codes = reinterpret_cast<const uint4*>(a.codes);
  constexpr int LANE_X_BYTES = GPL * EB + UNIT;      // padded stride between lanes' x runs

  for (int rbase = rb.row0 + warp * kGvRB; rbase < rb.row1; rbase += kGvWarps * kGvRB) {
    float res[kGvRB][TT];
#pragma unroll
    for (int r = 0; r < kGvRB; ++r)
#pragma unroll
      for (int t = 0; t < TT; ++t) res[r][t] = 0.f;

    for (int v = lane; v < vpr; v += 32) {
      uint4 cv[kGvRB];
#pragma unroll
      for (int r = 0; r < kGvRB; ++r) {
        const int row = rbase + r;
        cv[r] = row < rb.row1 ? ld_stream_16(codes + (size_t)row * vpr + v) : make_uint4(0, 0, 0, 0);
      }
      const unsigned char* xl = xs + (size_t)v * LANE_X_BYTES;
#pragma unroll
      for (int i = 0; i < GPL; ++i) {
        CW<D> xv[TT];
#pragma unroll
        for (int t = 0; t < TT; ++t) xv[t] = ld_cw_smem<D>(xl + (size_t)t * a.xs_stride + i * EB);
#pragma unroll
        for (int r = 0; r < kGvRB; ++r) {
          CW<D> cw[M];
#pragma unroll
          for (int m = 0; m < M; ++m) {
            uint32_t code = code_at<CodeT>(cv[r], i * M + m);
            code &= (uint32_t)a.E - 1u;
            const unsigned char* ent = cb + ((size_t)m * a.E + code) * CB_SHIFT_BYTES;
            cw[m] = (CBM == CB_GLOBAL) ? ld_cw_gmem<D>(ent) : ld_cw_smem<D>(ent);
          }
#pragma unroll
          for (int t = 0; t < TT; ++t) res[r][t] += dot_group<T, D, M>(cw, xv[t]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kGvRB; ++r)
#pragma unroll
      for (int t = 0; t < TT; ++t) res[r][t] = warp_sum(res[r][t]);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < kGvRB; ++r) {
        const int row = rbase + r;
        if (row >= rb.row1) continue;
        const float s = Traits<T>::to_f(reinterpret_cast<const T*>(a.scales)[row]);
        const float b = a.bias ? Traits<T>::to_f(reinterpret_cast<const T*>(a.bias)[row]) : 0.f;
#pragma unroll
        for (int t = 0; t < TT; ++t)
          if (t0 + t < a.T)
            reinterpret_cast<T*>(a.y)[(size_t)(t0 + t) * a.y_stride + row] =
                Traits<T>::from_f(fmaf(res[r][t], s, b));
      }
    }
  }
}

// Scalar fallback: any d, any m, any alignment.  One warp per row, lanes stride
// over groups, operands from global (L1/L2-cached), fp32 maths.
template <typename T, typename CodeT, int TT>
__global__ void __launch_bounds__(kGvThreads)
pq_gemv_scalar_kernel(const GvArgs a, int D, int M) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int t0 = blockIdx.y * TT;
  const RowBlock rb = locate_row_block(a.parts, blockIdx.x, a.rows_per_cta);
  const T* cb = reinterpret_cast<const T*>(a.codebooks) + (size_t)rb.part * M * a.E * D;
  const CodeT* codes = reinterpret_cast<const CodeT*>(a.codes);
  const T* x = reinterpret_cast<const T*>(a.x);
  pdl_wait();
  pdl_launch_dependents();
  for (int row = rb.row0 + warp; row < rb.row1; row += kGvWarps) {
    float res[TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) res[t] = 0.f;
    for (int g = lane; g < a.G; g += 32) {
      for (int j = 0; j < D; ++j) {
        float w = 0.f;
        for (int m = 0; m < M; ++m) {
          const uint32_t code = (uint32_t)codes[((size_t)row * a.G + g) * M + m] & ((uint32_t)a.E - 1u);
          w += Traits<T>::to_f(cb[((size_t)m * a.E + code) * D + j]);
        }
#pragma unroll
        for (int t = 0; t < TT; ++t)
          if (t0 + t < a.T)
            res[t] = fmaf(w, Traits<T>::to_f(x[(size_t)(t0 + t) * a.x_stride + (size_t)g * D + j]), res[t]);
      }
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) res[t] = warp_sum(res[t]);
    if (lane == 0) {
      const float s = Traits<T>::to_f(reinterpret_cast<const T*>(a.scales)[row]);
      const float b = a.bias ? Traits<T>::to_f(reinterpret_cast<const T*>(a.bias)[row]) : 0.f;
#pragma unroll
      for (int t = 0; t < TT; ++t)
        if (t0 + t < a.T)
          reinterpret_cast<T*>(a.y)[(size_t)(t0 + t) * a.y_stride + row] = Traits<T>::from_f(fmaf(res[t], s, b));
    }
  }
}

// ---- host side ---------------------------------------------------------------
extern bool g_use_pdl;

template <typename Kern, typename... Args>
static int launch_maybe_pdl(Kern kern, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 1 : 0;
  return (int)cudaLaunchKernelEx(&cfg, kern, args...);
}

template <typename T, int D, typename CodeT, int M, int TT>
static int launch_fast(GvArgs& a, int cbm, size_t smem, dim3 grid, cudaStream_t stream) {
  auto set = [&](auto kern) {
    // all three kernels share one function-pointer TYPE, so the flag is keyed by the codebook
    // mode explicitly (a static inside this generic lambda is one object for all of them)
    static bool attr_set[3][64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > 48 * 1024 && !attr_set[cbm][dev & 63]) {
      const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024);
      if (e != cudaSuccess) return (int)e;
      attr_set[cbm][dev & 63] = true;
    }
    return launch_maybe_pdl(kern, grid, dim3(kGvThreads), smem, stream, a);
  };
  if (cbm == CB_SMEM_REPL) return set(pq_gemv_direct_kernel<T, D, CodeT, M, TT, CB_SMEM_REPL>);
  if constexpr (sizeof(CodeT) == 2) {     // 8-bit codes always fit the replicated layout
    if (cbm == CB_SMEM_FLAT) return set(pq_gemv_direct_kernel<T, D, CodeT, M, TT, CB_SMEM_FLAT>);
    return set(pq_gemv_direct_kernel<T, D, CodeT, M, TT, CB_GLOBAL>);
  }
  return PQ_ERR_UNSUPPORTED;
}

template <typename T, int D, typename CodeT, int M>
static int launch_fast_tt(GvArgs& a, int tt, int cbm, size_t smem, dim3 grid, cudaStream_t stream) {
  switch (tt) {
    case 1: return launch_fast<T, D, CodeT, M, 1>(a, cbm, smem, grid, stream);
    case 2: return launch_fast<T, D, CodeT, M, 2>(a, cbm, smem, grid, stream);
    default: return launch_fast<T, D, CodeT, M, 4>(a, cbm, smem, grid, stream);
  }
}

template <typename T, typename CodeT>
static int gemv_direct_t(GvArgs& a, int d, int m, cudaStream_t stream,
                         const int64_t* part_sizes, int n_parts) {
  const int sms = sm_count();
  const size_t row_bytes = (size_t)a.G * m * sizeof(CodeT);
  const bool fast = (d == 4 || d == 8 || d == 16) && (row_bytes % 16 == 0) &&
                    ((sizeof(CodeT) == 1 && (m == 1 || m == 2)) || (sizeof(CodeT) == 2 && m == 1)) &&
                    ((uintptr_t)a.codes % 16 == 0) && (((size_t)a.x_stride * 2) % (d == 4 ? 8 : 16) == 0) &&
                    ((uintptr_t)a.x % 16 == 0);
  // token tile: as many tokens per pass as shared memory allows (<= 4)
  int tt = a.T >= 4 ? 4 : (a.T >= 2 ? 2 : 1);
  const int rows_gran = fast ? kGvWarps * kGvRB : kGvWarps;
  int rows_per_cta = (int)(((int64_t)a.N + (int64_t)sms * 8 - 1) / ((int64_t)sms * 8));
  rows_per_cta = (rows_per_cta + rows_gran - 1) / rows_gran * rows_gran;
  a.rows_per_cta = rows_per_cta;
  int rc = make_parts(a.parts, part_sizes, n_parts, a.N, rows_per_cta);
  if (rc) return rc;

  if (!fast) {
    dim3 grid(a.parts.cta_end[n_parts - 1], (a.T + tt - 1) / tt);
    if (tt == 1) return launch_maybe_pdl(pq_gemv_scalar_kernel<T, CodeT, 1>, grid, dim3(kGvThreads), 0, stream, a, d, m);
    if (tt == 2) return launch_maybe_pdl(pq_gemv_scalar_kernel<T, CodeT, 2>, grid, dim3(kGvThreads), 0, stream, a, d, m);
    return launch_maybe_pdl(pq_gemv_scalar_kernel<T, CodeT, 4>, grid, dim3(kGvThreads), 0, stream, a, d, m);
  }

  const int gpl = 16 / ((int)sizeof(CodeT) * m);
  const int unit = d == 4 ? 8 : 16;
  const int eb = d * 2;
  a.xs_stride = a.K * 2 + (a.G / gpl) * unit;
  a.xs_stride = (a.xs_stride + 15) / 16 * 16;
  int cbm;
  size_t cb_smem;
  if ((size_t)m * a.E * 128 <= 64 * 1024) { cbm = CB_SMEM_REPL; cb_smem = (size_t)m * a.E * 128; }
  else if ((size_t)m * a.E * eb <= 128 * 1024) { cbm = CB_SMEM_FLAT; cb_smem = (size_t)m * a.E * eb; }
  else { cbm = CB_GLOBAL; cb_smem = 0; }
  while (tt > 1 && cb_smem + (size_t)tt * a.xs_stride > 200 * 1024) tt >>= 1;
  const size_t smem = cb_smem + (size_t)tt * a.xs_stride;
  if (smem > 216 * 1024) return PQ_ERR_UNSUPPORTED;
  dim3 grid(a.parts.cta_end[n_parts - 1], (a.T + tt - 1) / tt);

#define PQ_DISPATCH_D(CT, MM)                                                           \
  switch (d) {                                                                          \
    case 4:  return launch_fast_tt<T, 4, CT, MM>(a, tt, cbm, smem, grid, stream);        \
    case 8:  return launch_fast_tt<T, 8, CT, MM>(a, tt, cbm, smem, grid, stream);        \
    default: return launch_fast_tt<T, 16, CT, MM>(a, tt, cbm, smem, grid, stream);       \
  }
  if constexpr (sizeof(CodeT) == 1) {
    if (m == 1) { PQ_DISPATCH_D(uint8_t, 1) }
    PQ_DISPATCH_D(uint8_t, 2)
  } else {
    PQ_DISPATCH_D(uint16_t, 1)
  }
#undef PQ_DISPATCH_D
}

#ifndef PQ_GV_DTYPE_BF16
bool g_use_pdl = false;
int gemv_direct_bf16(GvArgs& a, int d, int m, int nbits, cudaStream_t stream, const int64_t* part_sizes, int n_parts);

int gemv_direct_dispatch(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                         const void* scales, const void* bias, void* y, int64_t y_stride,
                         int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                         const int64_t* part_sizes, int n_parts, int dtype, cudaStream_t stream) {
  GvArgs a{};
  a.x = x; a.codes = codes; a.codebooks = codebooks; a.scales = scales; a.bias = bias; a.y = y;
  a.x_stride = x_stride; a.y_stride = y_stride;
  a.T = (int)t; a.N = (int)n; a.K = (int)k; a.G = (int)(k / d); a.E = 1 << nbits;
  if (dtype == PQ_F16)
    return nbits <= 8 ? gemv_direct_t<__half, uint8_t>(a, d, m, stream, part_sizes, n_parts)
                      : gemv_direct_t<__half, uint16_t>(a, d, m, stream, part_sizes, n_parts);
  return gemv_direct_bf16(a, d, m, nbits, stream, part_sizes, n_parts);
}
#else
int gemv_direct_bf16(GvArgs& a, int d, int m, int nbits, cudaStream_t stream, const int64_t* part_sizes, int n_parts) {
  return nbits <= 8 ? gemv_direct_t<__nv_bfloat16, uint8_t>(a, d, m, stream, part_sizes, n_parts)
                    : gemv_direct_t<__nv_bfloat16, uint16_t>(a, d, m, stream, part_sizes, n_parts);
}
#endif

}  // namespace pq
