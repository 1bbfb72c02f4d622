// Codebook-gather dequantisation: codes -> dense fp16/bf16 weights.
//
// Replaces aqlm_dequant (reference csrc/quantization/aqlm/gemm_kernels.cu:544-597,
// kernels Code1x16Dequant :185-232 and Code2x8Dequant :234-295) and extends it to
// the whole (nbits, d, m) grid that the reference only defines through
// dequantize_weight (vllm/model_executor/layers/quantization/aqlm.py:37-74).
//
// Semantics (bit-exact gate P1/P2):
//   w[n, g*d + j] = round_T( sum_{m, in order, fp32} cb[p*M + m][code(n,g,m)][j] )
//   optional:       round_T( float(w) * float(scales[n]) )          (aqlm.py:70-72)
//
// Roofline: HBM write-bound, 2 B/weight out vs <= 2/d B/weight in.  One thread per
// group: a warp writes 32*d*2 contiguous bytes (512 B for d=8) with 128-bit
// no-allocate stores; the flat (row, group) index is contiguous in both codes and
// output so no division is needed unless scales are applied.
#include "pq_common.cuh"

namespace pq {

template <int D> struct WVec { uint32_t w[D / 2]; };

template <int D>
__device__ __forceinline__ WVec<D> ld_entry_smem(const uint32_t* p) {
  WVec<D> v;
  if constexpr (D % 8 == 0) {
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      uint4 q = reinterpret_cast<const uint4*>(p)[i];
      v.w[4 * i] = q.x; v.w[4 * i + 1] = q.y; v.w[4 * i + 2] = q.z; v.w[4 * i + 3] = q.w;
    }
  } else if constexpr (D % 4 == 0) {
#pragma unroll
    for (int i = 0; i < D / 4; ++i) {
      uint2 q = reinterpret_cast<const uint2*>(p)[i];
      v.w[2 * i] = q.x; v.w[2 * i + 1] = q.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < D / 2; ++i) v.w[i] = p[i];
  }
  return v;
}

template <int D>
__device__ __forceinline__ WVec<D> ld_entry_gmem(const uint32_t* p) {
  WVec<D> v;
  if constexpr (D % 8 == 0) {
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      uint4 q = ld_cg_16(reinterpret_cast<const uint4*>(p) + i);   // L2 gather, bypass L1
      v.w[4 * i] = q.x; v.w[4 * i + 1] = q.y; v.w[4 * i + 2] = q.z; v.w[4 * i + 3] = q.w;
    }
  } else if constexpr (D % 4 == 0) {
#pragma unroll
    for (int i = 0; i < D / 4; ++i) {
      uint2 q = ld_cg_8(reinterpret_cast<const uint2*>(p) + i);
      v.w[2 * i] = q.x; v.w[2 * i + 1] = q.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < D / 2; ++i) v.w[i] = __ldcg(p + i);
  }
  return v;
}

template <int D>
__device__ __forceinline__ void st_group(void* p, const WVec<D>& v) {
  if constexpr (D % 8 == 0) {
#pragma unroll
    for (int i = 0; i < D / 8; ++i)
      st_stream_16(reinterpret_cast<uint4*>(p) + i,
                   make_uint4(v.w[4 * i], v.w[4 * i + 1], v.w[4 * i + 2], v.w[4 * i + 3]));
  } else if constexpr (D % 4 == 0) {
#pragma unroll
    for (int i = 0; i < D / 4; ++i)
      reinterpret_cast<uint2*>(p)[i] = make_uint2(v.w[2 * i], v.w[2 * i + 1]);
  } else {
#pragma unroll
    for (int i = 0; i < D / 2; ++i) reinterpret_cast<uint32_t*>(p)[i] = v.w[i];
  }
}

constexpr int kDqThreads = 256;
constexpr int kDqUnroll = 4;

template <typename T, int D, typename CodeT, bool CB_SMEM>
__global__ void __launch_bounds__(kDqThreads)
pq_dequant_kernel(const CodeT* __restrict__ codes, const T* __restrict__ codebooks,
                  const T* __restrict__ scales, T* __restrict__ out,
                  int G, int M, int E, Parts parts, int rows_per_cta) {
  extern __shared__ uint4 dq_smem[];
  const RowBlock rb = locate_row_block(parts, blockIdx.x, rows_per_cta);
  const uint32_t* cb =
      reinterpret_cast<const uint32_t*>(codebooks + (size_t)rb.part * M * E * D);
  if constexpr (CB_SMEM) {
    const int n16 = M * E * D * 2 / 16;
    const uint4* src = reinterpret_cast<const uint4*>(cb);
    for (int i = threadIdx.x; i < n16; i += kDqThreads) dq_smem[i] = src[i];
    __syncthreads();
    cb = reinterpret_cast<const uint32_t*>(dq_smem);
  }
  const uint32_t emask = (uint32_t)E - 1u;
  const size_t i0 = (size_t)rb.row0 * G, i1 = (size_t)rb.row1 * G;

  for (size_t base = i0 + threadIdx.x; base < i1; base += (size_t)kDqThreads * kDqUnroll) {
    float acc[kDqUnroll][D];
#pragma unroll
    for (int u = 0; u < kDqUnroll; ++u)
#pragma unroll
      for (int j = 0; j < D; ++j) acc[u][j] = 0.f;

    for (int m = 0; m < M; ++m) {
      uint32_t code[kDqUnroll];
#pragma unroll
      for (int u = 0; u < kDqUnroll; ++u) {
        const size_t i = base + (size_t)u * kDqThreads;
        code[u] = i < i1 ? ((uint32_t)codes[i * M + m] & emask) : 0u;
      }
#pragma unroll
      for (int u = 0; u < kDqUnroll; ++u) {
        const uint32_t* ent = cb + ((size_t)m * E + code[u]) * (D / 2);
        WVec<D> v = CB_SMEM ? ld_entry_smem<D>(ent) : ld_entry_gmem<D>(ent);
#pragma unroll
        for (int j = 0; j < D / 2; ++j) {
          float2 f = Traits<T>::to_f2(v.w[j]);
          acc[u][2 * j] += f.x;          // fp32, codebook order (aqlm.py:60-65)
          acc[u][2 * j + 1] += f.y;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kDqUnroll; ++u) {
      const size_t i = base + (size_t)u * kDqThreads;
      if (i >= i1) continue;
      WVec<D> o;
      if (scales != nullptr) {
        const float s = Traits<T>::to_f(scales[i / G]);
#pragma unroll
        for (int j = 0; j < D / 2; ++j) {
          // round the sum first, then the (exact in fp32) product: aqlm.py:70-72
          float2 r = Traits<T>::to_f2(Traits<T>::pack(acc[u][2 * j], acc[u][2 * j + 1]));
          o.w[j] = Traits<T>::pack(r.x * s, r.y * s);
        }
      } else {
#pragma unroll
        for (int j = 0; j < D / 2; ++j) o.w[j] = Traits<T>::pack(acc[u][2 * j], acc[u][2 * j + 1]);
      }
      st_group<D>(out + i * D, o);
    }
  }
}

// Any even d (scalar pairs); not a performance path.
template <typename T, typename CodeT>
__global__ void __launch_bounds__(kDqThreads)
pq_dequant_generic_kernel(const CodeT* __restrict__ codes, const T* __restrict__ codebooks,
                          const T* __restrict__ scales, T* __restrict__ out,
                          int G, int D, int M, int E, Parts parts, int rows_per_cta) {
  const RowBlock rb = locate_row_block(parts, blockIdx.x, rows_per_cta);
  const T* cb = codebooks + (size_t)rb.part * M * E * D;
  const uint32_t emask = (uint32_t)E - 1u;
  const size_t e0 = (size_t)rb.row0 * G * D, e1 = (size_t)rb.row1 * G * D;
  for (size_t e = e0 + threadIdx.x; e < e1; e += kDqThreads) {
    const size_t grp = e / D;
    const int j = (int)(e - grp * D);
    float acc = 0.f;
    for (int m = 0; m < M; ++m) {
      const uint32_t code = (uint32_t)codes[grp * M + m] & emask;
      acc += Traits<T>::to_f(cb[((size_t)m * E + code) * D + j]);
    }
    T w = Traits<T>::from_f(acc);
    if (scales != nullptr)
      w = Traits<T>::from_f(Traits<T>::to_f(w) * Traits<T>::to_f(scales[grp / G]));
    out[e] = w;
  }
}

template <typename T, int D, typename CodeT>
static int launch_dequant_d(const void* codes, const void* codebooks, const void* scales, void* out,
                            int64_t n, int G, int M, int E, const int64_t* part_sizes, int n_parts,
                            cudaStream_t stream) {
  const int sms = sm_count();
  const size_t cb_bytes = (size_t)M * E * D * 2;
  const bool cb_smem = cb_bytes <= 100 * 1024;      // >= 2 CTAs / SM
  const int ctas_per_sm = cb_smem ? (cb_bytes > 48 * 1024 ? 2 : 4) : 8;
  int rows_per_cta = (int)((n + (int64_t)sms * ctas_per_sm - 1) / ((int64_t)sms * ctas_per_sm));
  if (rows_per_cta < 1) rows_per_cta = 1;
  Parts ps;
  int rc = make_parts(ps, part_sizes, n_parts, n, rows_per_cta);
  if (rc) return rc;
  const int grid = ps.cta_end[n_parts - 1];
  if (cb_smem) {
    auto kern = pq_dequant_kernel<T, D, CodeT, true>;
    if (cb_bytes > 48 * 1024)
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cb_bytes);
    kern<<<grid, kDqThreads, cb_bytes, stream>>>((const CodeT*)codes, (const T*)codebooks,
                                                 (const T*)scales, (T*)out, G, M, E, ps, rows_per_cta);
  } else {
    pq_dequant_kernel<T, D, CodeT, false><<<grid, kDqThreads, 0, stream>>>(
        (const CodeT*)codes, (const T*)codebooks, (const T*)scales, (T*)out, G, M, E, ps, rows_per_cta);
  }
  return (int)cudaGetLastError();
}

template <typename T, typename CodeT>
static int launch_dequant_t(const void* codes, const void* codebooks, const void* scales, void* out,
                            int64_t n, int64_t k, int d, int M, int E, const int64_t* part_sizes,
                            int n_parts, cudaStream_t stream) {
  const int G = (int)(k / d);
  switch (d) {
    case 4:  return launch_dequant_d<T, 4, CodeT>(codes, codebooks, scales, out, n, G, M, E, part_sizes, n_parts, stream);
    case 8:  return launch_dequant_d<T, 8, CodeT>(codes, codebooks, scales, out, n, G, M, E, part_sizes, n_parts, stream);
    case 16: return launch_dequant_d<T, 16, CodeT>(codes, codebooks, scales, out, n, G, M, E, part_sizes, n_parts, stream);
    default: break;
  }
  const int sms = sm_count();
  int rows_per_cta = (int)((n + (int64_t)sms * 8 - 1) / ((int64_t)sms * 8));
  if (rows_per_cta < 1) rows_per_cta = 1;
  Parts ps;
  int rc = make_parts(ps, part_sizes, n_parts, n, rows_per_cta);
  if (rc) return rc;
  pq_dequant_generic_kernel<T, CodeT><<<ps.cta_end[n_parts - 1], kDqThreads, 0, stream>>>(
      (const CodeT*)codes, (const T*)codebooks, (const T*)scales, (T*)out, G, d, M, E, ps, rows_per_cta);
  return (int)cudaGetLastError();
}

int dequant_dispatch(const void* codes, const void* codebooks, const void* scales, void* out,
                     int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                     int n_parts, int dtype, cudaStream_t stream) {
  const int E = 1 << nbits;
  if (dtype == PQ_F16) {
    return nbits <= 8
        ? launch_dequant_t<__half, uint8_t>(codes, codebooks, scales, out, n, k, d, m, E, part_sizes, n_parts, stream)
        : launch_dequant_t<__half, uint16_t>(codes, codebooks, scales, out, n, k, d, m, E, part_sizes, n_parts, stream);
  }
  return nbits <= 8
      ? launch_dequant_t<__nv_bfloat16, uint8_t>(codes, codebooks, scales, out, n, k, d, m, E, part_sizes, n_parts, stream)
      : launch_dequant_t<__nv_bfloat16, uint16_t>(codes, codebooks, scales, out, n, k, d, m, E, part_sizes, n_parts, stream);
}

// ---- embedding lookup over PQ weights ----------------------------------------
// out[i, :] = scaled dequantised row ids[i] of a [n, k] PQ matrix - what QuantizeMethodBase.embedding
// has to return for a quantised VocabParallelEmbedding (base_config.py:30-35,
// vocab_parallel_embedding.py:41-43,229-246: F.embedding over the weight).  The same bits as the
// corresponding row of pq_b200_dequant with scales (fp32 sum over the M codewords, one rounding,
// scale multiply rounded once more: aqlm.py:60-72).  Ids outside [0, n) give a zero row (the caller
// masks the ids of other vocabulary shards, vocab_parallel_embedding.py:395-404).  One CTA per id,
// one group per thread; codewords come straight from global memory / L2: a decode step looks up
// a handful of rows, staging a codebook in shared memory would cost more than it saves.
template <typename T, int D, typename CodeT>
__global__ void __launch_bounds__(256)
pq_embedding_kernel(const long long* __restrict__ ids, const CodeT* __restrict__ codes, const T* __restrict__ codebooks,
                    const T* __restrict__ scales, T* __restrict__ out, long long out_stride, int n, int G, int M, int E) {
  const long long id = ids[blockIdx.x];
  T* o = out + (size_t)blockIdx.x * out_stride;
  const bool ok = id >= 0 && id < n;
  const uint32_t emask = (uint32_t)E - 1u;
  const float sc = ok ? Traits<T>::to_f(scales[id]) : 0.f;
  for (int g = threadIdx.x; g < G; g += 256) {
    WVec<D> r;
#pragma unroll
    for (int j = 0; j < D / 2; ++j) r.w[j] = 0u;
    if (ok) {
      float acc[D];
#pragma unroll
      for (int j = 0; j < D; ++j) acc[j] = 0.f;
      const CodeT* c = codes + ((size_t)id * G + g) * M;
      for (int m = 0; m < M; ++m) {
        const uint32_t code = (uint32_t)c[m] & emask;
        const WVec<D> v = ld_entry_gmem<D>(reinterpret_cast<const uint32_t*>(codebooks) + ((size_t)m * E + code) * (D / 2));
#pragma unroll
        for (int j = 0; j < D / 2; ++j) {
          const float2 f = Traits<T>::to_f2(v.w[j]);
          acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
        }
      }
#pragma unroll
      for (int j = 0; j < D / 2; ++j) {
        // round the sum (aqlm.py:60-65), then the scale multiply in the weight dtype (aqlm.py:70-72)
        const float2 w = Traits<T>::to_f2(Traits<T>::pack(acc[2 * j], acc[2 * j + 1]));
        r.w[j] = Traits<T>::pack(w.x * sc, w.y * sc);
      }
    }
    st_group<D>(o + (size_t)g * D, r);
  }
}

template <typename T, typename CodeT>
static int launch_embedding(const long long* ids, int64_t n_ids, const void* codes, const void* codebooks,
                            const void* scales, void* out, int64_t out_stride, int64_t n, int64_t k, int d, int nbits,
                            int m, cudaStream_t stream) {
  const int G = (int)(k / d), E = 1 << nbits;
#define PQ_EMB(DD)                                                                                              \
  pq_embedding_kernel<T, DD, CodeT><<<(unsigned)n_ids, 256, 0, stream>>>(                                       \
      ids, (const CodeT*)codes, (const T*)codebooks, (const T*)scales, (T*)out, (long long)out_stride, (int)n, G, m, E)
  switch (d) {
    case 4: PQ_EMB(4); break;
    case 8: PQ_EMB(8); break;
    case 16: PQ_EMB(16); break;
    default: return PQ_ERR_FORMAT;
  }
#undef PQ_EMB
  return (int)cudaGetLastError();
}

int embedding_dispatch(const long long* ids, int64_t n_ids, const void* codes, const void* codebooks,
                       const void* scales, void* out, int64_t out_stride, int64_t n, int64_t k, int d, int nbits,
                       int m, int dtype, cudaStream_t stream) {
  if (dtype == PQ_F16)
    return nbits <= 8 ? launch_embedding<__half, uint8_t>(ids, n_ids, codes, codebooks, scales, out, out_stride, n, k, d, nbits, m, stream)
                      : launch_embedding<__half, uint16_t>(ids, n_ids, codes, codebooks, scales, out, out_stride, n, k, d, nbits, m, stream);
  return nbits <= 8 ? launch_embedding<__nv_bfloat16, uint8_t>(ids, n_ids, codes, codebooks, scales, out, out_stride, n, k, d, nbits, m, stream)
                    : launch_embedding<__nv_bfloat16, uint16_t>(ids, n_ids, codes, codebooks, scales, out, out_stride, n, k, d, nbits, m, stream);
}

}  // namespace pq
