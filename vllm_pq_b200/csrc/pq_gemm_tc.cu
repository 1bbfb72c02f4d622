// Prefill-path PQ GEMM on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
//   y[t, n] = scale[n] * sum_k x[t, k] * W[n, k] + bias[n],   W = codebook gather of `codes`
//
// Replaces aqlm_dequant + F.linear (+ the in-place scale) of optimized_dequantize_gemm
// (reference vllm/model_executor/layers/quantization/aqlm.py:132-156 and
// csrc/quantization/aqlm/gemm_kernels.cu:185-295,544-597): the dense weight never exists in
// HBM.  A persistent, warp-specialised CTA per SM works on tiles of 128 weight rows x BN tokens:
//   warp 0      TMA producer: x tile [BN tokens][64 k] bf16/fp16 -> shared memory (128-byte
//               swizzle) with cp.async.bulk.tensor, completion on an mbarrier
//   warps 4-7, 12-15   two groups of dequant producers that take alternate code chunks (one or two
//               64-wide k blocks), each filling its own pipeline stage: a group's k block is one
//               serial chain (codes -> gathers -> stores -> proxy fence -> arrive) that a single warp
//               per scheduler cannot hide (r1: ~1200 cycles per k block against 524 of MMA time);
//               two groups overlap two chains.  One weight row per thread; per 64-wide k block they read
//               the row's codes (8-16 B), gather the codewords (shared-memory codebooks, or L2
//               for 65536-entry books), sum additive codebooks in fp32 with ONE rounding - the
//               same bits pq_b200_dequant writes - and store them as the A operand in the
//               K-major 128-byte-swizzled layout the UMMA descriptor describes (16-byte chunk c
//               of row r at r*128 + ((c ^ (r & 7)) << 4)); fence.proxy.async, then mbarrier arrive
//   warp 1      one thread issues tcgen05.mma (M = 128 weight rows, N = BN tokens, K = 16) with
//               the fp32 accumulator in tensor memory; tcgen05.commit frees the stage / publishes
//               the accumulator
//   warps 8-11  epilogue: tcgen05.ld (lane = weight row, column = token), scale and bias in fp32,
//               one rounding, stores of 32 consecutive outputs per token per warp
// Two accumulator stages (2 x BN <= 512 TMEM columns) overlap the epilogue of one tile with the
// main loop of the next.  Weight rows are the MMA M operand so that decode-sized token counts
// (BN = 16..64) still fill the 128-lane datapath.
#include <cuda.h>

#include <cstdlib>
#include <type_traits>

#include "pq_common.cuh"

namespace pq {

constexpr int kTcGroups = 2;               // dequant warp groups (4 warps each), alternating code chunks
constexpr int kTcThreads = 384 + 128 * (kTcGroups - 1);
constexpr int kTcBM = 128;                 // weight rows per tile = UMMA M
constexpr int kTcBK = 64;                  // k per stage: 128 bytes = one swizzle atom
constexpr int kTcMaxStages = 8;
constexpr int kTcABytes = kTcBM * kTcBK * 2;
constexpr int kTcTmemCols = 512;
constexpr int kTcMaxSplit = 8;             // K splits of a few-tile problem
constexpr int kTcCodeRing = 4;             // 16..64-byte code chunks in flight per dequant thread
constexpr int kTcRingVec = 128 * kTcGroups * 16;   // one 16-byte vector per dequant thread

struct TcArgs {
  const void* codes; const void* codebooks; const void* scales; const void* bias; void* y;
  long long y_stride;
  int T, N, K, E, M;
  int BN;                                  // tokens per tile (UMMA N), multiple of 16
  int n_tok_tiles, n_row_tiles;
  int stages;
  int groups;                              // dequant groups at work (1 or kTcGroups), see tc_plan
  int ksplit;                              // CTAs along K per (token tile, row tile): few-tile problems (decode-sized
                                           // batches on the small-N layers) otherwise leave most SMs idle
  float* ws_part;                          // ksplit > 1: [ksplit][T][N] fp32 partial sums (workspace)
  int* ws_tickets;                         // ksplit > 1: two counters per (token tile, row tile, epilogue warp); zero between launches
  unsigned long long* trace;               // PQ_TC_TRACE builds: 16 stamps per CTA
  int dense_w;                             // 1: the A operand is TMA-loaded from a dense [N, K] weight (tmap_w) that
                                           // pq_b200_dequant wrote into the workspace; the dequant warps idle
  int cb_smem_bytes;                       // 0: gather codewords from global memory (L2)
  int cb_stride;                           // bytes between codebook entries as staged (128 when replicated)
  int row_code_bytes;                      // bytes of codes per weight row
  Parts parts;                             // cta_end counts 128-row tiles
};

// ---- PTX wrappers ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t tc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld_x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte swizzle, dense 8-row groups (1024 B apart): LBO = 1 (unused), SBO = 64,
// descriptor version 1 (sm_100), layout type 2 (SWIZZLE_128B).  cute/arch/mma_sm100_desc.hpp.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor: fp32 accumulate, A and B K-major, M = 128, N = bn.
__device__ __forceinline__ uint32_t umma_idesc(int bn, bool bf16) {
  const uint32_t fmt = bf16 ? 1u : 0u;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(kTcBM >> 4) << 24);
}

// ---- dequant producer helpers -----------------------------------------------------------------
template <int D> struct TcCw { uint32_t w[D / 2]; };

// Codeword fetch: explicit shared-space vector loads (a generic pointer makes the compiler split the
// 16-byte access into three narrower ones), or L2 (ld.global.cg) for books that do not fit.
template <int D>
__device__ __forceinline__ TcCw<D> tc_ld_cw(const unsigned char* p, bool in_smem) {
  TcCw<D> v;
  if (in_smem) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(p);
    if constexpr (D == 4) {
      asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.w[0]), "=r"(v.w[1]) : "r"(sa));
    } else {
#pragma unroll
      for (int i = 0; i < D / 8; ++i)
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v.w[4 * i]), "=r"(v.w[4 * i + 1]), "=r"(v.w[4 * i + 2]), "=r"(v.w[4 * i + 3])
                     : "r"(sa + 16 * i));
    }
  } else {
    if constexpr (D == 4) {
      uint2 q = ld_cg_8(p);
      v.w[0] = q.x; v.w[1] = q.y;
    } else {
#pragma unroll
      for (int i = 0; i < D / 8; ++i) {
        uint4 q = ld_cg_16(reinterpret_cast<const uint4*>(p) + i);
        v.w[4 * i] = q.x; v.w[4 * i + 1] = q.y; v.w[4 * i + 2] = q.z; v.w[4 * i + 3] = q.w;
      }
    }
  }
  return v;
}

// Store one group's D elements into the K-major 128B-swizzled A tile: row r, k offset g*D.
template <int D>
__device__ __forceinline__ void tc_st_group(uint32_t a_row, int r7, int g, const TcCw<D>& v) {
  if constexpr (D == 8) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(a_row + (uint32_t)((g ^ r7) << 4)),
                 "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3]) : "memory");
  } else if constexpr (D == 4) {
    asm volatile("st.shared.v2.u32 [%0], {%1,%2};" :: "r"(a_row + (uint32_t)((((g >> 1) ^ r7) << 4) + ((g & 1) << 3))),
                 "r"(v.w[0]), "r"(v.w[1]) : "memory");
  } else {
#pragma unroll
    for (int h = 0; h < 2; ++h)
      asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(a_row + (uint32_t)(((2 * g + h) ^ r7) << 4)),
                   "r"(v.w[4 * h]), "r"(v.w[4 * h + 1]), "r"(v.w[4 * h + 2]), "r"(v.w[4 * h + 3]) : "memory");
  }
}

template <int D>
__device__ __forceinline__ void tc_zero_block(uint32_t a_row, int r7) {
  TcCw<D> z;
#pragma unroll
  for (int j = 0; j < D / 2; ++j) z.w[j] = 0u;
#pragma unroll
  for (int g = 0; g < kTcBK / D; ++g) tc_st_group<D>(a_row, r7, g, z);
}

template <typename CodeT>
__device__ __forceinline__ uint32_t tc_code_at(const uint32_t* cw, int idx) {   // idx constant after unrolling
  if constexpr (sizeof(CodeT) == 1) return (cw[idx >> 2] >> (8 * (idx & 3))) & 0xffu;
  else return (cw[idx >> 1] >> (16 * (idx & 1))) & 0xffffu;
}

// Sum of the M codewords of one group in fp32 (codebook order), ONE RNE rounding: the bits
// pq_b200_dequant produces.
template <typename T, int D>
__device__ __forceinline__ void tc_accumulate(float (&acc)[D], const TcCw<D>& v) {
#pragma unroll
  for (int j = 0; j < D / 2; ++j) {
    const float2 f = Traits<T>::to_f2(v.w[j]);
    acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
  }
}

// One weight row, one 64-wide k block, the block's (64 / D) * M codes already in registers.
template <typename T, int D, typename CodeT, int M>
__device__ __forceinline__ void tc_dequant_regs(uint32_t a_row, int r7, const uint32_t* cw, int E,
                                                const unsigned char* cb, uint32_t cb_stride, bool cb_in_smem) {
  constexpr int GPB = kTcBK / D;
  const uint32_t emask = (uint32_t)E - 1u;
  // all gathers of a batch are issued before the first store: the loads overlap instead of each
  // store waiting a full shared-memory round trip (the dominant stall in the first ncu capture)
  constexpr int GB = (M == 1) ? (GPB < 8 ? GPB : 8) : (8 / M);      // groups per batch
#pragma unroll
  for (int g0 = 0; g0 < GPB; g0 += GB) {
    TcCw<D> v[GB][M];
#pragma unroll
    for (int g = 0; g < GB; ++g)
#pragma unroll
      for (int m = 0; m < M; ++m) {
        const uint32_t code = tc_code_at<CodeT>(cw, (g0 + g) * M + m) & emask;
        v[g][m] = tc_ld_cw<D>(cb + ((size_t)m * E + code) * cb_stride, cb_in_smem);
      }
#pragma unroll
    for (int g = 0; g < GB; ++g) {
      if constexpr (M == 1) {
        tc_st_group<D>(a_row, r7, g0 + g, v[g][0]);
      } else if constexpr (M == 2) {
        // two codewords: the packed add IS the fp32 sum rounded once (a sum of two fp16 / bf16 values is
        // exact in fp32) - the same bits as the general path, a fifth of its instructions
        using T2 = typename Traits<T>::T2;
        TcCw<D> o;
#pragma unroll
        for (int j = 0; j < D / 2; ++j) {
          const T2 r2 = __hadd2(*reinterpret_cast<const T2*>(&v[g][0].w[j]), *reinterpret_cast<const T2*>(&v[g][1].w[j]));
          o.w[j] = *reinterpret_cast<const uint32_t*>(&r2);
        }
        tc_st_group<D>(a_row, r7, g0 + g, o);
      } else {
        float acc[D];
#pragma unroll
        for (int j = 0; j < D; ++j) acc[j] = 0.f;
#pragma unroll
        for (int m = 0; m < M; ++m) tc_accumulate<T, D>(acc, v[g][m]);
        TcCw<D> o;
#pragma unroll
        for (int j = 0; j < D / 2; ++j) o.w[j] = Traits<T>::pack(acc[2 * j], acc[2 * j + 1]);
        tc_st_group<D>(a_row, r7, g0 + g, o);
      }
    }
  }
}

// Any M, codes read one by one from global memory (rows that are not 16-byte multiples, M = 3, ...).
template <typename T, int D, typename CodeT>
__device__ __forceinline__ void tc_dequant_generic(uint32_t a_row, int r7, const CodeT* cp, int M, int E,
                                                   const unsigned char* cb, uint32_t cb_stride, bool cb_in_smem) {
  constexpr int GPB = kTcBK / D;
  const uint32_t emask = (uint32_t)E - 1u;
#pragma unroll 2
  for (int g = 0; g < GPB; ++g) {
    float acc[D];
#pragma unroll
    for (int j = 0; j < D; ++j) acc[j] = 0.f;
    TcCw<D> v;
    for (int m = 0; m < M; ++m) {
      const uint32_t code = (uint32_t)cp[g * M + m] & emask;
      v = tc_ld_cw<D>(cb + ((size_t)m * E + code) * cb_stride, cb_in_smem);
      tc_accumulate<T, D>(acc, v);
    }
    if (M > 1) {
#pragma unroll
      for (int j = 0; j < D / 2; ++j) v.w[j] = Traits<T>::pack(acc[2 * j], acc[2 * j + 1]);
    }
    tc_st_group<D>(a_row, r7, g, v);
  }
}

#ifdef PQ_TC_TRACE
extern unsigned long long* g_lut_trace;
#define TC_STAMP(buf, slot)                                                            \
  do { if ((buf) != nullptr) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); \
       (buf)[blockIdx.x * 16 + (slot)] = t_; } } while (0)
#else
#define TC_STAMP(buf, slot) do {} while (0)
#endif

// ---- the kernel ---------------------------------------------------------------------------------
// MT: codebooks per group known at compile time (1, 2, 4: codes prefetched through a per-thread
// cp.async ring) or 0 (any M, codes read from global memory as needed).
template <typename T, int D, typename CodeT, int MT>
__global__ void __launch_bounds__(kTcThreads, 1)
pq_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                  const TcArgs a) {
  extern __shared__ __align__(1024) unsigned char tc_smem[];
  __shared__ __align__(8) uint64_t bars[3 * kTcMaxStages + 4];
  __shared__ uint32_t tmem_base_s;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) TC_STAMP(a.trace, 0);
  // dynamic shared memory: [A stages | B stages | codebooks], 1024-byte aligned tiles
  const uint32_t smem0 = (tc_smem_u32(tc_smem) + 1023u) & ~1023u;
  const uint32_t b_bytes = (uint32_t)a.BN * 128u;
  const uint32_t a_base = smem0;
  const uint32_t b_base = smem0 + (uint32_t)a.stages * kTcABytes;
  unsigned char* cb_smem = tc_smem + (smem0 - tc_smem_u32(tc_smem)) + (size_t)a.stages * (kTcABytes + b_bytes);

  const uint32_t bar0 = tc_smem_u32(bars);
  auto full_a = [&](int s) { return bar0 + 8u * s; };
  auto full_b = [&](int s) { return bar0 + 8u * (kTcMaxStages + s); };
  auto empty = [&](int s) { return bar0 + 8u * (2 * kTcMaxStages + s); };
  auto tmem_full = [&](int i) { return bar0 + 8u * (3 * kTcMaxStages + i); };
  auto tmem_empty = [&](int i) { return bar0 + 8u * (3 * kTcMaxStages + 2 + i); };

  if (threadIdx.x == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full_a(s), a.dense_w ? 1 : 4); mbar_init(full_b(s), 1); mbar_init(empty(s), 1);
    }
    for (int i = 0; i < 2; ++i) { mbar_init(tmem_full(i), 1); mbar_init(tmem_empty(i), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" :: "l"(&tmap_x) : "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 :: "r"(tc_smem_u32(&tmem_base_s)), "n"(kTcTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  if (threadIdx.x == 0) TC_STAMP(a.trace, 1);

  const int KB = a.K / kTcBK / a.ksplit;            // k blocks per tile (per K split)
  const int tiles_per_split = a.n_tok_tiles * a.n_row_tiles;
  const int total_tiles = tiles_per_split * a.ksplit;   // tile = (ks * n_tok_tiles + tt) * n_row_tiles + rt

  if (warp == 0) {
    // ===== TMA producer: x tiles =====
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int tt = (tile / a.n_row_tiles) % a.n_tok_tiles, kb0 = (tile / tiles_per_split) * KB;
        const int w_row0 = a.dense_w ? locate_row_block(a.parts, tile % a.n_row_tiles, kTcBM).row0 : 0;
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(empty(s), ph ^ 1u);
          mbar_arrive_expect_tx(full_b(s), b_bytes);
          tma_load_2d(b_base + (uint32_t)s * b_bytes, &tmap_x, (kb0 + kb) * kTcBK, tt * a.BN, full_b(s));
          if (a.dense_w) {                           // the weight tile too: [128 rows][64 k], same swizzle the dequant warps write
            mbar_arrive_expect_tx(full_a(s), kTcABytes);
            tma_load_2d(a_base + (uint32_t)s * kTcABytes, &tmap_w, (kb0 + kb) * kTcBK, w_row0, full_a(s));
          }
          if (++s == a.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(a.BN, std::is_same<T, __nv_bfloat16>::value);
      int s = 0; uint32_t ph = 0; int acc = 0; uint32_t acc_ph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(tmem_empty(acc), acc_ph ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)acc * 256u;
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(full_a(s), ph);
          if (kb == 0) TC_STAMP(a.trace, 2);
          mbar_wait(full_b(s), ph);
          if (kb == 0) TC_STAMP(a.trace, 3);
          tc_fence_after();
          const uint64_t ad = umma_desc_sw128(a_base + (uint32_t)s * kTcABytes);
          const uint64_t bd = umma_desc_sw128(b_base + (uint32_t)s * b_bytes);
#pragma unroll
          for (int k = 0; k < kTcBK / 16; ++k)       // +32 bytes (2 x 16 B) per K = 16 step inside the swizzle atom
            umma_f16(d_tmem, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(empty(s));                     // frees the stage when these MMAs retire
          if (++s == a.stages) { s = 0; ph ^= 1u; }
        }
        umma_commit(tmem_full(acc));                 // accumulator complete
        TC_STAMP(a.trace, 4);
        if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
      }
    }
  } else if ((warp >= 4 && warp < 8) || warp >= 12) {
    // ===== dequant producers: one weight row per thread, group `grp` takes chunks grp, grp + G, ... =====
    const int grp = warp >= 12 ? 1 + ((warp - 12) >> 2) : 0;
    const int G = a.dense_w ? 0 : a.groups;        // dense A operand: nothing to dequantise
    if (grp < G) {
    constexpr int GPB = kTcBK / D;
    constexpr int CBYTES = MT > 0 ? GPB * MT * (int)sizeof(CodeT) : 16;   // code bytes per row and k block
    constexpr int CHB = CBYTES > 16 ? CBYTES : 16;                         // ring chunk bytes per thread
    constexpr int CH = CHB / CBYTES;                                       // k blocks per chunk
    constexpr int NV = CHB / 16;
    constexpr int NW = CBYTES / 4;
    const int r = (threadIdx.x & 127);               // 128-thread groups start at multiples of 128
    const int r7 = r & 7;
    const bool cb_in_smem = a.cb_smem_bytes > 0;
    const uint32_t cb_stride = (uint32_t)a.cb_stride;
    const int repl = a.cb_stride / (D * 2);          // copies of each entry in its 128-byte line (1: flat)
    const uint32_t lane_off = (uint32_t)(lane & (repl - 1)) * (D * 2);
    const size_t part_cb_bytes = (size_t)a.M * a.E * D * 2;
    const uint32_t ring0 = tc_smem_u32(cb_smem + a.cb_smem_bytes) + (uint32_t)(grp * 128 + r) * 16u;
    const int n_chunks = KB / CH;
    const int gtid = grp * 128 + r;                  // index among all dequant threads
    int cur_part = -1;

    // ---- issue cursor of the code ring: this group's chunks, kTcCodeRing ahead of the consumer.
    // Chunks are numbered through the CTA's tiles; group g owns numbers g, g + G, g + 2G, ...
    int itile = blockIdx.x, ichunk = grp, islot = 0;
    const unsigned char* irow = nullptr;
    auto seek = [&]() {
      while (itile < total_tiles && ichunk >= n_chunks) { ichunk -= n_chunks; itile += gridDim.x; }
      irow = nullptr;
      if (itile < total_tiles) {
        const RowBlock rb = locate_row_block(a.parts, itile % a.n_row_tiles, kTcBM);
        if (rb.row0 + r < rb.row1)
          irow = reinterpret_cast<const unsigned char*>(a.codes) + (size_t)(rb.row0 + r) * a.row_code_bytes +
                 (size_t)(itile / tiles_per_split) * n_chunks * CHB;      // this K split's first chunk
      }
    };
    auto issue = [&]() {
      if constexpr (MT > 0) {
        if (itile < total_tiles) {
          if (irow != nullptr) {
#pragma unroll
            for (int v = 0; v < NV; ++v)
              asm volatile("cp.async.cg.shared.global [%0], [%1], 16;"
                           :: "r"(ring0 + (uint32_t)((islot * NV + v) * kTcRingVec)),
                              "l"(irow + (size_t)ichunk * CHB + 16 * v) : "memory");
          }
          ichunk += G;
          if (ichunk >= n_chunks) seek();
          islot = (islot + 1 == kTcCodeRing) ? 0 : islot + 1;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      }
    };
    seek();
#pragma unroll
    for (int j = 0; j < kTcCodeRing; ++j) issue();
    int cslot = 0;

    // ---- consumer: walk every tile of this CTA (all groups meet at the codebook reloads), dequantise
    // this group's chunks.  (s, ph) is the pipeline stage of k block `cur`, counted through the tiles.
    int s = 0; uint32_t ph = 0; int cur = 0;
    int ch = grp, tile_k0 = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, tile_k0 += KB) {
      const int rt = tile % a.n_row_tiles;
      const RowBlock rb = locate_row_block(a.parts, rt, kTcBM);
      const int row = rb.row0 + r;
      const bool valid = row < rb.row1;
      const unsigned char* cb_g = reinterpret_cast<const unsigned char*>(a.codebooks) + (size_t)rb.part * part_cb_bytes;
      if (cb_in_smem && rb.part != cur_part) {
        asm volatile("bar.sync 1, %0;" :: "r"(128 * G) : "memory");            // everyone is done with the old books
        // one 16-byte slot of the staged books per thread and step: slot i holds entry i / repl (its
        // `repl` copies are neighbours), so a warp writes 512 consecutive bytes.  (One thread per
        // ENTRY, writing its copies word by word at a 128-byte stride, was a 32-way bank conflict on
        // every store: 4.5 us for 256 entries, 10 us for 512 - most of a decode-sized launch; r2_24.)
        for (int i = gtid; i < a.M * a.E * repl; i += 128 * G) {
          const TcCw<D> v = tc_ld_cw<D>(cb_g + (size_t)(i / repl) * (D * 2), false);
          uint32_t* dst = reinterpret_cast<uint32_t*>(cb_smem + (size_t)i * (D * 2));
          if constexpr (D == 8) {
            *reinterpret_cast<uint4*>(dst) = make_uint4(v.w[0], v.w[1], v.w[2], v.w[3]);
          } else {
#pragma unroll
            for (int j = 0; j < D / 2; ++j) dst[j] = v.w[j];
          }
        }
        asm volatile("bar.sync 1, %0;" :: "r"(128 * G) : "memory");
        cur_part = rb.part;
        if (gtid == 0) TC_STAMP(a.trace, 7);
      }
      const unsigned char* cb = (cb_in_smem ? cb_smem : cb_g) + lane_off;
      const CodeT* crow = reinterpret_cast<const CodeT*>(
          reinterpret_cast<const unsigned char*>(a.codes) + (size_t)(valid ? row : rb.row0) * a.row_code_bytes) +
          (size_t)(tile / tiles_per_split) * KB * GPB * a.M;              // this K split's first code
      for (; ch < n_chunks; ch += G) {
        {                                              // move the stage cursor to this chunk's first k block
          s += tile_k0 + ch * CH - cur;
          cur = tile_k0 + ch * CH;
          while (s >= a.stages) { s -= a.stages; ph ^= 1u; }
        }
        if constexpr (MT > 0) asm volatile("cp.async.wait_group %0;" :: "n"(kTcCodeRing - 1) : "memory");
#pragma unroll
        for (int sub = 0; sub < CH; ++sub) {
          const int kb = ch * CH + sub;
          uint32_t cw[NW];
          if constexpr (MT > 0) {
            if (valid) {
              if constexpr (CBYTES >= 16) {
#pragma unroll
                for (int v = 0; v < NV; ++v)
                  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                               : "=r"(cw[4 * v]), "=r"(cw[4 * v + 1]), "=r"(cw[4 * v + 2]), "=r"(cw[4 * v + 3])
                               : "r"(ring0 + (uint32_t)((cslot * NV + v) * kTcRingVec)));
              } else if constexpr (CBYTES == 8) {
                asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(cw[0]), "=r"(cw[1])
                             : "r"(ring0 + (uint32_t)(cslot * kTcRingVec + sub * 8)));
              } else {
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(cw[0]) : "r"(ring0 + (uint32_t)(cslot * kTcRingVec + sub * 4)));
              }
            }
          }
          mbar_wait(empty(s), ph ^ 1u);
          const uint32_t a_row = a_base + (uint32_t)s * kTcABytes + (uint32_t)r * 128u;
          if (!valid) tc_zero_block<D>(a_row, r7);
          else if constexpr (MT > 0) tc_dequant_regs<T, D, CodeT, (MT > 0 ? MT : 1)>(a_row, r7, cw, a.E, cb, cb_stride, cb_in_smem);
          else tc_dequant_generic<T, D, CodeT>(a_row, r7, crow + (size_t)kb * GPB * a.M, a.M, a.E, cb, cb_stride, cb_in_smem);
          fence_proxy_async();                         // generic-proxy stores -> visible to the tensor core
          __syncwarp();
          if (lane == 0) mbar_arrive(full_a(s));
          if (gtid == 0 && cur == 0) TC_STAMP(a.trace, 8);
          ++cur;
          if (++s == a.stages) { s = 0; ph ^= 1u; }
        }
        if constexpr (MT > 0) { issue(); cslot = (cslot + 1 == kTcCodeRing) ? 0 : cslot + 1; }
      }
      ch -= n_chunks;                                  // this group's first chunk of the next tile
    }
    }  // grp < G
  } else if (warp >= 8) {
    // ===== epilogue: TMEM -> registers -> y =====
    const int w = warp - 8;                          // == warp % 4: the TMEM lanes this warp may read
    int acc = 0; uint32_t acc_ph = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int rt = tile % a.n_row_tiles, tt = (tile / a.n_row_tiles) % a.n_tok_tiles, ks = tile / tiles_per_split;
      const RowBlock rb = locate_row_block(a.parts, rt, kTcBM);
      const int row = rb.row0 + 32 * w + lane;
      const bool valid = row < rb.row1;
      float sc = 0.f, bi = 0.f;
      if (valid) {
        sc = Traits<T>::to_f(reinterpret_cast<const T*>(a.scales)[row]);
        if (a.bias) bi = Traits<T>::to_f(reinterpret_cast<const T*>(a.bias)[row]);
      }
      mbar_wait(tmem_full(acc), acc_ph);
      if (w == 0 && lane == 0) TC_STAMP(a.trace, 5);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)acc * 256u + ((uint32_t)(32 * w) << 16);
      T* yrow = reinterpret_cast<T*>(a.y) + row;
      if (a.ksplit == 1) {
        for (int c0 = 0; c0 < a.BN; c0 += 32) {
          uint32_t v[32];
          tmem_ld_x32(taddr + (uint32_t)c0, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int tok = tt * a.BN + c0 + j;
            if (valid && c0 + j < a.BN && tok < a.T)
              yrow[(size_t)tok * a.y_stride] = Traits<T>::from_f(fmaf(__uint_as_float(v[j]), sc, bi));
          }
        }
      } else {
        // split K: this CTA's fp32 partial sums go to the workspace (lanes = 32 consecutive rows: one
        // 128-byte line per token)
        float* part = a.ws_part + (size_t)ks * a.T * a.N + row;
        for (int c0 = 0; c0 < a.BN; c0 += 32) {
          uint32_t v[32];
          tmem_ld_x32(taddr + (uint32_t)c0, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int tok = tt * a.BN + c0 + j;
            if (valid && c0 + j < a.BN && tok < a.T) __stcg(part + (size_t)tok * a.N, __uint_as_float(v[j]));
          }
        }
        __threadfence();
        __syncwarp();
        // All `ksplit` CTAs of a tile are resident at once (split launches have one tile per CTA and at
        // most one CTA per SM), so they can wait for each other: a ticket counts arrivals, every
        // split then reduces ITS share of the tile's tokens - ksplit warps add and store in parallel
        // instead of the last arriver doing all of it (that serial tail was 15 us at 32 tokens, r2_24).
        // The partials are added in split order: deterministic.  The last to leave zeroes the counters.
        int* ticket = a.ws_tickets + (((size_t)tt * a.n_row_tiles + rt) * 4 + w) * 2;
        if (lane == 0) {
          atomicAdd(ticket, 1);
          while (*reinterpret_cast<volatile int*>(ticket) < a.ksplit) {}
        }
        __syncwarp();
        __threadfence();
        const int per = (a.BN + a.ksplit - 1) / a.ksplit;
        const int j_end = min(a.BN, (ks + 1) * per);
        for (int j0 = ks * per; j0 < j_end; j0 += 8) {
          float pv[kTcMaxSplit][8];
#pragma unroll
          for (int q = 0; q < kTcMaxSplit; ++q)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int tok = tt * a.BN + j0 + j;
              pv[q][j] = (q < a.ksplit && valid && j0 + j < j_end && tok < a.T)
                             ? __ldcg(a.ws_part + ((size_t)q * a.T + tok) * a.N + row) : 0.f;
            }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int tok = tt * a.BN + j0 + j;
            float sum = 0.f;
#pragma unroll
            for (int q = 0; q < kTcMaxSplit; ++q) sum += pv[q][j];        // split order; absent splits add 0
            if (valid && j0 + j < j_end && tok < a.T)
              yrow[(size_t)tok * a.y_stride] = Traits<T>::from_f(fmaf(sum, sc, bi));
          }
        }
        __syncwarp();
        if (lane == 0 && atomicAdd(ticket + 1, 1) == a.ksplit - 1) { ticket[0] = 0; ticket[1] = 0; }
      }
      if (w == 0 && lane == 0) TC_STAMP(a.trace, 6);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty(acc));
      if (++acc == 2) { acc = 0; acc_ph ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) TC_STAMP(a.trace, 9);
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "n"(kTcTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);

static PFN_tmapEncodeTiled tmap_encoder() {
  static PFN_tmapEncodeTiled fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (PFN_tmapEncodeTiled)p;
  }();
  return fn;
}

// x [T, K] row-major (row stride x_stride elements) as a 2-D tensor map: box = 64 k x bn tokens,
// 128-byte swizzle, out-of-bounds tokens read as zero.
static int make_x_tmap(CUtensorMap* map, const void* x, int64_t x_stride, int64_t t, int64_t k, int bn, int dtype) {
  PFN_tmapEncodeTiled enc = tmap_encoder();
  if (!enc) return PQ_ERR_UNSUPPORTED;
  const cuuint64_t gdim[2] = {(cuuint64_t)k, (cuuint64_t)t};
  const cuuint64_t gstride[1] = {(cuuint64_t)x_stride * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kTcBK, (cuuint32_t)bn};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, dtype == PQ_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                         const_cast<void*>(x), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PQ_OK : PQ_ERR_ALIGN;
}

struct TcPlan { int bn, n_tok_tiles, stages, groups, cb_smem, cb_stride, mt, ring_bytes, smem_total, ch; };

static bool tc_plan(TcPlan& pl, int64_t t, int64_t k, int d, int nbits, int m) {
  const int tok_tiles = (int)((t + 255) / 256);
  int bn = (int)((t + tok_tiles - 1) / tok_tiles);
  bn = (bn + 15) / 16 * 16;
  pl.bn = bn;
  pl.n_tok_tiles = (int)((t + bn - 1) / bn);
  // codebooks: replicated to one 128-byte line per entry (conflict-free gathers) if that fits in
  // 64 KiB, else flat in shared memory, else gathered from L2
  const int64_t entries = (int64_t)m * (1 << nbits);
  pl.cb_stride = d * 2; pl.cb_smem = 0;
  if (entries * 128 <= 64 * 1024) { pl.cb_smem = (int)(entries * 128); pl.cb_stride = 128; }
  else if (entries * d * 2 <= 64 * 1024) pl.cb_smem = (int)(entries * d * 2);
  // code ring (compile-time M): rows must be whole 16-byte vectors
  const int code_bytes = nbits <= 8 ? 1 : 2;
  const int64_t row_bytes = k / d * m * code_bytes;
  const int cbytes = kTcBK / d * m * code_bytes;
  pl.mt = ((m == 1 || m == 2 || m == 4) && row_bytes % 16 == 0 && cbytes <= 64) ? m : 0;
  pl.ring_bytes = pl.mt ? kTcCodeRing * (cbytes > 16 ? cbytes : 16) * 128 * kTcGroups : 0;
  const int stage = kTcABytes + bn * 128;
  int stages = (int)((225 * 1024 - 1024 - pl.cb_smem - pl.ring_bytes) / stage);
  if (stages > kTcMaxStages) stages = kTcMaxStages;
  if (stages > k / kTcBK) stages = (int)(k / kTcBK) < 2 ? 2 : (int)(k / kTcBK);
  pl.stages = stages;
  // Two dequant groups take alternate chunks of CH k blocks.  A group's consecutive k blocks are then up
  // to (G - 1) * CH + 1 apart; the mbarrier parity wait on a stage is only unambiguous if that gap does
  // not exceed the ring depth (a waiter may never be two phases ahead), else one group works alone.
  const int ch = pl.mt ? ((cbytes > 16 ? cbytes : 16) / cbytes) : 1;
  static const int force_groups = [] { const char* e = getenv("PQ_TC_GROUPS"); return e ? atoi(e) : 0; }();
  pl.ch = ch;
  pl.groups = ((kTcGroups - 1) * ch + 1 <= stages) ? kTcGroups : 1;
  if (force_groups == 1) pl.groups = 1;
  pl.smem_total = stages * stage + pl.cb_smem + pl.ring_bytes + 1024;
  return stages >= 2;
}

bool gemm_tc_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (t < 1 || n < 1 || k < kTcBK || k % kTcBK) return false;      // whole 64-wide k blocks
  if (d != 4 && d != 8 && d != 16) return false;
  if (m < 1 || m > 8 || nbits < 1 || nbits > 16) return false;
  TcPlan pl;
  return tc_plan(pl, t, k, d, nbits, m);
}

constexpr int kTcTicketBytes = 64 * 1024;          // head of the workspace (include/pq_b200.h), shared with the table GEMV

// K splits for problems with few tiles: the largest split <= SMs / tiles that keeps whole code chunks
// and at least four k blocks per CTA.
static int tc_ksplit(const TcPlan& pl, int64_t tiles, int64_t k) {
  static const int force = [] { const char* e = getenv("PQ_TC_KSPLIT"); return e ? atoi(e) : 0; }();
  const int kb = (int)(k / kTcBK);
  int best = 1;
  for (int ks = 2; ks <= kTcMaxSplit; ++ks) {
    if (kb % ks || (kb / ks) % pl.ch || kb / ks < 4) continue;
    if (tiles * ks <= sm_count() && (!force || ks <= force)) best = ks;   // one tile per CTA: the splits wait for each other
  }
  return force == 1 ? 1 : best;
}

// Only decode-sized batches split K (a prefill has tiles to spare): tickets + kTcMaxSplit partial sums.
// 16-bit codes at prefill: the 65536-entry books live in L2 and every codeword is gathered once per weight
// group AND token tile (8 times at 2048 tokens) - the kernel is L2-gather-bound at a third of its 8-bit
// rate.  With a workspace for the dense weight, pq_b200_dequant writes W once (one gather per weight
// group, the bits the in-kernel dequant would produce) and this kernel TMA-loads it as its A operand:
// what the reference does with aqlm_dequant + cuBLAS (aqlm.py:132-156), minus the library GEMM.
static bool tc_dense_prefill(int64_t t, int nbits) {
  static const int off = [] { const char* e = getenv("PQ_TC_DENSE"); return e && atoi(e) == 0; }();
  return !off && nbits > 8 && t >= 1024;
}

int64_t gemm_tc_workspace(int64_t t, int64_t n, int64_t k, int, int nbits, int, int) {
  if (tc_dense_prefill(t, nbits)) return kTcTicketBytes + n * k * 2;    // W behind the ticket area, which stays zero
  return t <= 256 ? kTcTicketBytes + (int64_t)kTcMaxSplit * t * n * 4 : 0;
}

int dequant_dispatch(const void* codes, const void* codebooks, const void* scales, void* out, int64_t n, int64_t k,
                     int d, int nbits, int m, const int64_t* part_sizes, int n_parts, int dtype, cudaStream_t stream);

// dense [N, K] weight as a 2-D tensor map: box = 64 k x 128 rows, 128-byte swizzle (the A operand's layout)
static int make_w_tmap(CUtensorMap* map, const void* w, int64_t n, int64_t k, int dtype) {
  PFN_tmapEncodeTiled enc = tmap_encoder();
  if (!enc) return PQ_ERR_UNSUPPORTED;
  const cuuint64_t gdim[2] = {(cuuint64_t)k, (cuuint64_t)n};
  const cuuint64_t gstride[1] = {(cuuint64_t)k * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kTcBK, (cuuint32_t)kTcBM};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, dtype == PQ_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                         const_cast<void*>(w), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PQ_OK : PQ_ERR_ALIGN;
}

template <typename T, int D, typename CodeT, int MT>
static int launch_tc(const CUtensorMap& map, const CUtensorMap& wmap, const TcArgs& a, const TcPlan& pl, cudaStream_t stream) {
  auto kern = pq_gemm_tc_kernel<T, D, CodeT, MT>;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr_set[dev & 63] = true;
  }
  const int tiles = a.n_tok_tiles * a.n_row_tiles * a.ksplit;
  const int grid = tiles < sm_count() ? tiles : sm_count();
  kern<<<grid, kTcThreads, pl.smem_total, stream>>>(map, wmap, a);
  return (int)cudaGetLastError();
}

template <typename T, typename CodeT>
static int launch_tc_d(const CUtensorMap& map, const CUtensorMap& wmap, const TcArgs& a, const TcPlan& pl, int d,
                       cudaStream_t stream) {
#define PQ_TC_MT(DD)                                                          \
  switch (pl.mt) {                                                            \
    case 1: return launch_tc<T, DD, CodeT, 1>(map, wmap, a, pl, stream);            \
    case 2: return launch_tc<T, DD, CodeT, 2>(map, wmap, a, pl, stream);            \
    case 4: return launch_tc<T, DD, CodeT, 4>(map, wmap, a, pl, stream);            \
    default: return launch_tc<T, DD, CodeT, 0>(map, wmap, a, pl, stream);           \
  }
  switch (d) {
    case 4: PQ_TC_MT(4)
    case 8: PQ_TC_MT(8)
    default: PQ_TC_MT(16)
  }
#undef PQ_TC_MT
}

int gemm_tc_dispatch(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                     const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                     int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                     int n_parts, int dtype, void* workspace, int64_t workspace_bytes, cudaStream_t stream) {
  if ((uintptr_t)x % 16 || (x_stride * 2) % 16 || (uintptr_t)codes % 16) return PQ_ERR_ALIGN;
  TcPlan pl;
  if (!tc_plan(pl, t, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
  TcArgs a{};
  a.codes = codes; a.codebooks = codebooks; a.scales = scales; a.bias = bias; a.y = y; a.y_stride = y_stride;
  a.T = (int)t; a.N = (int)n; a.K = (int)k; a.E = 1 << nbits; a.M = m;
  a.BN = pl.bn; a.n_tok_tiles = pl.n_tok_tiles; a.stages = pl.stages; a.groups = pl.groups; a.cb_smem_bytes = pl.cb_smem; a.cb_stride = pl.cb_stride;
  a.row_code_bytes = (int)(k / d * m * (nbits <= 8 ? 1 : 2));
  int rc = make_parts(a.parts, part_sizes, n_parts, n, kTcBM);
  if (rc) return rc;
  a.n_row_tiles = a.parts.cta_end[n_parts - 1];
#ifdef PQ_TC_TRACE
  a.trace = g_lut_trace;
#endif
  a.ksplit = 1;
  const int64_t tiles = (int64_t)a.n_tok_tiles * a.n_row_tiles;
  if (workspace && t <= 256 && tiles * 4 * 2 * 4 <= kTcTicketBytes) {
    const int ks = tc_ksplit(pl, tiles, k);
    if (ks > 1 && workspace_bytes >= kTcTicketBytes + (int64_t)ks * t * n * 4) {
      a.ksplit = ks;
      a.ws_tickets = reinterpret_cast<int*>(workspace);
      a.ws_part = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + kTcTicketBytes);
    }
  }
  CUtensorMap map, wmap;
  rc = make_x_tmap(&map, x, x_stride, t, k, pl.bn, dtype);
  if (rc) return rc;
  wmap = map;                                        // unused unless dense_w
  if (tc_dense_prefill(t, nbits) && workspace && workspace_bytes >= kTcTicketBytes + n * k * 2 &&
      ((uintptr_t)workspace & 127) == 0) {
    void* w_dense = reinterpret_cast<unsigned char*>(workspace) + kTcTicketBytes;   // the first 64 KiB are tickets
    rc = dequant_dispatch(codes, codebooks, nullptr, w_dense, n, k, d, nbits, m, part_sizes, n_parts, dtype, stream);
    if (rc) return rc;
    rc = make_w_tmap(&wmap, w_dense, n, k, dtype);
    if (rc) return rc;
    a.dense_w = 1;
    // no codebooks and no code ring in shared memory: deeper pipeline
    const int stage = kTcABytes + pl.bn * 128;
    int stages = (225 * 1024 - 1024) / stage;
    if (stages > kTcMaxStages) stages = kTcMaxStages;
    if (stages > k / kTcBK) stages = (int)(k / kTcBK);
    if (stages >= 2) { pl.stages = stages; pl.cb_smem = 0; pl.ring_bytes = 0; pl.smem_total = stages * stage + 1024; }
    a.stages = pl.stages; a.cb_smem_bytes = pl.cb_smem;
  }
  if (dtype == PQ_F16)
    return nbits <= 8 ? launch_tc_d<__half, uint8_t>(map, wmap, a, pl, d, stream)
                      : launch_tc_d<__half, uint16_t>(map, wmap, a, pl, d, stream);
  return nbits <= 8 ? launch_tc_d<__nv_bfloat16, uint8_t>(map, wmap, a, pl, d, stream)
                    : launch_tc_d<__nv_bfloat16, uint16_t>(map, wmap, a, pl, d, stream);
}

}  // namespace pq
