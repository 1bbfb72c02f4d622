// Decode-path PQ GEMV, lookup-table form (nbits <= 8): the bandwidth path.
//
//   y[t, n] = scale[n] * sum_q LUT_t[q][code[n, q]] + bias[n]
//   LUT_t[q][e] = <codebook[p*M + q%M][e][:], x[t, (q/M)*d : (q/M+1)*d]>        q = g*M + m
//
// A "position" q is one code slot of a row (G*M per row), so additive codebooks (M > 1)
// cost nothing extra in the consumer.  Per code the consumer does ONE 4-byte shared-memory
// lookup and ONE add, against 16 B of codebook + 8 FMAs in the direct form - at 1-2 bits per
// weight that is the difference between shared-memory-bound and HBM-bound.
// There is no reference kernel of this shape; semantics are aqlm.py:77-127.
//
// Two launches per linear:
//  1. pq_lut_build_kernel  builds the tables once per token into an L2-resident workspace in
//     the consumer's tile layout [E][32 positions] (entry-major), and zeroes the split-K
//     counters.  P*T*G*M*E*d FMAs: ~1 M FMAs for a 4096-wide layer.
//  2. pq_gemv_lut_kernel   CTA = (row block, 32-position tile, token tile).  The tile's tables
//     (E*128 B per token) arrive by one cp.async.bulk (TMA bulk copy, mbarrier complete_tx);
//     codes stream with 128-bit no-allocate loads issued BEFORE griddepcontrol.wait - they do
//     not depend on the previous kernel.  A lane owns 16 consecutive positions of one row;
//     table column = position mod 32 = shared-memory bank, and each lane walks its 16 bytes
//     starting at byte (lane/2) so that at every step the 32 lanes of a warp hit 32 distinct
//     banks (conflict-free lookups).  Split-K partial sums go to the workspace in a fixed
//     layout; the last CTA of a row block (atomic ticket) adds them in tile order - a
//     deterministic sum, unlike atomics - applies scale and bias in fp32 and writes y.
#include "pq_common.cuh"

namespace pq {

constexpr int kLutThreads = 256;
constexpr int kLutWarps = kLutThreads / 32;
constexpr int kLutTile = 32;                       // positions per tile = banks
constexpr int kLutRowsPerIter = kLutWarps * 16;    // 128 rows per CTA iteration
constexpr int kLutInflight = 4;                    // code vectors in flight per lane

struct LutArgs {
  const void* x; const unsigned char* codes; const void* codebooks; const void* scales; const void* bias;
  void* y;
  float* lut; float* partials; unsigned int* counters;
  long long x_stride, y_stride;
  int T, N, K, G, E, M, d;
  int Gp;                 // positions (= code bytes) per row
  int n_tiles;            // Gp / 32
  int t_pad;              // T rounded up to the token tile
  int rows_per_cta;
  int n_counters;
  int P;
  Parts parts;
};

// ------------------------------------------------------------------ table build
template <typename T, int D, int TT>
__global__ void __launch_bounds__(kLutThreads)
pq_lut_build_kernel(const LutArgs a) {
  const int tile = blockIdx.x, p = blockIdx.y, t = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q = tile * kLutTile + lane;
  const int g = q / a.M, m = q - g * a.M;
  pdl_wait();                                      // x comes from the previous kernel
  pdl_launch_dependents();
  if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0)
    for (int i = threadIdx.x; i < a.n_counters; i += kLutThreads) a.counters[i] = 0u;

  float xg[D];
  {
    const T* xp = reinterpret_cast<const T*>(a.x) + (size_t)t * a.x_stride + (size_t)g * D;
#pragma unroll
    for (int j = 0; j < D; ++j) xg[j] = Traits<T>::to_f(xp[j]);
  }
  const uint32_t* cb = reinterpret_cast<const uint32_t*>(
      reinterpret_cast<const T*>(a.codebooks) + ((size_t)(p * a.M + m) * a.E) * D);
  float* out = a.lut + ((((size_t)(t / TT) * a.P + p) * a.n_tiles + tile) * TT + (t % TT)) * a.E * kLutTile;
  for (int e = warp; e < a.E; e += kLutWarps) {
    const uint32_t* ent = cb + (size_t)e * (D / 2);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < D / 2; ++j) {
      const float2 c = Traits<T>::to_f2(__ldg(ent + j));
      s = fmaf(c.x, xg[2 * j], s);
      s = fmaf(c.y, xg[2 * j + 1], s);
    }
    out[(size_t)e * kLutTile + lane] = s;          // 128 B per warp store
  }
}

// ------------------------------------------------------------------ consumer
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// rotate the 16 bytes of v left by r bytes: out.byte[i] = v.byte[(i + r) % 16]
__device__ __forceinline__ void rotate_bytes(const uint4& v, int r, uint32_t emask4, uint32_t (&u)[4]) {
  uint32_t w0 = v.x, w1 = v.y, w2 = v.z, w3 = v.w;
  if (r & 4) { uint32_t t = w0; w0 = w1; w1 = w2; w2 = w3; w3 = t; }
  if (r & 8) { uint32_t t0 = w0, t1 = w1; w0 = w2; w1 = w3; w2 = t0; w3 = t1; }
  const int sh = (r & 3) * 8;
  u[0] = __funnelshift_r(w0, w1, sh) & emask4;    // mask all four codes of a word at once
  u[1] = __funnelshift_r(w1, w2, sh) & emask4;
  u[2] = __funnelshift_r(w2, w3, sh) & emask4;
  u[3] = __funnelshift_r(w3, w0, sh) & emask4;
}

template <typename T, int TT>
__global__ void __launch_bounds__(kLutThreads, 2)
pq_gemv_lut_kernel(const LutArgs a) {
  extern __shared__ __align__(128) unsigned char lut_smem[];
  __shared__ __align__(8) unsigned long long bar;
  __shared__ int is_last;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tile = blockIdx.y, z = blockIdx.z;
  const RowBlock rb = locate_row_block(a.parts, blockIdx.x, a.rows_per_cta);
  const uint32_t lut_bytes = (uint32_t)TT * a.E * kLutTile * 4;

  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }

  const int r = lane >> 1, half = lane & 1;
  const int n_iter = (rb.row1 - rb.row0 + kLutRowsPerIter - 1) / kLutRowsPerIter;
  const unsigned char* cbase = a.codes + (size_t)tile * kLutTile + half * 16;
  const int row_in_iter = warp * 16 + r;

  uint4 cv[kLutInflight];
  auto load_batch = [&](int it0) {
#pragma unroll
    for (int i = 0; i < kLutInflight; ++i) {
      const int row = rb.row0 + (it0 + i) * kLutRowsPerIter + row_in_iter;
      cv[i] = (it0 + i < n_iter && row < rb.row1)
                  ? ld_stream_16(cbase + (size_t)row * a.Gp) : make_uint4(0, 0, 0, 0);
    }
  };
  load_batch(0);                                   // weights: independent of the previous kernel

  pdl_wait();                                      // tables come from pq_lut_build_kernel
  pdl_launch_dependents();
  __syncthreads();                                 // barrier init visible to all threads
  if (threadIdx.x == 0) {
    const float* src = a.lut + (((size_t)z * a.P + rb.part) * a.n_tiles + tile) * TT * a.E * kLutTile;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 :: "r"(smem_u32(&bar)), "r"(lut_bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(lut_smem)), "l"(src), "r"(lut_bytes), "r"(smem_u32(&bar)) : "memory");
  }
  {
    uint32_t done = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\t"
                   "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                   "selp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    }
  }

  // lane-constant column addresses: step i reads column 16*half + (i + r) % 16
  uint32_t col[16];
  const uint32_t lut_base = smem_u32(lut_smem);
#pragma unroll
  for (int i = 0; i < 16; ++i) col[i] = lut_base + (uint32_t)(16 * half + ((i + r) & 15)) * 4u;
  const uint32_t emask4 = ((uint32_t)a.E - 1u) * 0x01010101u;
  const uint32_t t_stride = (uint32_t)a.E * kLutTile * 4u;
  float* part = a.partials + ((size_t)tile * a.t_pad + (size_t)z * TT) * a.N;

  for (int it0 = 0; it0 < n_iter; it0 += kLutInflight) {
    uint4 cur[kLutInflight];
#pragma unroll
    for (int i = 0; i < kLutInflight; ++i) cur[i] = cv[i];
    if (it0 + kLutInflight < n_iter) load_batch(it0 + kLutInflight);
#pragma unroll
    for (int i = 0; i < kLutInflight; ++i) {
      if (it0 + i >= n_iter) break;
      uint32_t u[4];
      rotate_bytes(cur[i], r, emask4, u);
      float acc[TT];
#pragma unroll
      for (int t = 0; t < TT; ++t) acc[t] = 0.f;
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const uint32_t code = __byte_perm(u[k >> 2], 0u, 0x4440u + (k & 3));   // one PRMT
        const uint32_t addr = col[k] + code * 128u;                               // one IMAD
#pragma unroll
        for (int t = 0; t < TT; ++t) {
          float v;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr + t * t_stride));
          acc[t] += v;
        }
      }
      const int row = rb.row0 + (it0 + i) * kLutRowsPerIter + row_in_iter;
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], 1);       // the row's other 16 positions
        if (half == 0 && row < rb.row1) part[(size_t)t * a.N + row] = acc[t];
      }
    }
  }

  // ---- split-K epilogue: the last CTA of this (row block, token tile) reduces ----
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(a.counters + (size_t)z * gridDim.x + blockIdx.x, 1u);
    is_last = (prev == (unsigned int)a.n_tiles - 1u);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int rows = rb.row1 - rb.row0;
  for (int i = threadIdx.x; i < rows * TT; i += kLutThreads) {
    const int t = i / rows, row = rb.row0 + (i - t * rows);
    const int tg = z * TT + t;
    if (tg >= a.T) continue;
    const float* pp = a.partials + (size_t)tg * a.N + row;
    float s = 0.f;
    for (int k = 0; k < a.n_tiles; ++k) s += __ldcg(pp + (size_t)k * a.t_pad * a.N);   // fixed order
    const float sc = Traits<T>::to_f(reinterpret_cast<const T*>(a.scales)[row]);
    const float b = a.bias ? Traits<T>::to_f(reinterpret_cast<const T*>(a.bias)[row]) : 0.f;
    reinterpret_cast<T*>(a.y)[(size_t)tg * a.y_stride + row] = Traits<T>::from_f(fmaf(s, sc, b));
  }
}

// ------------------------------------------------------------------ host side
extern bool g_use_pdl;

static int lut_token_tile(int64_t t) { return t >= 4 ? 4 : (t >= 2 ? 2 : 1); }

bool gemv_lut_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (nbits > 8 || t > 8) return false;
  if (d != 4 && d != 8 && d != 16) return false;
  if (k % d) return false;
  const int64_t gp = k / d * m;
  if (gp % kLutTile) return false;                  // whole 32-position tiles, 16-byte aligned rows
  const int tt = lut_token_tile(t);
  if ((int64_t)tt * (1 << nbits) * kLutTile * 4 > 200 * 1024) return false;
  (void)n;
  return true;
}

// PQ_ALGO_AUTO heuristic: tiny problems are better served by one direct-kernel launch than by
// two launches plus a split-K reduction.
bool gemv_lut_preferred(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  return gemv_lut_supported(t, n, k, d, nbits, m) && n * (k / d * m) >= (1 << 20);
}

struct LutPlan {
  int tt, t_pad, n_tiles, rows_per_cta, n_row_blocks, n_counters;
  size_t lut_bytes, partial_bytes, counter_bytes;
};

static LutPlan lut_plan(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts) {
  LutPlan pl{};
  pl.tt = lut_token_tile(t);
  pl.t_pad = (int)((t + pl.tt - 1) / pl.tt * pl.tt);
  pl.n_tiles = (int)(k / d * m / kLutTile);
  // rows per CTA: large enough to amortise the tile's table load (E*128 B per token against
  // 32 B of codes per row), small enough that the grid covers the machine twice
  const int sms = sm_count();
  int64_t want_ctas = (int64_t)sms * 2;
  int64_t rows = (n * pl.n_tiles + want_ctas - 1) / want_ctas;
  rows = (rows + kLutRowsPerIter - 1) / kLutRowsPerIter * kLutRowsPerIter;
  if (rows < 4 * kLutRowsPerIter) rows = 4 * kLutRowsPerIter;
  if (rows > 16 * kLutRowsPerIter) rows = 16 * kLutRowsPerIter;
  pl.rows_per_cta = (int)rows;
  pl.n_row_blocks = (int)((n + rows - 1) / rows) + n_parts;      // upper bound (partition-aligned blocks)
  pl.n_counters = pl.n_row_blocks * (pl.t_pad / pl.tt);
  pl.lut_bytes = (size_t)pl.t_pad * n_parts * pl.n_tiles * (1 << nbits) * kLutTile * 4;
  pl.partial_bytes = (size_t)pl.n_tiles * pl.t_pad * n * 4;
  pl.counter_bytes = ((size_t)pl.n_counters * 4 + 255) / 256 * 256;
  return pl;
}

int64_t gemv_lut_workspace(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts) {
  if (!gemv_lut_supported(t, n, k, d, nbits, m)) return 0;
  const LutPlan pl = lut_plan(t, n, k, d, nbits, m, n_parts);
  return (int64_t)(pl.lut_bytes + pl.partial_bytes + pl.counter_bytes + 512);
}

template <typename Kern>
static int launch_pdl(Kern kern, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, const LutArgs& a) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 1 : 0;
  return (int)cudaLaunchKernelEx(&cfg, kern, a);
}

template <typename T, int TT>
static int launch_lut(LutArgs& a, cudaStream_t stream) {
  dim3 bgrid(a.n_tiles, a.P, a.T);
  int rc;
  switch (a.d) {
    case 4:  rc = launch_pdl(pq_lut_build_kernel<T, 4, TT>, bgrid, dim3(kLutThreads), 0, stream, a); break;
    case 8:  rc = launch_pdl(pq_lut_build_kernel<T, 8, TT>, bgrid, dim3(kLutThreads), 0, stream, a); break;
    default: rc = launch_pdl(pq_lut_build_kernel<T, 16, TT>, bgrid, dim3(kLutThreads), 0, stream, a); break;
  }
  if (rc) return rc;
  const size_t smem = (size_t)TT * a.E * kLutTile * 4;
  auto kern = pq_gemv_lut_kernel<T, TT>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  dim3 grid(a.parts.cta_end[a.parts.n_parts - 1], a.n_tiles, a.t_pad / TT);
  return launch_pdl(kern, grid, dim3(kLutThreads), smem, stream, a);
}

int gemv_lut_dispatch(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                      const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                      int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                      int n_parts, int dtype, void* workspace, int64_t workspace_bytes,
                      cudaStream_t stream) {
  if ((uintptr_t)codes % 16 || (uintptr_t)workspace % 256) return PQ_ERR_ALIGN;
  const LutPlan pl = lut_plan(t, n, k, d, nbits, m, n_parts);
  LutArgs a{};
  a.x = x; a.codes = (const unsigned char*)codes; a.codebooks = codebooks; a.scales = scales; a.bias = bias;
  a.y = y; a.x_stride = x_stride; a.y_stride = y_stride;
  a.T = (int)t; a.N = (int)n; a.K = (int)k; a.G = (int)(k / d); a.E = 1 << nbits; a.M = m; a.d = d;
  a.Gp = a.G * m; a.n_tiles = pl.n_tiles; a.t_pad = pl.t_pad; a.rows_per_cta = pl.rows_per_cta;
  a.P = n_parts;
  int rc = make_parts(a.parts, part_sizes, n_parts, n, pl.rows_per_cta);
  if (rc) return rc;
  const int row_blocks = a.parts.cta_end[n_parts - 1];
  if (row_blocks > pl.n_row_blocks) return PQ_ERR_WORKSPACE;
  a.n_counters = pl.n_counters;
  unsigned char* ws = (unsigned char*)workspace;
  a.counters = (unsigned int*)ws;
  a.lut = (float*)(ws + pl.counter_bytes);
  a.partials = (float*)(ws + pl.counter_bytes + pl.lut_bytes);
  if ((int64_t)(pl.counter_bytes + pl.lut_bytes + pl.partial_bytes) > workspace_bytes) return PQ_ERR_WORKSPACE;

#define PQ_LUT_TT(TYPE)                                                    \
  switch (pl.tt) {                                                         \
    case 1: return launch_lut<TYPE, 1>(a, stream);                         \
    case 2: return launch_lut<TYPE, 2>(a, stream);                         \
    default: return launch_lut<TYPE, 4>(a, stream);                        \
  }
  if (dtype == PQ_F16) { PQ_LUT_TT(__half) }
  PQ_LUT_TT(__nv_bfloat16)
#undef PQ_LUT_TT
}

}  // namespace pq
