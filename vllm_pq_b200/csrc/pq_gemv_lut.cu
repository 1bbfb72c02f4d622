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
// ONE launch per linear.  CTA = (row block, 32-position tile, token tile):
//  1. the first code vectors are requested (128-bit no-allocate loads) BEFORE
//     griddepcontrol.wait: weights do not depend on the previous kernel, so under programmatic
//     dependent launch the HBM stream starts while the previous layer is still finishing;
//  2. the CTA builds its own table tile [E entries][32 positions] per token in shared memory.
//     That is a tiny GEMM (E x d) . (d x 32) and runs on the tensor cores (mma.sync m16n8k8 /
//     k16, fp32 accumulate): ~8 MMAs per warp instead of ~450 FMAs per thread, no table
//     kernel, no table traffic through L2;
//  3. lookups: a lane owns 16 consecutive positions of one row; table column = position mod 32
//     = shared-memory bank, and each lane walks its 16 bytes starting at byte (lane/2) so that
//     at every step the 32 lanes of a warp hit 32 distinct banks (conflict-free);
//  4. split-K: partial sums go to the workspace in a fixed layout; the last CTA of a row block
//     (self-resetting ticket counter) adds them in tile order - a deterministic sum, unlike
//     atomics - applies scale and bias in fp32 and writes y.
#include <type_traits>

#include "pq_common.cuh"

namespace pq {

constexpr int kLutThreads = 256;
constexpr int kLutWarps = kLutThreads / 32;
constexpr int kLutTile = 32;                       // positions per tile = banks
constexpr int kLutRowsPerIter = kLutWarps * 16;    // 128 rows per CTA iteration
constexpr int kLutInflight = 4;                    // code vectors in flight per lane
constexpr int kLutCounterBytes = 64 * 1024;        // zero-initialised, self-resetting tickets
constexpr int kLutMaxCounters = kLutCounterBytes / 4;

struct LutArgs {
  const void* x; const unsigned char* codes; const void* codebooks; const void* scales; const void* bias;
  void* y;
  float* partials; unsigned int* counters;
  long long x_stride, y_stride;
  int T, N, K, G, E, M;
  int Gp;                 // positions (= code bytes) per row
  int n_tiles;            // Gp / 32
  int t_pad;              // T rounded up to the token tile
  int rows_per_cta;
  Parts parts;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

template <typename T>
__device__ __forceinline__ void mma_k8(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
  if constexpr (std::is_same<T, __half>::value)
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
  else
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
}
template <typename T>
__device__ __forceinline__ void mma_k16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                        uint32_t b0, uint32_t b1) {
  if constexpr (std::is_same<T, __half>::value)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
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

// Table tile for one token: lut[e][q] = <cb[book(q)][e], x[group(q)]>, e < E, q < 32.
// MMA mapping (m16n8): rows = 16 entries, columns = 8 positions of ONE codebook.
// n-block j (0..3) uses book j % M; its column c = 2*tq + i is position 8*tq + slot(j, i) with
// slot(j, i) = j % M + M * (2 * (j / M) + i), so that after the four n-blocks lane (g, tq) holds
// positions 8*tq .. 8*tq + 7 of entries g and g + 8: two 128-bit stores per entry row.
template <typename T, int D, int M>
__device__ __forceinline__ void build_lut_tile(float* lut, const LutArgs& a, int part, int tile, int t_glob,
                                               int warp, int lane) {
  const int g = lane >> 2, tq = lane & 3;
  constexpr int EW = D / 2;                       // u32 words per codebook entry / x group
  const uint32_t* cb32 = reinterpret_cast<const uint32_t*>(
      reinterpret_cast<const T*>(a.codebooks) + (size_t)part * M * a.E * D);
  const uint32_t* x32 = reinterpret_cast<const uint32_t*>(
      reinterpret_cast<const T*>(a.x) + (size_t)t_glob * a.x_stride);
  const bool t_ok = t_glob < a.T;

  // B fragments: column n = g of n-block j is position 8*(g/2) + slot(j, g%2)
  uint32_t b0[4], b1[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int slot = (j % M) + M * (2 * (j / M) + (g & 1));
    const int q = tile * kLutTile + 8 * (g >> 1) + slot;
    const int grp = q / M;
    b0[j] = 0u; b1[j] = 0u;
    if (t_ok) {
      if constexpr (D == 4) { if (tq < 2) b0[j] = __ldg(x32 + (size_t)grp * EW + tq); }
      else if constexpr (D == 8) { b0[j] = __ldg(x32 + (size_t)grp * EW + tq); }
      else { b0[j] = __ldg(x32 + (size_t)grp * EW + tq); b1[j] = __ldg(x32 + (size_t)grp * EW + tq + 4); }
    }
  }
  // Each warp owns entry blocks {warp, warp + 8} (+16, ... for E > 256): all fragment loads of a
  // pair are issued before the first MMA so the pair costs one L2 round trip, not several.
  for (int ebase = warp; ebase * 16 < a.E; ebase += 2 * kLutWarps) {
    const bool two = (ebase + kLutWarps) * 16 < a.E;
    uint32_t fa[2][M][4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int e0 = (ebase + h * kLutWarps) * 16 + g;
#pragma unroll
      for (int m = 0; m < M; ++m) {
        const uint32_t* ca = cb32 + ((size_t)m * a.E + e0) * EW;
        const uint32_t* cbp = ca + 8 * EW;
        const bool ok = (h == 0 || two) && (D != 4 || tq < 2);
        fa[h][m][0] = ok ? __ldg(ca + tq) : 0u;
        fa[h][m][1] = ok ? __ldg(cbp + tq) : 0u;
        if constexpr (D == 16) {
          fa[h][m][2] = ok ? __ldg(ca + tq + 4) : 0u;
          fa[h][m][3] = ok ? __ldg(cbp + tq + 4) : 0u;
        }
      }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (h == 1 && !two) break;
      const int e0 = (ebase + h * kLutWarps) * 16 + g, e1 = e0 + 8;
      float out0[8], out1[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int m = j % M;
        float c[4] = {0.f, 0.f, 0.f, 0.f};
        if constexpr (D == 16) mma_k16<T>(c, fa[h][m][0], fa[h][m][1], fa[h][m][2], fa[h][m][3], b0[j], b1[j]);
        else mma_k8<T>(c, fa[h][m][0], fa[h][m][1], b0[j]);
        const int s0 = (j % M) + M * (2 * (j / M)), s1 = s0 + M;
        out0[s0] = c[0]; out0[s1] = c[1]; out1[s0] = c[2]; out1[s1] = c[3];
      }
      float4* r0 = reinterpret_cast<float4*>(lut + (size_t)e0 * kLutTile + 8 * tq);
      float4* r1 = reinterpret_cast<float4*>(lut + (size_t)e1 * kLutTile + 8 * tq);
      r0[0] = make_float4(out0[0], out0[1], out0[2], out0[3]);
      r0[1] = make_float4(out0[4], out0[5], out0[6], out0[7]);
      r1[0] = make_float4(out1[0], out1[1], out1[2], out1[3]);
      r1[1] = make_float4(out1[4], out1[5], out1[6], out1[7]);
    }
  }
}

template <typename T, int D, int TT>
__global__ void __launch_bounds__(kLutThreads, TT == 1 ? 3 : 2)
pq_gemv_lut_kernel(const LutArgs a) {
  extern __shared__ __align__(128) unsigned char lut_smem[];
  __shared__ int is_last;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tile = blockIdx.y, z = blockIdx.z;
  const RowBlock rb = locate_row_block(a.parts, blockIdx.x, a.rows_per_cta);

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

  pdl_wait();                                      // x comes from the previous kernel
  pdl_launch_dependents();
  const uint32_t t_stride = (uint32_t)a.E * kLutTile * 4u;
#pragma unroll
  for (int t = 0; t < TT; ++t) {
    float* lt = reinterpret_cast<float*>(lut_smem + (size_t)t * t_stride);
    if (a.M == 1) build_lut_tile<T, D, 1>(lt, a, rb.part, tile, z * TT + t, warp, lane);
    else if (a.M == 2) build_lut_tile<T, D, 2>(lt, a, rb.part, tile, z * TT + t, warp, lane);
    else build_lut_tile<T, D, 4>(lt, a, rb.part, tile, z * TT + t, warp, lane);
  }
  __syncthreads();

  // lane-constant column addresses: step i reads column 16*half + (i + r) % 16
  uint32_t col[16];
  const uint32_t lut_base = smem_u32(lut_smem);
#pragma unroll
  for (int i = 0; i < 16; ++i) col[i] = lut_base + (uint32_t)(16 * half + ((i + r) & 15)) * 4u;
  const uint32_t emask4 = ((uint32_t)a.E - 1u) * 0x01010101u;
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
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();                               // cumulative over the CTA's stores (ordered by the barrier)
    unsigned int* ctr = a.counters + (size_t)z * gridDim.x + blockIdx.x;
    const unsigned int prev = atomicAdd(ctr, 1u);
    is_last = (prev == (unsigned int)a.n_tiles - 1u);
    if (is_last) *ctr = 0u;                        // leave the ticket zero for the next launch
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

struct LutPlan {
  int tt, t_pad, n_tiles, rows_per_cta, n_row_blocks;
  size_t partial_bytes;
};

static LutPlan lut_plan(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts) {
  LutPlan pl{};
  pl.tt = lut_token_tile(t);
  pl.t_pad = (int)((t + pl.tt - 1) / pl.tt * pl.tt);
  pl.n_tiles = (int)(k / d * m / kLutTile);
  // rows per CTA: the whole grid should be resident at once (one wave - a partial second wave
  // costs a full CTA latency on these microsecond kernels) unless that needs > 2048 rows per CTA
  const int sms = sm_count();
  const size_t smem = (size_t)pl.tt * (1 << nbits) * kLutTile * 4;
  int resident = (int)((200 * 1024) / (smem + 1024));
  const int reg_limit = pl.tt == 1 ? 3 : 2;          // __launch_bounds__ of pq_gemv_lut_kernel
  if (resident > reg_limit) resident = reg_limit;
  if (resident < 1) resident = 1;
  const int64_t z_tiles = pl.t_pad / pl.tt;
  const int64_t want_ctas = (int64_t)sms * resident;
  int64_t rows = (n * pl.n_tiles * z_tiles + want_ctas - 1) / want_ctas;
  rows = (rows + kLutRowsPerIter - 1) / kLutRowsPerIter * kLutRowsPerIter;
  if (rows < kLutRowsPerIter) rows = kLutRowsPerIter;
  for (; rows < 16 * kLutRowsPerIter; rows += kLutRowsPerIter) {
    // partition-aligned row blocks: at most n_parts - 1 extra blocks
    const int64_t blocks = (n + rows - 1) / rows + (n_parts - 1);
    if (blocks * pl.n_tiles * z_tiles <= want_ctas) break;
  }
  pl.rows_per_cta = (int)rows;
  pl.n_row_blocks = (int)((n + rows - 1) / rows) + n_parts;      // upper bound (partition-aligned blocks)
  pl.partial_bytes = (size_t)pl.n_tiles * pl.t_pad * n * 4;
  return pl;
}

bool gemv_lut_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (nbits > 8 || nbits < 4 || t > 8 || t < 1) return false;       // E >= 16: whole 16-entry MMA blocks
  if (d != 4 && d != 8 && d != 16) return false;
  if (m != 1 && m != 2 && m != 4) return false;
  if (k % d) return false;
  const int64_t gp = k / d * m;
  if (gp % kLutTile) return false;                  // whole 32-position tiles, 16-byte aligned rows
  const int tt = lut_token_tile(t);
  if ((int64_t)tt * (1 << nbits) * kLutTile * 4 > 200 * 1024) return false;
  const LutPlan pl = lut_plan(t, n, k, d, nbits, m, PQ_MAX_PARTS);
  if ((int64_t)pl.n_row_blocks * (pl.t_pad / pl.tt) > kLutMaxCounters) return false;
  return true;
}

// PQ_ALGO_AUTO heuristic: tiny problems are better served by the direct kernel (no split-K pass).
bool gemv_lut_preferred(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  return gemv_lut_supported(t, n, k, d, nbits, m) && n * (k / d * m) >= (1 << 20);
}

// Workspace = [64 KiB ticket counters | split-K partials].  The counters must be zero before the
// first call; every call leaves them zero (the python side keeps one zero-initialised buffer).
int64_t gemv_lut_workspace(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts) {
  if (!gemv_lut_supported(t, n, k, d, nbits, m)) return 0;
  const LutPlan pl = lut_plan(t, n, k, d, nbits, m, n_parts);
  return (int64_t)(kLutCounterBytes + pl.partial_bytes);
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

template <typename T, int D, int TT>
static int launch_lut(LutArgs& a, cudaStream_t stream) {
  const size_t smem = (size_t)TT * a.E * kLutTile * 4;
  auto kern = pq_gemv_lut_kernel<T, D, TT>;
  static bool attr_set[64] = {};                   // per instantiation, per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (smem > 48 * 1024 && !attr_set[dev & 63]) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set[dev & 63] = true;
  }
  dim3 grid(a.parts.cta_end[a.parts.n_parts - 1], a.n_tiles, a.t_pad / TT);
  return launch_pdl(kern, grid, dim3(kLutThreads), smem, stream, a);
}

template <typename T, int D>
static int launch_lut_tt(LutArgs& a, int tt, cudaStream_t stream) {
  switch (tt) {
    case 1: return launch_lut<T, D, 1>(a, stream);
    case 2: return launch_lut<T, D, 2>(a, stream);
    default: return launch_lut<T, D, 4>(a, stream);
  }
}

template <typename T>
static int launch_lut_d(LutArgs& a, int d, int tt, cudaStream_t stream) {
  switch (d) {
    case 4: return launch_lut_tt<T, 4>(a, tt, stream);
    case 8: return launch_lut_tt<T, 8>(a, tt, stream);
    default: return launch_lut_tt<T, 16>(a, tt, stream);
  }
}

int gemv_lut_dispatch(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                      const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                      int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                      int n_parts, int dtype, void* workspace, int64_t workspace_bytes,
                      cudaStream_t stream) {
  if ((uintptr_t)codes % 16 || (uintptr_t)workspace % 256 || (uintptr_t)x % 4 || (x_stride * 2) % 4)
    return PQ_ERR_ALIGN;
  const LutPlan pl = lut_plan(t, n, k, d, nbits, m, n_parts);
  LutArgs a{};
  a.x = x; a.codes = (const unsigned char*)codes; a.codebooks = codebooks; a.scales = scales; a.bias = bias;
  a.y = y; a.x_stride = x_stride; a.y_stride = y_stride;
  a.T = (int)t; a.N = (int)n; a.K = (int)k; a.G = (int)(k / d); a.E = 1 << nbits; a.M = m;
  a.Gp = a.G * m; a.n_tiles = pl.n_tiles; a.t_pad = pl.t_pad; a.rows_per_cta = pl.rows_per_cta;
  int rc = make_parts(a.parts, part_sizes, n_parts, n, pl.rows_per_cta);
  if (rc) return rc;
  const int row_blocks = a.parts.cta_end[n_parts - 1];
  if ((int64_t)row_blocks * (pl.t_pad / pl.tt) > kLutMaxCounters) return PQ_ERR_UNSUPPORTED;
  if ((int64_t)(kLutCounterBytes + pl.partial_bytes) > workspace_bytes) return PQ_ERR_WORKSPACE;
  unsigned char* ws = (unsigned char*)workspace;
  a.counters = (unsigned int*)ws;
  a.partials = (float*)(ws + kLutCounterBytes);
  return dtype == PQ_F16 ? launch_lut_d<__half>(a, d, pl.tt, stream)
                         : launch_lut_d<__nv_bfloat16>(a, d, pl.tt, stream);
}

}  // namespace pq
