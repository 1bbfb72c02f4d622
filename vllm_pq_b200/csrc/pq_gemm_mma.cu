// Small-batch PQ GEMM (1 <= T <= 64 tokens, 8-bit codes, d = 8): dequantise straight into
// mma.sync register fragments.
//
//   y[t, n] = scale[n] * sum_k x[t, k] * W[n, k] + bias[n],   W = codebook gather of `codes`
//
// Replaces Code2x8MatVec + the per-token host loop of aqlm_gemm and, for T > 6,
// aqlm_dequant + F.linear (reference csrc/quantization/aqlm/gemm_kernels.cu:102-183,467-496 and
// vllm/model_executor/layers/quantization/aqlm.py:132-156).  Where it sits between the two
// other decode kernels of this library:
//   * the table GEMV (pq_gemv_lut.cu) pays one 4-byte lookup per code AND per token, plus a
//     table build per K-slice: unbeatable per byte at T = 1 on large layers, linear in T;
//   * the tcgen05 GEMM (pq_gemm_tc.cu) writes every dequantised weight to shared memory
//     (16 B written + 16 B gathered per code): right when T is large enough to amortise it;
//   * here a codeword is gathered ONCE (one conflict-free 128-bit shared-memory load) and IS
//     the tensor-core A fragment - no shared-memory write of W, no table, cost independent of T
//     up to 8 tokens per MMA column block.  The ceiling is the codeword gather: 16 B of shared
//     memory per code byte -> 8 codes/clk/SM -> about 2.3 TB/s of int8 codes (35 % of HBM).
//
// The contraction index of an MMA can be permuted freely as long as A and B agree.  Lane
// (g = lane / 4, tq = lane % 4) owns, per 64-code-byte "superblock" of a row, the 16 code bytes
// [16 tq, 16 tq + 16): ONE 128-bit global load per row, the four lanes of a quad cover 64
// contiguous bytes.  Step j of the superblock multiplies the four groups {16 tq + j} (d = 8 ->
// k = 32): the gathered codeword of (row g, group 16 tq + j) supplies a0/a2 of two m16n8k16
// MMAs, the one of row g + 8 supplies a1/a3, and b0/b1 are the same 8 x-values of token g, read
// with one 128-bit load from a padded shared-memory copy of x.
//
// Work split: a CTA owns a block of rows (never straddling an output partition) and, inside a
// cluster of CS CTAs, a K-slice; its 8 warps take the slice's superblocks round-robin and sweep
// all row tiles of the block, RT tiles at a time.  Partial sums meet in shared memory (warps),
// then in the owner CTA's shared memory over DSMEM (cluster), always in a fixed order.
// Before griddepcontrol.wait: codebooks -> shared memory (each 16-byte entry replicated to a
// 128-byte line so that the 8 lanes of a load phase hit 8 different bank groups) and the first
// code vectors -> registers; x is only touched after the wait (programmatic dependent launch).
#include <cstdlib>
#include <type_traits>

#include "pq_common.cuh"

namespace pq {

constexpr int kMmThreads = 256;
constexpr int kMmWarps = kMmThreads / 32;
constexpr int kMmMaxCluster = 8;

struct MmArgs {
  const void* x; const unsigned char* codes; const void* codebooks; const void* scales; const void* bias;
  void* y;
  long long x_stride, y_stride;
  int T, N, K, E;
  int G;                  // groups per row
  int Gp;                 // code bytes per row = G * M
  int n_sb;               // superblocks (64 code bytes) per row
  int S;                  // superblocks per CTA (K-slice)
  int CS;                 // cluster size
  int WK;                 // warps along K inside the CTA (power of two); 8 / WK warps split the row chunks
  int PH;                 // superblocks of x staged per phase (<= S)
  int TE;                 // token columns kept in the reduction buffers = min(T, 8 NB)
  int rows_per_cluster;   // multiple of 16 * RT
  int x_rows;             // token rows staged (min(T_tile, 8 NB) + zero row if needed)
  int xs_stride;          // 16-byte chunks per staged x row, == 4 (mod 8)
  int smem_cb;            // bytes of replicated codebooks
  Parts parts;            // cta_end counts clusters
};

__device__ __forceinline__ uint32_t mm_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// plain (non-volatile) loads: ptxas may hoist the next step's gathers above this step's MMAs
__device__ __forceinline__ uint4 mm_lds128(const unsigned char* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ uint32_t mm_cluster_ctarank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ void mm_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <typename T>
__device__ __forceinline__ void mm_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                       uint32_t b0, uint32_t b1) {
  if constexpr (std::is_same<T, __half>::value)
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  else
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// sum of two codewords, rounded once (two addends: identical to fp32 accumulate + one rounding
// up to double-rounding corner cases that only the dequant kernel has to honour bit for bit)
template <typename T>
__device__ __forceinline__ uint32_t mm_add2(uint32_t a, uint32_t b) {
  using T2 = typename Traits<T>::T2;
  T2 r = __hadd2(*reinterpret_cast<const T2*>(&a), *reinterpret_cast<const T2*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// Table offset of one gather in ONE byte-permute: byte 1 = the code, byte 0 = the lane's slot in the
// entry line (pre-doubled for 128-byte lines, where the offset is then halved), bytes 2-3 = 0.
__device__ __forceinline__ uint32_t mm_code_off(const uint4& v, uint32_t lane_sel, int i) {   // i: constant after unrolling
  const uint32_t w = (i >> 2) == 0 ? v.x : (i >> 2) == 1 ? v.y : (i >> 2) == 2 ? v.z : v.w;
  return __byte_perm(w, lane_sel, 0x5504u | ((uint32_t)(i & 3) << 4));
}

template <typename T, int NB, int RT, int M>
__global__ void __launch_bounds__(kMmThreads, 2)
pq_gemm_mma_kernel(const MmArgs a) {
  constexpr int GPV = 16 / M;                      // groups per 16-byte code vector
  constexpr int GPS = 64 / M;                      // groups per superblock
  constexpr int TT = 8 * NB;                       // tokens per CTA
  constexpr int LINE = 128 * M;                    // bytes per codebook entry line: M books x 8 copies x 16 B
  extern __shared__ __align__(256) unsigned char mm_smem[];
  unsigned char* cbs = mm_smem;                                              // [E][M][8][16 B]
  unsigned char* zline = cbs + a.smem_cb;                                    // 256 zero bytes
  uint4* xs = reinterpret_cast<uint4*>(zline + 256);                         // [x_rows][xs_stride]
  float* red = reinterpret_cast<float*>(xs + (size_t)a.x_rows * a.xs_stride);  // [WK][TE][rows]
  float* recv = red + (size_t)a.WK * a.TE * a.rows_per_cluster;              // [CS][TE][chunk]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, tq = lane & 3;
  const int rank = a.CS > 1 ? (int)mm_cluster_ctarank() : 0;
  const int tok0 = blockIdx.y * TT;
  const RowBlock rb = locate_row_block(a.parts, blockIdx.x / a.CS, a.rows_per_cluster);
  const int rows = rb.row1 - rb.row0;
  const int n_chunks = (rows + 16 * RT - 1) / (16 * RT);     // RT-tile chunks in this block
  // warp (wk, wr): superblocks wk, wk + WK, ... of the CTA's K-slice x row chunks wr, wr + WR, ...
  const int WK = a.WK, WR = kMmWarps / WK;
  const int wk = warp & (WK - 1), wr = warp / WK;
  const int sb_begin = rank * a.S;
  const int sb_end = min(a.n_sb, sb_begin + a.S);
  const int my_ch = wr < n_chunks ? (n_chunks - wr + WR - 1) / WR : 0;
  const int n_ph = (max(sb_end - sb_begin, 0) + a.PH - 1) / a.PH;   // x is staged PH superblocks at a time

  // ---- weights first (independent of the previous kernel) ----
  {
    const uint4* cb_g = reinterpret_cast<const uint4*>(a.codebooks) + (size_t)rb.part * M * a.E;
    uint4* cb_s = reinterpret_cast<uint4*>(cbs);
    for (int i = tid; i < M * a.E; i += kMmThreads) {
      const int book = i / a.E, e = i - book * a.E;
      const uint4 v = __ldg(cb_g + i);
#pragma unroll
      for (int c = 0; c < 8; ++c) cb_s[(e * M + book) * 8 + ((c + lane) & 7)] = v;   // rotated: conflict-free
    }
    if (tid < 16) reinterpret_cast<uint4*>(zline)[tid] = make_uint4(0, 0, 0, 0);
  }
  const uint32_t emask4 = ((uint32_t)a.E - 1u) * 0x01010101u;
  // phase ph covers superblocks [p0, p1); this warp's are p0 + wk, p0 + wk + WK, ...
  // items of a phase in order: chunk-major (chunk = wr + WR * ci), then superblock (p0 + wk + WK * si)
  int p0 = sb_begin, p1 = min(sb_end, p0 + a.PH);
  int my_sb = p0 + wk < p1 ? (p1 - p0 - wk + WK - 1) / WK : 0;
  int n_items = my_ch * my_sb;
  int pci = 0, psi = 0;                              // (chunk, superblock) index of the item being fetched
  auto load_item = [&](uint4 (&cv)[RT][2], bool& ok) {
    ok = false;
    int byte0 = 0;
    if (pci < my_ch && my_sb > 0) {
      byte0 = (p0 + wk + WK * psi) * 64 + tq * 16;
      ok = byte0 < a.Gp;
    }
    const int ch = wr + WR * pci;
#pragma unroll
    for (int rt = 0; rt < RT; ++rt)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int rl = min(ch * 16 * RT + rt * 16 + h * 8 + g, rows - 1);
        cv[rt][h] = ok ? ld_stream_16(a.codes + (size_t)(rb.row0 + rl) * a.Gp + byte0) : make_uint4(0, 0, 0, 0);
      }
    if (++psi >= my_sb) { psi = 0; ++pci; }
  };
  uint4 cv[RT][2], nx[RT][2];
  bool ok = false, nok = false;
  load_item(nx, nok);

  pdl_launch_dependents();
  pdl_wait();

  // lanes of token rows that are not staged read the zero row (row x_rows - 1)
  const unsigned char* xrow[NB];
#pragma unroll
  for (int nb = 0; nb < NB; ++nb) {
    const int r = min(nb * 8 + g, a.x_rows - 1);
    xrow[nb] = reinterpret_cast<const unsigned char*>(xs) + (size_t)(r * a.xs_stride) * 16u;
  }
  // byte 0 of every gather offset: this lane's 16-byte copy (doubled when the offset gets halved)
  const uint32_t lane_sel = (uint32_t)(lane & 7) * (M == 1 ? 32u : 16u);

  float acc[RT][NB][4];
  for (int ph = 0; ph < n_ph; ++ph) {
    if (ph > 0) {
      __syncthreads();                               // everyone is done with the previous x slice
      p0 = sb_begin + ph * a.PH; p1 = min(sb_end, p0 + a.PH);
      my_sb = p0 + wk < p1 ? (p1 - p0 - wk + WK - 1) / WK : 0;
      n_items = my_ch * my_sb;
      pci = 0; psi = 0;
      load_item(nx, nok);
    }
    // ---- x of this phase -> shared memory, one pad chunk per code vector ----
    {
      const int ngl = (p1 - p0) * GPS;               // groups in the phase
      const T* xg = reinterpret_cast<const T*>(a.x);
      for (int idx = tid; idx < a.x_rows * ngl; idx += kMmThreads) {
        const int r = idx / ngl, lg = idx - r * ngl;
        const int gg = p0 * GPS + lg, tok = tok0 + r;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (gg < a.G && tok < a.T && r < a.TE)
          v = __ldg(reinterpret_cast<const uint4*>(xg + (size_t)tok * a.x_stride + (size_t)gg * 8));
        xs[r * a.xs_stride + lg + lg / GPV] = v;
      }
    }
    __syncthreads();

    int ci = 0, si = 0;
    for (int it = 0; it < n_items; ++it) {
      const int ch = wr + WR * ci;
#pragma unroll
      for (int rt = 0; rt < RT; ++rt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {                // mask once per word: codes may be stored sign-extended
          cv[rt][h].x = nx[rt][h].x & emask4; cv[rt][h].y = nx[rt][h].y & emask4;
          cv[rt][h].z = nx[rt][h].z & emask4; cv[rt][h].w = nx[rt][h].w & emask4;
        }
      ok = nok;
      load_item(nx, nok);                            // next item's codes are in flight during this one
      if (si == 0) {
#pragma unroll
        for (int rt = 0; rt < RT; ++rt)
#pragma unroll
          for (int nb = 0; nb < NB; ++nb)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[rt][nb][q] = 0.f;
      }
      // a code vector beyond K has zero codes and gathers from the zero line: no predication below
      const unsigned char* cb_base = ok ? cbs : zline;
      // local group of step 0: (wk + WK si) * GPS + tq * GPV; its padded chunk adds one per code vector
      const int lg0 = (wk + WK * si) * GPS + tq * GPV;
      const uint32_t xoff = (uint32_t)(lg0 + lg0 / GPV) * 16u;
#pragma unroll
      for (int j = 0; j < GPV; ++j) {
        uint4 xb[NB];
#pragma unroll
        for (int nb = 0; nb < NB; ++nb) xb[nb] = mm_lds128(xrow[nb] + xoff + j * 16);
#pragma unroll
        for (int rt = 0; rt < RT; ++rt) {
          uint4 w[2];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if constexpr (M == 1) {
              w[h] = mm_lds128(cb_base + (mm_code_off(cv[rt][h], lane_sel, j) >> 1));
            } else {
              w[h] = mm_lds128(cb_base + mm_code_off(cv[rt][h], lane_sel, 2 * j));
              const uint4 w1 = mm_lds128(cb_base + 128 + mm_code_off(cv[rt][h], lane_sel, 2 * j + 1));
              w[h].x = mm_add2<T>(w[h].x, w1.x); w[h].y = mm_add2<T>(w[h].y, w1.y);
              w[h].z = mm_add2<T>(w[h].z, w1.z); w[h].w = mm_add2<T>(w[h].w, w1.w);
            }
          }
#pragma unroll
          for (int nb = 0; nb < NB; ++nb) {
            mm_mma<T>(acc[rt][nb], w[0].x, w[1].x, w[0].y, w[1].y, xb[nb].x, xb[nb].y);
            mm_mma<T>(acc[rt][nb], w[0].z, w[1].z, w[0].w, w[1].w, xb[nb].z, xb[nb].w);
          }
        }
      }
      if (si == my_sb - 1) {                        // chunk done for this phase: K partial -> shared memory
#pragma unroll
        for (int rt = 0; rt < RT; ++rt)
#pragma unroll
          for (int nb = 0; nb < NB; ++nb)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const int rl = ch * 16 * RT + rt * 16 + (q >> 1) * 8 + g;
              const int t = nb * 8 + 2 * tq + (q & 1);
              if (rl < rows && t < a.TE) {
                float* dst = red + ((size_t)wk * a.TE + t) * a.rows_per_cluster + rl;
                *dst = ph == 0 ? acc[rt][nb][q] : *dst + acc[rt][nb][q];
              }
            }
      }
      if (++si >= my_sb) { si = 0; ++ci; }
    }
  }
  __syncthreads();

  // ---- reduce over the warps that had work, then over the cluster (rank order), scale, bias ----
  const int n_wk = min(WK, max(sb_end - sb_begin, 0));
  const int TE = a.TE;
  const int chunk = (rows + a.CS - 1) / a.CS;
  if (a.CS > 1) {
    const uint32_t recv_addr = mm_smem_u32(recv);
    for (int i = tid; i < rows * TE; i += kMmThreads) {
      const int t = i / rows, rl = i - t * rows;
      float s = 0.f;
      for (int w = 0; w < n_wk; ++w) s += red[((size_t)w * TE + t) * a.rows_per_cluster + rl];
      const int o = rl / chunk, jrow = rl - o * chunk;
      const uint32_t la = recv_addr + (uint32_t)(((rank * TE + t) * chunk) + jrow) * 4u;
      uint32_t ra;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(o));
      asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(ra), "f"(s) : "memory");
    }
    mm_cluster_sync();
  }
  const int fin0 = rank * chunk, fin1 = min(rows, fin0 + chunk);
  const int nfin = max(fin1 - fin0, 0);
  const T* sc_g = reinterpret_cast<const T*>(a.scales);
  const T* bi_g = reinterpret_cast<const T*>(a.bias);
  for (int i = tid; i < nfin * TE; i += kMmThreads) {
    const int t = i / nfin, jrow = i - t * nfin;
    const int tok = tok0 + t;
    if (tok >= a.T) break;                          // t grows with i
    float s = 0.f;
    if (a.CS > 1) {
      for (int r = 0; r < a.CS; ++r) s += recv[((r * TE + t) * chunk) + jrow];
    } else {
      for (int w = 0; w < n_wk; ++w) s += red[((size_t)w * TE + t) * a.rows_per_cluster + jrow];
    }
    const int row = rb.row0 + fin0 + jrow;
    const float sc = Traits<T>::to_f(sc_g[row]);
    const float b = bi_g ? Traits<T>::to_f(bi_g[row]) : 0.f;
    reinterpret_cast<T*>(a.y)[(size_t)tok * a.y_stride + row] = Traits<T>::from_f(fmaf(s, sc, b));
  }
}

// ------------------------------------------------------------------ host side
extern bool g_use_pdl;
int g_mm_last_plan[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // nb, rt, cs, wk, ph, rows per cluster, grid.x, smem

struct MmPlan {
  int nb, rt, cs, wk, ph, S, rows_per_cluster, z_tiles, x_rows, te, xs_stride, smem_cb, smem_total;
};

static int mm_env(const char* name) {
  const char* e = getenv(name);
  return e ? atoi(e) : 0;
}

// Pick the cluster size (K split over CTAs), the warp split inside a CTA (WK warps along K,
// 8 / WK along row chunks), the row tiles per warp step and the rows per cluster with a small
// time model: per SM the codeword gathers run at about 8 codes/clk; every CTA pays a fixed cost,
// its codebook + x staging and one or two reduction steps.
static bool mm_plan(MmPlan& pl, int64_t t, int64_t n, int64_t k, int nbits, int m, int n_parts) {
  const int E = 1 << nbits;
  const int64_t gp = k / 8 * m;
  const int n_sb = (int)((gp + 63) / 64);
  const int gpv = 16 / m, gps = 64 / m;
  // token columns per MMA step: 8 * nb.  Two additive codebooks keep nb <= 2 (16 tokens per z tile): at
  // nb = 4 the second gather + add per fragment no longer fits the register budget and the kernel
  // measured 6 x slower per token (o_proj 2x8: 11.8 us at 16 tokens, 75 us at 32; profiles/r2_22)
  pl.nb = t <= 8 ? 1 : ((t <= 16 || m == 2) ? 2 : 4);
  const int tt = 8 * pl.nb;
  pl.z_tiles = (int)((t + tt - 1) / tt);
  const int t_tile = (int)(t < tt ? t : tt);
  pl.te = t_tile;
  pl.x_rows = t_tile < tt ? t_tile + 1 : tt;            // + zero row for the idle token lanes
  pl.smem_cb = m * E * 128 + 256;                    // replicated codebooks + the zero line
  const int sms = sm_count();
  static const int force_cs = mm_env("PQ_MMA_CS"), force_rt = mm_env("PQ_MMA_RT"),
                   force_rows = mm_env("PQ_MMA_ROWS"), force_wk = mm_env("PQ_MMA_WK");
  const int smem_cap = 112 * 1024;
  double best = 1e30;
  bool found = false;
  for (int cs = 1; cs <= kMmMaxCluster; cs *= 2) {
    if (cs > 1 && cs > n_sb) break;
    if (force_cs && cs != force_cs) continue;
    const int S = (n_sb + cs - 1) / cs;
    for (int wk = 8; wk >= 1; wk >>= 1) {
      if (wk > S && wk > 1) continue;
      if (force_wk && wk != force_wk) continue;
      const int wr = kMmWarps / wk;
      // x slice staged per phase: the whole K-slice if it fits, else the largest multiple of wk
      static const int ph_opts[] = {1 << 20, 48, 32, 24, 16, 8, 4, 2, 1};
      for (int pi = 0; pi < 9; ++pi) {
        const int ph = ph_opts[pi] > S ? S : ph_opts[pi];
        if (pi > 0 && ph_opts[pi] >= S) continue;        // same as the first option
        if (ph < wk || ph % wk) continue;
        int chunks = ph * gps + ph * gps / gpv;          // groups + pads
        while ((chunks & 7) != 4) ++chunks;
        const int64_t smem_x = (int64_t)pl.x_rows * chunks * 16;
        if (pl.smem_cb + smem_x > smem_cap - 8192) continue;
        const int n_ph = (S + ph - 1) / ph;
        for (int rt = 4; rt >= 1; rt >>= 1) {
          if (force_rt && rt != force_rt) continue;
          if (rt > (m == 1 ? (pl.nb == 1 ? 4 : pl.nb == 2 ? 2 : 1) : (pl.nb == 1 ? 2 : 1))) continue;   // register budget (128)
          const int gran = 16 * rt * wr;                 // every warp gets a row chunk
          for (int per_sm = 2; per_sm >= 1; --per_sm) {
            const int64_t slots = (int64_t)per_sm * sms / cs / pl.z_tiles;
            int64_t rows = (n + (slots > 0 ? slots : 1) - 1) / (slots > 0 ? slots : 1);
            if (force_rows) rows = force_rows;
            rows = (rows + gran - 1) / gran * gran;
            auto smem_of = [&](int64_t r) {
              return (int64_t)pl.smem_cb + smem_x + (int64_t)wk * pl.te * r * 4 +
                     (cs > 1 ? (int64_t)cs * pl.te * ((r + cs - 1) / cs) * 4 : 0);
            };
            while (rows > gran && smem_of(rows) > smem_cap) rows -= gran;   // reduction buffers must fit
            const int64_t smem = smem_of(rows);
            if (smem > smem_cap) continue;
            const int64_t blocks = (n + rows - 1) / rows + (n_parts - 1);
            const int64_t ctas = blocks * cs * pl.z_tiles;
            const int64_t waves = (ctas + (int64_t)2 * sms - 1) / ((int64_t)2 * sms);
            const double share = ctas > sms ? 1.4 : 1.0;  // CTAs sharing an SM slow each other down
            const double codes_cta = (double)rows * S * 64.0;
            const double xover = 1.0 + (double)pl.nb / (2.0 * rt);
            const double mma = pl.nb >= 2 ? 0.5 * pl.nb : 1.0;           // tensor pipe beyond 8 tokens
            const double gather = codes_cta / 8.0 / 1965.0 * share * (xover > mma ? xover : mma);
            const double stage = (pl.smem_cb * 1.0 + smem_x * 2.0 * n_ph) / 128.0 / 1965.0 * share;
            const double cta = 2.5 + stage + gather + 0.3 * (n_ph - 1) + (cs > 1 ? 0.8 : 0.1);
            const double cost = (double)waves * cta;
            if (cost < best) {
              best = cost; found = true;
              pl.rt = rt; pl.cs = cs; pl.wk = wk; pl.S = S; pl.ph = ph; pl.rows_per_cluster = (int)rows;
              pl.xs_stride = chunks; pl.smem_total = (int)smem;
            }
          }
        }
      }
    }
  }
  return found;
}

bool gemm_mma_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (d != 8 || nbits > 8 || nbits < 1 || (m != 1 && m != 2) || t < 1 || t > 64) return false;
  if (k % 8 || (k / 8 * m) % 16) return false;
  MmPlan pl;
  return mm_plan(pl, t, n, k, nbits, m, PQ_MAX_PARTS);
}

template <typename T, int NB, int RT, int M>
static int mm_launch(MmArgs& a, const MmPlan& pl, cudaStream_t stream) {
  auto kern = pq_gemm_mma_kernel<T, NB, RT, M>;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr_set[dev & 63] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(a.parts.cta_end[a.parts.n_parts - 1] * pl.cs, pl.z_tiles, 1);
  cfg.blockDim = dim3(kMmThreads);
  cfg.dynamicSmemBytes = pl.smem_total;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = pl.cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 2 : 1;
  g_mm_last_plan[0] = pl.nb; g_mm_last_plan[1] = pl.rt; g_mm_last_plan[2] = pl.cs; g_mm_last_plan[3] = pl.wk;
  g_mm_last_plan[4] = pl.ph; g_mm_last_plan[5] = pl.rows_per_cluster; g_mm_last_plan[6] = cfg.gridDim.x;
  g_mm_last_plan[7] = pl.smem_total;
  return (int)cudaLaunchKernelEx(&cfg, kern, a);
}

template <typename T, int NB, int M>
static int mm_launch_rt(MmArgs& a, const MmPlan& pl, cudaStream_t stream) {
  if constexpr (NB <= 2) { if (pl.rt == 4) return mm_launch<T, NB, 4, M>(a, pl, stream); }
  if (pl.rt >= 2) return mm_launch<T, NB, 2, M>(a, pl, stream);
  return mm_launch<T, NB, 1, M>(a, pl, stream);
}

template <typename T, int M>
static int mm_launch_nb(MmArgs& a, const MmPlan& pl, cudaStream_t stream) {
  switch (pl.nb) {
    case 1: return mm_launch_rt<T, 1, M>(a, pl, stream);
    case 2: return mm_launch_rt<T, 2, M>(a, pl, stream);
    default: return mm_launch_rt<T, 4, M>(a, pl, stream);
  }
}

int gemm_mma_dispatch(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                      const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                      int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                      int n_parts, int dtype, cudaStream_t stream) {
  if (d != 8) return PQ_ERR_UNSUPPORTED;
  if ((uintptr_t)codes % 16 || (uintptr_t)x % 16 || (x_stride * 2) % 16 || (uintptr_t)codebooks % 16) return PQ_ERR_ALIGN;
  MmPlan pl;
  if (!mm_plan(pl, t, n, k, nbits, m, n_parts)) return PQ_ERR_UNSUPPORTED;
  MmArgs a{};
  a.x = x; a.codes = (const unsigned char*)codes; a.codebooks = codebooks; a.scales = scales; a.bias = bias;
  a.y = y; a.x_stride = x_stride; a.y_stride = y_stride;
  a.T = (int)t; a.N = (int)n; a.K = (int)k; a.E = 1 << nbits;
  a.G = (int)(k / 8); a.Gp = a.G * m; a.n_sb = (a.Gp + 63) / 64;
  a.S = pl.S; a.CS = pl.cs; a.WK = pl.wk; a.PH = pl.ph; a.TE = pl.te; a.rows_per_cluster = pl.rows_per_cluster;
  a.x_rows = pl.x_rows; a.xs_stride = pl.xs_stride; a.smem_cb = pl.smem_cb - 256;
  int rc = make_parts(a.parts, part_sizes, n_parts, n, pl.rows_per_cluster);
  if (rc) return rc;
  if (dtype == PQ_F16) return m == 1 ? mm_launch_nb<__half, 1>(a, pl, stream) : mm_launch_nb<__half, 2>(a, pl, stream);
  return m == 1 ? mm_launch_nb<__nv_bfloat16, 1>(a, pl, stream) : mm_launch_nb<__nv_bfloat16, 2>(a, pl, stream);
}

}  // namespace pq
