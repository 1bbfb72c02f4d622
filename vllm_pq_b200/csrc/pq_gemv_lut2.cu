// Decode-path PQ GEMV, lookup-table form, second generation (nbits <= 8): TMA-fed, one row per lane.
//
//   y[t, n] = scale[n] * sum_q LUT_t[q][code[n, q]] + bias[n]          q = g*M + m  ("position")
//   LUT_t[q][e] = <codebook[p*M + q%M][e][:], x[t, (q/M)*d : (q/M+1)*d]>
//
// Semantics are aqlm.py:77-127 (dequantize + F.linear); there is no reference kernel of this shape.
// What changed against pq_gemv_lut.cu (kept for T-pairs and as a cross-check):
//  * codes arrive by TMA (cp.async.bulk.tensor.2d, boxes of 64 rows x 64 code bytes) issued by a
//    producer warp into a 4-stage ring, with a rolling cp.async.bulk.prefetch.tensor into L2
//    several row iterations ahead.  No LSU issue slots and no per-thread address arithmetic are
//    spent on the stream, and the producer never waits for x: under programmatic dependent launch
//    it runs ahead of griddepcontrol.wait.
//  * ONE LANE OWNS ONE ROW: no shuffles in the main loop, the per-row partial sum is a register.
//  * the bank-conflict-free walk needs no register rotation: per code PRMT + LDS + FADD and nothing
//    else (the old kernel spent ~2 more instructions per code on SEL/SHF/LOP rotations).
//    Lane l (bits l4..l0) visits the 64 positions of its row in its own order; the order is built
//    from three free things: the ADDRESS it loads its 16-byte code chunks from (chunk order and
//    8-byte halves swapped by lane bits), the PRMT SELECTOR (a per-lane register: which byte of
//    a code word goes first) and the TABLE COLUMN PERMUTATION below.
//
// Table layout: lut[e][col], 64 fp32 columns = 256 B per entry, so (code << 8 | column byte) is
// ONE PRMT.  Position pos = 16 c + 4 j + b  (c: 16-byte chunk, j: word in chunk, b: byte in word)
// lives in column  col = 32 (j & 1) + 16 (j >> 1) + 4 c + b,  bank = col % 32.
// At step (v, j, b) of the walk lane l reads
//      chunk c = (v + l[2:1]) & 3,  word j' = j ^ 2 l[3],  byte b' = (b + 2 l[4] + l[0]) & 3
// so bank = 16 ((j>>1) ^ l[3]) + 4 ((v + l[2:1]) & 3) + ((b + 2 l[4] + l[0]) & 3): a bijection of
// the 32 lanes onto the 32 banks at every step; the 128 (j & 1) part is an immediate offset.
// The code words themselves are read with two LDS.64 per chunk from a dense [256 rows][64 B]
// stage: address = row * 64 + 16 c + 8 (l[3] or !l[3]) - 16 distinct 8-byte bank pairs per
// half-warp, conflict-free too.
//
// Split-K: a cluster of CS CTAs owns a block of rows; CTA `rank` owns position tiles rank, rank+CS..
// Partial row sums are pushed through distributed shared memory and added in rank order
// (deterministic); scale and bias are applied in fp32 before the single rounding of y.
#include <cuda.h>

#include <cstdlib>
#include <type_traits>

#include "pq_common.cuh"

namespace pq {

// Two CTA shapes (template parameter CW = consumer warps; one lane owns one row, so a row
// iteration covers CW * 32 rows):
//   CW = 16: one CTA per SM, 96 registers, 140-215 KB of shared memory - the throughput shape
//            (four warps per scheduler: the lookups are latency-bound);
//   CW = 8:  at most 112 registers and 113 KB, so TWO CTAs fit one SM: under programmatic
//            dependent launch the next layer's CTA is co-resident and runs its whole prologue
//            (codebooks, barriers, the first code stages, lane constants) while this layer
//            computes - the latency shape for the Llama-2-7B sized layers.
constexpr int kL2ProdWarps = 4;                     // the first issues the TMA (~70 cycles per issue); all four copy
                                                    // when the code pitch rules TMA out (COPY instantiation)
constexpr int kL2Tile = 64;                         // positions per table = code bytes per row per box
constexpr int kL2BoxRows = 64;
constexpr int kL2BoxBytes = kL2BoxRows * kL2Tile;   // 4 KiB
constexpr int kL2MaxStages = 6;
constexpr int kL2MaxCluster = 16;                   // 16 needs cudaFuncAttributeNonPortableClusterSizeAllowed
constexpr int kL2MaxT = 8;

struct Lut2Args {
  const void* x; const void* codebooks; const void* scales; const void* bias; void* y;
  long long x_stride, y_stride;
  int T, N, K, E, M;
  int Gp;                 // positions (= code bytes) per row
  const unsigned char* codes;
  int copy_w;             // 0: TMA.  4 / 8: the code pitch is not a multiple of 16 bytes (TP shards of
                          // down_proj) - the producer warp copies with cp.async of this width instead
  int n_tiles;            // ceil(Gp / 64)
  int S;                  // tiles per CTA = ceil(n_tiles / CS)
  int CS;                 // cluster size (a power of two)
  int cs_log2, m_log2;    // the prologue is on every layer's critical path: shifts, not divisions
  int rows_per_cluster;   // multiple of 64
  int KG;                 // clusters along K per row block (each owns CS * S consecutive tiles)
  float* ws_part;         // KG > 1: [T][KG][N] fp32 partial sums (workspace)
  int* ws_tickets;        // KG > 1: one counter per (token, row block, rank); zero between launches
  int stages;
  int prefetch;           // row iterations prefetched into L2 ahead of the ring
  int smem_cb, smem_x;    // bytes
  int mask;               // nbits < 8: codes are masked to nbits (aqlm.py:29-34 takes them mod 2^nbits)
  unsigned long long* trace;
  int xact;               // 1: x is [T, 2K] and the input is SiLU(x[:, :K]) * x[:, K:]
  int early_ld;           // griddepcontrol.launch_dependents at kernel entry (experiment knob)
  int trace_mask;         // which trace slots stamp (a %globaltimer read is slow: few stamps perturb less)
  // ---- fused epilogue of the row-parallel layers: all-reduce over NVLink peer memory,
  // residual add, sum of squares for the RMSNorm that follows ----
  int ep_flags;           // 1: all-reduce (ar_world > 1), 2: residual, 4: ssq - 0 on the plain path
  int ar_world, ar_rank;  // ar_world > 1: y = sum over ranks of this rank's scaled partial rows
  unsigned long long* ar_peer[PQ_AR_MAX_WORLD];   // every rank's receive buffer: [2][world][ar_slot] packets
  unsigned int* ar_state; // this rank's {epoch counter, CTA ticket}
  long long ar_slot;      // packets per (parity, source rank)
  const void* resid;      // [T, N] residual stream or NULL: y = T(T(sum) + resid)   (layernorm_kernels.cu:75-80)
  long long resid_stride;
  float* ssq;             // [T, gridDim.x] or NULL: every CTA's sum of y^2 over the rows it stored
  // ---- RMSNorm on the x load (xact == 2): x = T(T(z * rstd) * w), rstd from the producer's ssq ----
  const float* nrm_ssq;   // [T, nrm_slots]
  const void* nrm_w;      // [K]
  int nrm_slots;
  float nrm_eps;
  Parts parts;            // cta_end counts clusters
};

// One 8-byte packet = {fp32 partial row sum, epoch}: data and flag travel in ONE store, so a reader
// that sees the epoch sees the value (the LL protocol of NCCL) - no fence, no separate flag write.
__device__ __forceinline__ void l2_st_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long l2_ld_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// This rank's scaled partial sum of (tok, row) -> the sum over all ranks, identical bits on every rank
// (fp32, rank order 0..world-1: custom_all_reduce.cuh:204-212).  Replaces the store of y, the
// all-reduce launch and its staging copy (custom_all_reduce.cuh:214-237, custom_all_reduce.cu:75-77,
// linear.py:1209-1213).  Slot reuse: parity e & 1 was last used by launch e - 2, which every peer
// has completed before it could push launch e - 1 data, which this rank needed to finish e - 1.
__device__ __forceinline__ float l2_allreduce_row(const Lut2Args& a, int tok, int row, float v, uint32_t epoch) {
  const size_t idx = (size_t)tok * a.N + row;
  const size_t par = (size_t)(epoch & 1u) * a.ar_world;
  const unsigned long long pkt = ((unsigned long long)epoch << 32) | __float_as_uint(v);
#pragma unroll 1
  for (int p = 1; p < a.ar_world; ++p) {
    int peer = a.ar_rank + p;
    if (peer >= a.ar_world) peer -= a.ar_world;
    l2_st_sys_u64(a.ar_peer[peer] + (par + a.ar_rank) * a.ar_slot + idx, pkt);
  }
  const unsigned long long* mine = a.ar_peer[a.ar_rank] + par * a.ar_slot + idx;
  // all peers' packets are polled together: one L2 round trip per attempt, not one per peer
  unsigned long long got[PQ_AR_MAX_WORLD];
  bool all;
  do {
    all = true;
#pragma unroll
    for (int r = 0; r < PQ_AR_MAX_WORLD; ++r)
      if (r < a.ar_world && r != a.ar_rank) got[r] = l2_ld_sys_u64(mine + (size_t)r * a.ar_slot);
#pragma unroll
    for (int r = 0; r < PQ_AR_MAX_WORLD; ++r)
      if (r < a.ar_world && r != a.ar_rank) all = all && ((uint32_t)(got[r] >> 32) == epoch);
  } while (!all);
  float tot = 0.f;
#pragma unroll
  for (int r = 0; r < PQ_AR_MAX_WORLD; ++r)
    if (r < a.ar_world) tot += (r == a.ar_rank) ? v : __uint_as_float((uint32_t)got[r]);
  return tot;
}

constexpr int kL2TraceSlots = 16;
// Phase stamps (tools/trace_lut.py) are compiled into the TRACE instantiations only (fp16, d = 8, no
// mask): a dozen predicated stamp sequences on the latency-critical path are not free for 17 warps.
template <bool TRACE>
__device__ __forceinline__ void l2_trace(const Lut2Args& a, int slot) {
  if constexpr (!TRACE) return;
  if (a.trace != nullptr && threadIdx.x == 0 && ((a.trace_mask >> slot) & 1)) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[(((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * kL2TraceSlots + slot] = t;
  }
}

// BAR.SYNC does not block at issue: a stamp right after a barrier shows when this warp ARRIVED.
// This one waits for `dep` (a value read after the barrier from another warp's data) first.
template <bool TRACE>
__device__ __forceinline__ void l2_trace_dep(const Lut2Args& a, int slot, const volatile uint32_t* p) {
  if constexpr (!TRACE) return;
  if (a.trace != nullptr && threadIdx.x == 0 && ((a.trace_mask >> slot) & 1)) {
    const uint32_t dep = *p;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "r"(dep));
    a.trace[(((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * kL2TraceSlots + slot] = t;
  }
}

// ---- PTX wrappers ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t l2_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void l2_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count));
}
__device__ __forceinline__ void l2_mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void l2_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void l2_mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void l2_tma_load(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void l2_tma_prefetch(const CUtensorMap* map, int c0, int c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];"
               :: "l"(map), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void l2_cp_async_16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}
template <int CONS> __device__ __forceinline__ void l2_cons_sync() {      // the consumer warps only
  asm volatile("bar.sync 1, %0;" :: "n"(CONS) : "memory");
}
__device__ __forceinline__ uint32_t l2_cluster_rank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ void l2_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t l2_prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel)); return d;
}
template <int OFF> __device__ __forceinline__ float l2_lds(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(OFF));
  return v;
}
__device__ __forceinline__ void l2_lds64(uint32_t addr, uint32_t& a, uint32_t& b) {
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr));
}

template <typename T>
__device__ __forceinline__ void l2_mma_k16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
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

// x of one tile lives in shared memory in mma B-fragment order: word (hb, jn, lane) is what lane (g, tq)
// feeds to n-block jn of half hb (column n = g: pos = 16 (g >> 1) + 4 (hb + 2 (g & 1)) + jn, k = 2 tq,
// 2 tq + 1); D = 16 keeps the k + 8 words 256 further.  Conflict-free for every (D, M), and every
// warp of the builder reads the same 8 words.
template <int D> constexpr int l2_x_words() { return D == 16 ? 512 : 256; }
// Element offset (inside the CTA's first tile, tile index `tile`) of fragment word f, or -1 for the
// zero words of D = 4.  Computed before the dependency wait; after it each thread only adds the
// tile step, loads 4 bytes and stores them.
template <int D>
__device__ __forceinline__ int l2_x_offset(int f, int tile, int m_log2) {
  const int gpt = kL2Tile >> m_log2;               // x groups per tile
  const int half = f >> 8, r = f & 255, hb = r >> 7, jn = (r >> 5) & 3, ln = r & 31;
  const int g = ln >> 2, tq = ln & 3;
  const int pos = 16 * (g >> 1) + 4 * (hb + 2 * (g & 1)) + jn;
  if (D == 4 && tq >= 2) return -1;
  return (tile * gpt + (pos >> m_log2)) * D + 2 * (tq + 4 * half);
}

// Half a table (the 32 positions with j & 1 == hb of one 64-position tile) on the tensor cores:
// a tiny (E x D).(D x 32) GEMM.  cb: this partition's codebooks [M][E][D] in shared memory;
// xt: the tile's x in B-fragment order (l2_x_offset; zero beyond K) in shared memory.
// MMA mapping (m16n8): rows = 16 entries; n-block jn (0..3) is byte b = jn of every word, so all
// 8 columns of an n-block share codebook jn % M; its column cn is (chunk c = cn >> 1, word
// j = hb + 2 (cn & 1)):  pos = 16 (cn >> 1) + 4 (hb + 2 (cn & 1)) + jn.
// After the four n-blocks lane (g, tq) holds, for entries g and g + 8, the four bytes of word
// (c = tq, j = hb) and of word (c = tq, j = hb + 2): columns 32 hb + 4 tq + {0..3} and
// 32 hb + 16 + 4 tq + {0..3} - two 128-bit stores per entry row; odd g stores them in the other
// order so that the 8 lanes of a store phase cover all 32 banks.
template <typename T, int D, int M, int CW>
__device__ __forceinline__ void l2_build_half(float* lut, const T* cb, const T* xt, int E, int hb,
                                              int warp, int lane) {
  const int g = lane >> 2, tq = lane & 3;
  constexpr int EW = D / 2;                        // u32 words per codebook entry / x group
  const uint32_t* cb32 = reinterpret_cast<const uint32_t*>(cb);
  const uint32_t* x32 = reinterpret_cast<const uint32_t*>(xt);

  // B fragments: staged in fragment order (l2_x_offset), the same for every warp and entry block
  uint32_t b0[4], b1[4];
#pragma unroll
  for (int jn = 0; jn < 4; ++jn) {
    b0[jn] = x32[(hb * 4 + jn) * 32 + lane];
    b1[jn] = D == 16 ? x32[256 + (hb * 4 + jn) * 32 + lane] : 0u;
  }
  for (int eb = warp; eb * 16 < E; eb += CW) {
    const int e0 = eb * 16 + g, e1 = e0 + 8;
    uint32_t fa[M][4];
#pragma unroll
    for (int m = 0; m < M; ++m) {
      const uint32_t* ca = cb32 + (m * E + e0) * EW;
      const uint32_t* cbp = ca + 8 * EW;
      const bool ok = (D != 4 || tq < 2);
      fa[m][0] = ok ? ca[tq] : 0u;
      fa[m][1] = ok ? cbp[tq] : 0u;
      fa[m][2] = 0u; fa[m][3] = 0u;
      if constexpr (D == 16) { fa[m][2] = ca[tq + 4]; fa[m][3] = cbp[tq + 4]; }
    }
    float a0[4], bb0[4], a1[4], bb1[4];            // word (tq, hb) / (tq, hb + 2) of entries e0 / e1
#pragma unroll
    for (int jn = 0; jn < 4; ++jn) {
      const int m = jn % M;
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      if constexpr (D == 16) l2_mma_k16<T>(c, fa[m][0], fa[m][1], fa[m][2], fa[m][3], b0[jn], b1[jn]);
      // D <= 8: the k16 shape with a zero upper half - HMMA.16816 is the native shape on sm_100a,
      // HMMA.1688 measured ~0.2 us slower per table (profiles/r1_12_trace_lut2.txt)
      else l2_mma_k16<T>(c, fa[m][0], fa[m][1], 0u, 0u, b0[jn], 0u);
      a0[jn] = c[0]; bb0[jn] = c[1]; a1[jn] = c[2]; bb1[jn] = c[3];
    }
    float4* r0 = reinterpret_cast<float4*>(lut + e0 * kL2Tile + 32 * hb + 4 * tq);
    float4* r1 = reinterpret_cast<float4*>(lut + e1 * kL2Tile + 32 * hb + 4 * tq);
    if (g & 1) {
      r0[4] = make_float4(bb0[0], bb0[1], bb0[2], bb0[3]);
      r0[0] = make_float4(a0[0], a0[1], a0[2], a0[3]);
      r1[4] = make_float4(bb1[0], bb1[1], bb1[2], bb1[3]);
      r1[0] = make_float4(a1[0], a1[1], a1[2], a1[3]);
    } else {
      r0[0] = make_float4(a0[0], a0[1], a0[2], a0[3]);
      r0[4] = make_float4(bb0[0], bb0[1], bb0[2], bb0[3]);
      r1[0] = make_float4(a1[0], a1[1], a1[2], a1[3]);
      r1[4] = make_float4(bb1[0], bb1[1], bb1[2], bb1[3]);
    }
  }
}

template <typename T, int D, int CW>
__device__ __forceinline__ void l2_build_table(float* lut, const T* cb, const T* xt, int E, int M,
                                               int warp, int lane) {
#pragma unroll
  for (int hb = 0; hb < 2; ++hb) {
    if (M == 1) l2_build_half<T, D, 1, CW>(lut, cb, xt, E, hb, warp, lane);
    else if (M == 2) l2_build_half<T, D, 2, CW>(lut, cb, xt, E, hb, warp, lane);
    else l2_build_half<T, D, 4, CW>(lut, cb, xt, E, hb, warp, lane);
  }
}

// 16 codes (one chunk) of one row: PRMT + LDS + FADD each.  cw[j] holds word j ^ 2 l[3] of the chunk,
// colreg[jh] the four column bytes {0,4,8,12} + 16 c + 64 (jh ^ l[3]), sel[b] the lane's selectors.
// Four independent add chains (one per byte step) keep the FADD latency off the critical path.
template <bool MASK>
__device__ __forceinline__ void l2_lookup16(float (&acc)[4], uint32_t lut_base, const uint32_t (&cw)[4],
                                            const uint32_t (&colreg)[2], const uint32_t (&sel)[4], uint32_t emask4) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t w = MASK ? (cw[j] & emask4) : cw[j];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const uint32_t off = l2_prmt(colreg[j >> 1], w, sel[b]);
      const float v = (j & 1) ? l2_lds<128>(lut_base + off) : l2_lds<0>(lut_base + off);
      acc[b] += v;
    }
  }
}

// MASK: nbits < 8 - codes are masked to nbits (aqlm.py:29-34 takes them mod 2^nbits).
// COPY: the code pitch is not a multiple of 16 bytes - four producer warps copy with cp.async instead
// of one issuing TMA.  A separate instantiation on purpose: the lookup loop sits at the edge of its
// 96-register budget, and any extra live value in the kernel makes ptxas rematerialise the column
// constants inside it (measured: 10 % slower on the large layers, profiles/r1_28).
template <int CW, int COPY> constexpr int l2_threads() { return CW * 32 + 32 * (COPY ? kL2ProdWarps : 1); }
// FUSED: the instantiation that carries the fused prologue / epilogue (RMSNorm on the x load, all-reduce,
// residual add, sums of squares).  A template parameter, not run-time flags: with run-time flags
// alone the plain launch measured 0.3 us slower (profiles/r2_12_ab.txt) - set-up code of the fused
// steps hoisted onto the latency-critical path after the dependency wait and before the stores.
template <typename T, int D, bool MASK, bool COPY, int CW, bool FUSED, bool TRACE = false>
__global__ void __launch_bounds__(l2_threads<CW, COPY>(), CW <= 8 ? 2 : 1)
pq_gemv_lut2_kernel(const __grid_constant__ CUtensorMap tmap, const Lut2Args a) {
  constexpr int kL2ConsWarps = CW;
  constexpr int kL2Cons = CW * 32;                   // consumer threads = rows per iteration
  constexpr int kL2StageBoxes = kL2Cons / kL2BoxRows;
  constexpr int kL2StageBytes = kL2Cons * kL2Tile;   // CW = 16: 32 KiB = 8 boxes; CW = 8: 16 KiB
  extern __shared__ __align__(1024) unsigned char l2_smem_raw[];
  float* lut = reinterpret_cast<float*>(l2_smem_raw);
  unsigned char* ring = l2_smem_raw + (size_t)a.E * (kL2Tile * 4);
  T* cbs = reinterpret_cast<T*>(ring + (size_t)a.stages * kL2StageBytes);
  T* xs = reinterpret_cast<T*>(reinterpret_cast<unsigned char*>(cbs) + a.smem_cb);
  float* rowsum = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(xs) + a.smem_x);
  float* recv = rowsum + a.rows_per_cluster;       // [CS][chunk] partial sums pushed by the peers
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(recv + a.rows_per_cluster + kL2MaxCluster);
  const uint32_t full0 = l2_smem(bars), empty0 = full0 + 8 * kL2MaxStages;
  unsigned int* epoch_smem = reinterpret_cast<unsigned int*>(bars + 2 * kL2MaxStages) + 1;   // word 0: K-group ticket flag

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (a.early_ld) pdl_launch_dependents();
  const bool clustered = a.CS > 1;
  const int rank = clustered ? (int)l2_cluster_rank() : 0;
  // grid = (CS * KG, row blocks, tokens): no division to find this CTA's place
  const int tok = blockIdx.z;
  l2_trace<TRACE>(a, 0);
  // every CTA of the cluster has started once this barrier phase completes (waited for just before
  // the first distributed-shared-memory store, long after everyone arrived)
  if (clustered) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  const int block = blockIdx.y, kg = blockIdx.x >> a.cs_log2;   // cluster: (row block, K group)
  const int tile0 = kg * a.CS * a.S + rank;        // this CTA's tiles: tile0, tile0 + CS, ... (S of them)
  RowBlock rb;
  if (a.parts.n_parts == 1) {                      // o_proj, down_proj: no partition walk
    rb.part = 0; rb.row0 = block * a.rows_per_cluster; rb.row1 = min(rb.row0 + a.rows_per_cluster, a.N);
  } else {
    rb = locate_row_block(a.parts, block, a.rows_per_cluster);
  }
  const int rows = rb.row1 - rb.row0;
  const int n_iter = (rows + kL2Cons - 1) / kL2Cons;
  // tiles of this CTA that exist (the last K group may leave some ranks with none)
  const int my_tiles = tile0 < a.n_tiles ? min(a.S, (a.n_tiles - tile0 + a.CS - 1) >> a.cs_log2) : 0;
  const int total = my_tiles * n_iter;

  // the rows this CTA finishes: scale and bias are weights too - fetch them now (raw bits:
  // converting here would stall on the load)
  const int chunk = (rows + a.CS - 1) >> a.cs_log2;
  const int fin0 = rank * chunk, fin1 = min(rows, fin0 + chunk);
  unsigned short sc0_raw = 0, bi0_raw = 0;
  if (fin0 + tid < fin1) {
    sc0_raw = __ldg(reinterpret_cast<const unsigned short*>(a.scales) + rb.row0 + fin0 + tid);
    if (a.bias) bi0_raw = __ldg(reinterpret_cast<const unsigned short*>(a.bias) + rb.row0 + fin0 + tid);
  }

  // this partition's codebooks: requested first of all, ahead of the bulk code stream in the
  // memory queues (they gate the table build); consumer threads only - each waits for its own
  if (tid < kL2Cons) {
    const unsigned char* cb_g = reinterpret_cast<const unsigned char*>(a.codebooks) + (size_t)rb.part * a.smem_cb;
    const uint32_t dst = l2_smem(cbs);
    for (int i = tid; i < a.smem_cb / 16; i += kL2Cons) l2_cp_async_16(dst + i * 16, cb_g + (size_t)i * 16);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }

  if (tid < 2 * kL2MaxStages) {                    // one barrier per thread (all slots: no loop, no stage count)
    l2_mbar_init(full0 + 8 * tid, tid < kL2MaxStages ? (COPY ? 32 * kL2ProdWarps : 1) : kL2ConsWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  l2_trace<TRACE>(a, 8);

  if (warp >= kL2ConsWarps) {
    // ================= producer warp: the whole code stream of this CTA, weights only ===========
    pdl_launch_dependents();
    if constexpr (COPY) {
      // ---- code pitch not a multiple of 16 bytes: no TMA.  All 128 producer threads copy the item's
      // [512 rows][64 B] with cp.async (8 or 4 bytes; consecutive lanes = consecutive bytes of a row)
      // and arrive on the stage's barrier when their copies land.  Bytes beyond the row / the
      // matrix are not copied: their table columns are all zero (x is zero beyond K) and rows
      // beyond the block are never used.
      const uint32_t ring_addr = l2_smem(ring);
      const int w = a.copy_w, per_row = kL2Tile / w;
      int stage = 0, round = 0;
      for (int i = 0; i < total; ++i) {
        const int s = i / n_iter, it = i - s * n_iter;
        if (round > 0) l2_mbar_wait(empty0 + 8 * stage, (round - 1) & 1);
        const int col0 = (tile0 + s * a.CS) * kL2Tile;
        const int nrow = min(kL2Cons, rows - it * kL2Cons);
        const unsigned char* src = a.codes + (size_t)(rb.row0 + it * kL2Cons) * a.Gp + col0;
        const uint32_t dst = ring_addr + stage * kL2StageBytes;
        for (int idx = tid - kL2Cons; idx < nrow * per_row; idx += 32 * kL2ProdWarps) {
          const int r = idx / per_row, c = (idx - r * per_row) * w;
          if (col0 + c + w <= a.Gp) {
            if (w == 8)
              asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst + r * kL2Tile + c), "l"(src + (size_t)r * a.Gp + c) : "memory");
            else
              asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dst + r * kL2Tile + c), "l"(src + (size_t)r * a.Gp + c) : "memory");
          }
        }
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(full0 + 8 * stage) : "memory");
        if (++stage == a.stages) { stage = 0; ++round; }
      }
      asm volatile("cp.async.wait_all;" ::: "memory");
    } else if (lane == 0) {
      const uint32_t ring_addr = l2_smem(ring);
      auto boxes_of = [&](int it) {
        return min(kL2StageBoxes, (rows - it * kL2Cons + kL2BoxRows - 1) / kL2BoxRows);
      };
      auto box_c0 = [&](int s, int) { return (tile0 + s * a.CS) * kL2Tile; };
      auto box_r0 = [&](int it, int b) { return rb.row0 + it * kL2Cons + b * kL2BoxRows; };
      auto prefetch_item = [&](int i) {
        const int s = i / n_iter, it = i - s * n_iter;
        const int nb = boxes_of(it);
        for (int b = 0; b < nb; ++b) l2_tma_prefetch(&tmap, box_c0(s, b), box_r0(it, b));
      };
      int stage = 0, round = 0;
      for (int i = 0; i < total; ++i) {
        const int s = i / n_iter, it = i - s * n_iter;
        if (round > 0) l2_mbar_wait(empty0 + 8 * stage, (round - 1) & 1);
        const int nb = boxes_of(it);
        const uint32_t bar = full0 + 8 * stage;
        l2_mbar_expect_tx(bar, nb * kL2BoxBytes);
        for (int b = 0; b < nb; ++b)
          l2_tma_load(ring_addr + stage * kL2StageBytes + b * kL2BoxBytes, &tmap, box_c0(s, b), box_r0(it, b), bar);
        // L2 prefetch runs a.prefetch items ahead of the ring (the ring itself holds `stages`); the
        // first burst goes out only after the ring's own loads (a TMA issue costs ~70 cycles)
        if (i == min(total, a.stages) - 1)
          for (int j = a.stages; j < min(total, i + a.stages + a.prefetch); ++j) prefetch_item(j);
        else if (i >= a.stages && i + a.stages + a.prefetch - 1 < total) prefetch_item(i + a.stages + a.prefetch - 1);
        if (++stage == a.stages) { stage = 0; ++round; }
      }
    }
    __syncwarp();
    if (clustered) asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  } else {
    // ================= consumer warps ===========================================================
    // ---- lane constants of the conflict-free walk ----
    const uint32_t lh = (lane >> 3) & 1, lc = (lane >> 1) & 3, lr = ((lane >> 4) << 1) | (lane & 1);
    uint32_t sel[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const uint32_t bp = (b + lr) & 3;            // byte of the code word AND of colreg
      sel[b] = ((8u | bp) << 12) | ((8u | bp) << 8) | ((4u + bp) << 4) | bp;   // {0, 0, code, column}
    }
    uint32_t colv[4][2];                           // column bytes of walk step v, word pair jh
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const uint32_t cbase = (16u * ((v + lc) & 3)) * 0x01010101u + 0x0C080400u;
      colv[v][0] = cbase + (64u * lh) * 0x01010101u;
      colv[v][1] = cbase + (64u * (lh ^ 1)) * 0x01010101u;
    }
    // row (inside the iteration) of the ring slot this thread reads
    int perm = tid;
    uint32_t lut_base = l2_smem(lut);
    uint32_t emask4 = ((uint32_t)a.E - 1u) * 0x01010101u;
    uint32_t recv_addr = l2_smem(recv);
    uint32_t caddr[4];                             // first 8-byte half of walk step v in stage 0
#pragma unroll
    for (int v = 0; v < 4; ++v)
      caddr[v] = l2_smem(ring) + (uint32_t)tid * kL2Tile + (((v + lc) & 3) << 4) + 8 * lh;
    // x: this thread's word of the B-fragment layout
    constexpr int XW = l2_x_words<D>();
    constexpr int XF = XW > kL2Cons ? XW / kL2Cons : 1;      // fragment words per thread
    int xoff[XF];
#pragma unroll
    for (int f = 0; f < XF; ++f) xoff[f] = tid + f * kL2Cons < XW ? l2_x_offset<D>(tid + f * kL2Cons, tile0, a.m_log2) : -1;
    int xstep = a.CS * (kL2Tile >> a.m_log2) * D;  // elements between this CTA's consecutive tiles
    const T* xg = reinterpret_cast<const T*>(a.x) + (size_t)tok * a.x_stride;
    uint32_t* xs_mine = reinterpret_cast<uint32_t*>(xs) + tid;
    // all of the above is integer work on lane ids and launch constants: make sure it is done
    // BEFORE the dependency wait (16 warps share 4 schedulers - every instruction after the wait
    // is on the layer's critical path four times over)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      asm volatile("" : "+r"(sel[i]), "+r"(colv[i][0]), "+r"(colv[i][1]), "+r"(caddr[i]));
    }
#pragma unroll
    for (int f = 0; f < XF; ++f) asm volatile("" : "+r"(xoff[f]));
    asm volatile("" : "+r"(lut_base), "+r"(emask4), "+r"(recv_addr), "+r"(xstep), "+r"(perm), "+l"(xg), "+l"(xs_mine));
    // what the few instructions right after the wait need to know, in ONE register read before it
    int post = a.xact == 1 ? 4 : 0;
    if constexpr (FUSED) post |= (a.ar_world > 1 && tid == 0 ? 1 : 0) | (a.xact == 2 ? 2 : 0);
    asm volatile("" : "+r"(post));

    // codes of one item (row iteration `it` of a tile) -> registers; the stage is released at once
    int stage = 0, round = 0;
    auto load_codes = [&](uint32_t (&cw)[4][4], int it) {
      l2_mbar_wait(full0 + 8 * stage, round & 1);
      if (it * kL2Cons + perm < rows) {
        const uint32_t so = stage * kL2StageBytes;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          l2_lds64(caddr[v] + so, cw[v][0], cw[v][1]);
          l2_lds64((caddr[v] ^ 8u) + so, cw[v][2], cw[v][3]);
        }
      }
      __syncwarp();
      if (lane == 0) l2_mbar_arrive(empty0 + 8 * stage);
      if (++stage == a.stages) { stage = 0; ++round; }
    };
    // 64 lookups of one row; on the CTA's last tile the row's partial sum goes straight to the
    // CTA that finishes the row (distributed shared memory), else it is kept in rowsum
    auto lookups = [&](const uint32_t (&cw)[4][4], int s, int it) {
      const int row_l = it * kL2Cons + perm;
      if (row_l < rows) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int v = 0; v < 4; ++v) l2_lookup16<MASK>(acc, lut_base, cw[v], colv[v], sel, emask4);
        float r = (acc[0] + acc[1]) + (acc[2] + acc[3]);
        if (s > 0) r += rowsum[row_l];
        if (s + 1 < my_tiles || !clustered) {
          rowsum[row_l] = r;
        } else {
          const int o = row_l / chunk, j = row_l - o * chunk;
          uint32_t ra;
          const uint32_t la = recv_addr + (uint32_t)(rank * chunk + j) * 4u;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(o));
          asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(ra), "f"(r) : "memory");
        }
      }
    };

    l2_trace<TRACE>(a, 1);
    pdl_launch_dependents();                       // the next layer may start fetching its weights
    pdl_wait();                                    // x comes from the previous kernel
    l2_trace<TRACE>(a, 2);
    // fused all-reduce: this launch's epoch.  The counter is advanced by the last CTA of the previous
    // fused launch, which has completed by now; parked in shared memory, not in a register that
    // would live across the lookup loop.
    if (FUSED && (post & 1))
      *epoch_smem = *reinterpret_cast<volatile unsigned int*>(a.ar_state) + 1u;
    // RMSNorm on the x load: rstd from the producer's per-CTA sums of squares, added in a fixed
    // order by every warp on its own (no barrier; layernorm_kernels.cu:82-88)
    float rstd = 1.f;
    if (FUSED && (post & 2)) {
      float ss = 0.f;
      for (int j = lane; j < a.nrm_slots; j += 32) ss += __ldcg(a.nrm_ssq + (size_t)tok * a.nrm_slots + j);
      rstd = rsqrtf(warp_sum(ss) / (float)a.K + a.nrm_eps);
    }

    // ---- x of this CTA's tiles -> shared memory, B-fragment order (zero beyond K) ----
#pragma unroll
    for (int f = 0; f < XF; ++f) {
      if (tid + f * kL2Cons < XW) {
        for (int s = 0; s < my_tiles; ++s) {
          const int elem = xoff[f] + s * xstep;
          uint32_t v = 0u;
          if (xoff[f] >= 0 && elem < a.K) {
            v = __ldg(reinterpret_cast<const uint32_t*>(xg + elem));
            // x is [T, 2K] = (gate | up): SiLU(gate) * up on the way in (llama.py:89-93)
            if (post & 4) v = silu_mul_pair<T>(v, __ldg(reinterpret_cast<const uint32_t*>(xg + a.K + elem)));
            // x is the residual stream z: RMSNorm(z) * w on the way in (layernorm_kernels.cu:91-97)
            else if (FUSED && (post & 2))
              v = rms_norm_pair<T>(v, rstd, __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const T*>(a.nrm_w) + elem)));
          }
          xs_mine[s * XW + f * kL2Cons] = v;
        }
      }
    }
    l2_trace<TRACE>(a, 11);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    l2_trace<TRACE>(a, 12);
    l2_cons_sync<kL2Cons>();
    l2_trace_dep<TRACE>(a, 3, reinterpret_cast<const volatile uint32_t*>(xs) + 255);

    for (int s = 0; s < my_tiles; ++s) {
      if (s > 0) l2_cons_sync<kL2Cons>();                   // everyone is done with the previous table
      l2_build_table<T, D, CW>(lut, cbs, xs + (size_t)s * l2_x_words<D>() * 2, a.E, a.M, warp, lane);
      if (s == 0) l2_trace<TRACE>(a, 10);
      l2_cons_sync<kL2Cons>();
      if (s == 0) l2_trace_dep<TRACE>(a, 4, reinterpret_cast<const volatile uint32_t*>(lut) + (a.E - 1) * kL2Tile + 63);
      if (s == 1) l2_trace_dep<TRACE>(a, 13, reinterpret_cast<const volatile uint32_t*>(lut) + (a.E - 1) * kL2Tile + 63);
      if (s + 1 == my_tiles && clustered) asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
      for (int it = 0; it < n_iter; ++it) {
        uint32_t cw[4][4];
        load_codes(cw, it);
        lookups(cw, s, it);
        if (s == 0 && it == 0) l2_trace<TRACE>(a, 9);
        if (s == 1 && it == 0) l2_trace<TRACE>(a, 14);
        if (s == 0 && it == n_iter - 1) l2_trace<TRACE>(a, 15);
      }
    }
    if (my_tiles == 0) {                           // no tile in this K group: contribute zeros
      if (clustered) {
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
        for (int i = tid; i < rows; i += kL2Cons) {
          const int o = i / chunk, j = i - o * chunk;
          uint32_t ra;
          const uint32_t la = recv_addr + (uint32_t)(rank * chunk + j) * 4u;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(o));
          asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(ra), "f"(0.f) : "memory");
        }
      } else {
        for (int i = tid; i < rows; i += kL2Cons) rowsum[i] = 0.f;
      }
    }
    l2_trace<TRACE>(a, 5);
  }

  // ---- split-K over the cluster: the partial sums of rows [o, o+1) * chunk were pushed into CTA
  // o's receive buffer; ONE cluster barrier, then each CTA adds the CS partials of its own rows in
  // rank order (deterministic), scales and stores ----
  const float* fin = rowsum;                       // CS == 1: the sums are already final
  int fin_parts = 1, part_stride = 0, fin_off = 0;
  if (clustered) {
    l2_cluster_sync();
    fin = recv; fin_parts = a.CS; part_stride = chunk; fin_off = -fin0;
  } else {
    __syncthreads();
  }
  l2_trace<TRACE>(a, 6);
  // One finished row: (all-reduce over the TP ranks,) one rounding of y, (residual add,) store.
  // (the plain launch pays one flag test here: the fused steps sit behind it, their loops not unrolled,
  // so that nothing of theirs is hoisted onto the plain path - measured 0.3 us per launch when it was)
  const int efl = FUSED ? a.ep_flags : 0;          // 1: all-reduce, 2: residual, 4: ssq
  const uint32_t epoch = (efl & 1) ? *epoch_smem : 0u;
  float ssq_local = 0.f;
  auto store_row = [&](int row, float v) {
    T* yp = reinterpret_cast<T*>(a.y) + (size_t)tok * a.y_stride + row;
    if (efl == 0) { *yp = Traits<T>::from_f(v); return; }
    float res = 0.f;
    if (efl & 2) res = Traits<T>::to_f(reinterpret_cast<const T*>(a.resid)[(size_t)tok * a.resid_stride + row]);
    if (efl & 1) v = l2_allreduce_row(a, tok, row, v, epoch);
    T out = Traits<T>::from_f(v);
    if (efl & 2) out = Traits<T>::from_f(Traits<T>::to_f(out) + res);   // z = x + residual in T (layernorm_kernels.cu:75-80)
    *yp = out;
    const float zf = Traits<T>::to_f(out);
    ssq_local = fmaf(zf, zf, ssq_local);
  };
  for (int i = fin0 + tid; i < fin1; i += l2_threads<CW, COPY>()) {
    const int row = rb.row0 + i;
    float sc = Traits<T>::to_f(*reinterpret_cast<const T*>(&sc0_raw));
    float b = Traits<T>::to_f(*reinterpret_cast<const T*>(&bi0_raw));
    if (i != fin0 + tid) {                         // chunk > blockDim: later rows fetch theirs here
      sc = Traits<T>::to_f(reinterpret_cast<const T*>(a.scales)[row]);
      b = a.bias ? Traits<T>::to_f(reinterpret_cast<const T*>(a.bias)[row]) : 0.f;
    }
    float sum = 0.f;
    for (int r = 0; r < fin_parts; ++r) sum += fin[r * part_stride + i + fin_off];
    if (a.KG == 1)
      store_row(row, fmaf(sum, sc, b));
    else
      __stcg(a.ws_part + ((size_t)tok * a.KG + kg) * a.N + row, sum);
  }
  // ---- split-K over the K groups (KG > 1): partial sums meet in the workspace; the last of the KG
  // CTAs that share these rows (a ticket) adds them in K-group order - deterministic - and stores ----
  if (a.KG > 1) {
    int* last_flag = reinterpret_cast<int*>(bars + 2 * kL2MaxStages);
    __threadfence();
    __syncthreads();
    int* ticket = a.ws_tickets + ((size_t)tok * gridDim.y + block) * a.CS + rank;
    if (tid == 0) {
      const int t = atomicAdd(ticket, 1);
      *last_flag = (t == a.KG - 1);
      if (t == a.KG - 1) *ticket = 0;              // leave the counter zero for the next launch
    }
    __syncthreads();
    if (*last_flag) {
      __threadfence();
      for (int i = fin0 + tid; i < fin1; i += l2_threads<CW, COPY>()) {
        const int row = rb.row0 + i;
        float sc = Traits<T>::to_f(*reinterpret_cast<const T*>(&sc0_raw));
        float b = Traits<T>::to_f(*reinterpret_cast<const T*>(&bi0_raw));
        if (i != fin0 + tid) {
          sc = Traits<T>::to_f(reinterpret_cast<const T*>(a.scales)[row]);
          b = a.bias ? Traits<T>::to_f(reinterpret_cast<const T*>(a.bias)[row]) : 0.f;
        }
        float sum = 0.f;
        for (int g = 0; g < a.KG; ++g) sum += __ldcg(a.ws_part + ((size_t)tok * a.KG + g) * a.N + row);
        store_row(row, fmaf(sum, sc, b));
      }
    }
  }
  // ---- this CTA's share of sum(y^2) for the RMSNorm the consumer applies on its x load: warps in
  // order, then thread 0 in warp order - a fixed order ----
  if ((efl & 4) && (a.KG == 1 || *reinterpret_cast<int*>(bars + 2 * kL2MaxStages))) {
    __syncthreads();                               // everyone is done reading rowsum / recv
    const float w = warp_sum(ssq_local);
    if (lane == 0) rowsum[warp] = w;
    __syncthreads();
    if (tid == 0) {
      float tot = 0.f;
      for (int i = 0; i < l2_threads<CW, COPY>() / 32; ++i) tot += rowsum[i];
      // slot = (row block, rank): with K groups only the CTA that finished the rows writes, so
      // every slot is written exactly once per launch, whichever K group came last
      a.ssq[((size_t)tok * gridDim.y + block) * a.CS + rank] = tot;
    }
  }
  // ---- fused all-reduce: the last CTA of the launch advances this rank's epoch ----
  if (efl & 1) {
    __syncthreads();
    if (tid == 0) {
      const unsigned int n_cta = gridDim.x * gridDim.y * gridDim.z;
      if (atomicAdd(a.ar_state + 1, 1u) == n_cta - 1) {
        a.ar_state[1] = 0u;
        __threadfence();
        *reinterpret_cast<volatile unsigned int*>(a.ar_state) = epoch;
      }
    }
  }
  l2_trace<TRACE>(a, 7);
}

// ------------------------------------------------------------------ host side
extern bool g_use_pdl;
extern unsigned long long* g_lut_trace;
extern int g_lut_last_grid[4];

struct Lut2Plan {
  int cw, cs, S, kg, n_tiles, rows_per_cluster, stages, smem_cb, smem_x, smem_total;
};

static int lut2_smem_bytes(int E, int stages, int cw, int smem_cb, int smem_x, int rows) {
  return E * kL2Tile * 4 + stages * cw * 32 * kL2Tile + smem_cb + smem_x + (2 * rows + kL2MaxCluster) * 4 +
         2 * 8 * kL2MaxStages + 16;                // mbarriers + the K-group ticket flag
}

constexpr int kL2TwoPerSm = 112 * 1024;            // two CTAs (this layer's + the next layer's) fit one SM:
                                                   // 228 KiB - 2 x (1 KiB static + 1 KiB reserved)

static int lut2_log2(int cs) { return cs == 1 ? 0 : cs == 2 ? 1 : cs == 4 ? 2 : cs == 8 ? 3 : 4; }

template <int CW>
static int lut2_query_clusters(int cs, int smem) {
  auto kern = pq_gemv_lut2_kernel<__half, 8, false, false, CW, false>;
  int n = -1;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024) != cudaSuccess) return -1;
  if (cs > 8 && cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) return -1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(cs * 1024, 1, 1); cfg.blockDim = dim3(l2_threads<CW, false>()); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) n = -1;
  return n;
}

// Co-resident clusters of `cs` CTAs of shape `cw` at `smem` bytes per CTA (GPC granularity strands
// a few SMs for cs >= 4).  Without a device: B200 figures.
static int lut2_cluster_slots(int cw, int cs, int smem) {
  static int cache[16][2][5][64] = {};             // device, shape, log2(cs), smem / 4 KiB
  int dev = 0;
  const bool have_dev = cudaGetDevice(&dev) == cudaSuccess && dev < 16;
  const int l2 = lut2_log2(cs), sb = min(63, smem / 4096), w = cw == 8 ? 0 : 1;
  if (have_dev && cache[dev][w][l2][sb] > 0) return cache[dev][w][l2][sb];
  int n = -1;
  if (have_dev) n = cw == 8 ? lut2_query_clusters<8>(cs, (sb + 1) * 4096) : lut2_query_clusters<16>(cs, (sb + 1) * 4096);
  if (n <= 0) {
    cudaGetLastError();
    static const int one[5] = {148, 74, 33, 15, 7}, two[5] = {296, 148, 71, 33, 15};
    n = (cw == 8 && smem <= kL2TwoPerSm) ? two[l2] : one[l2];
  }
  if (have_dev) cache[dev][w][l2][sb] = n;
  return n;
}

static int64_t lut2_blocks(const int64_t* part_sizes, int n_parts, int64_t n, int rows) {
  if (!part_sizes) return (n + rows - 1) / rows + (n_parts - 1);
  int64_t b = 0;
  for (int i = 0; i < n_parts; ++i) b += (part_sizes[i] + rows - 1) / rows;
  return b;
}

constexpr int kL2MaxKG = 8;                        // K groups (clusters along K) per row block
constexpr int kL2TicketBytes = 64 * 1024;          // head of the workspace (include/pq_b200.h)

static int lut2_env(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? (int)strtol(e, nullptr, 0) : dflt;
}

// CTA shape CW, cluster size CS, tiles per CTA S, K groups KG (CS * S * KG >= n_tiles) and rows per
// row block from a small cycle model calibrated on B200 phase traces (profiles/): a CTA pays a
// fixed cost, a cost per table build and one per row iteration of lookups; a cluster reduce costs
// ~1200 cycles, the cross-cluster (workspace + ticket) reduce ~3000.
//   CW = 16: one CTA per SM.   CW = 8: planned at `per_sm8` CTAs of this launch per SM (1: the
//   other half of the SM is left to the next layer's CTA, whose prologue then overlaps this
//   layer under programmatic dependent launch; 2: both halves work on this layer).
static bool lut2_plan(Lut2Plan& pl, int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                      const int64_t* part_sizes, int n_parts, bool allow_kg = true) {
  const int E = 1 << nbits;
  const int64_t gp = k / d * m;
  pl.n_tiles = (int)((gp + kL2Tile - 1) / kL2Tile);
  pl.smem_cb = m * E * d * 2;
  static const int force_cw = lut2_env("PQ_LUT2_CW", 0), force_cs = lut2_env("PQ_LUT2_CS", 0);
  static const int force_kg = lut2_env("PQ_LUT2_KG", 0), force_stages = lut2_env("PQ_LUT2_STAGES", 0);
  static const int per_sm8 = lut2_env("PQ_LUT2_PER_SM8", 1);
  static const int fixed8 = lut2_env("PQ_LUT2_FIXED8", 3000), fixed16 = lut2_env("PQ_LUT2_FIXED16", 5000);
  static const int build8 = lut2_env("PQ_LUT2_BUILD8", 2500), item8 = lut2_env("PQ_LUT2_ITEM8", 1100);
  const int sms = sm_count();
  double best = 1e30;
  bool found = false;
  for (int cw = 8; cw <= 16; cw += 8) {
    if (force_cw && cw != force_cw) continue;
    const int cons = cw * 32;
    const int smem_cap = cw == 8 ? kL2TwoPerSm : 220 * 1024;
    for (int cs = 1; cs <= kL2MaxCluster; cs *= 2) {
      if (cs > 1 && cs / 2 >= pl.n_tiles) break;   // up to the next power of two: the ranks left without a tile add zeros
      if (force_cs && cs != force_cs) continue;
      if (cs == 16 && !force_cs && cw == 16) continue;           // 16-CTA clusters: the small shape only
      int slots1 = lut2_cluster_slots(cw, cs, cw == 8 ? 100 * 1024 : 200 * 1024);
      const int per_sm = cw == 8 ? per_sm8 : 1;
      if (slots1 * cs > sms * per_sm) slots1 = sms * per_sm / cs;
      if (slots1 < 1) continue;
      for (int kg = 1; kg <= (allow_kg ? kL2MaxKG : 1); ++kg) {
        if (force_kg && kg != force_kg && allow_kg) continue;
        const int S = (pl.n_tiles + cs * kg - 1) / (cs * kg);
        if ((kg - 1) * cs * S >= pl.n_tiles) continue;           // the last K group would be empty
        const int smem_x = S * (d == 16 ? 512 : 256) * 4;        // B-fragment order, see l2_x_offset
        if (smem_x > 32 * 1024) continue;
        int target = slots1 / (int)(t * kg);                     // row blocks in one wave
        if (target < 1) target = 1;
        int rows = (int)((n + target - 1) / target);
        rows = (rows + kL2BoxRows - 1) / kL2BoxRows * kL2BoxRows;
        const int max_rows = 4096;
        if (rows > max_rows) rows = max_rows;
        while (rows < max_rows && lut2_blocks(part_sizes, n_parts, n, rows) > target) rows += kL2BoxRows;
        const int n_iter = (rows + cons - 1) / cons;
        // measured on B200 (gpurun r2_01, profiles/r2_01_*): the 8-warp CTA wins only where a CTA has a
        // single item (o_proj: 4.6 against 5.0 us); with more items its half-size lookup crew loses
        // more than the co-resident prologue gains (qkv 6.7 / 6.3, gate_up 8.7 / 7.6, down 9.2 / 8.1 us)
        // ... except with a 16-CTA cluster (8-warp shape only) that gives every CTA ONE tile: saving the
        // second table build outweighs a second row iteration (down_proj shard at TP = 2, 11 tiles:
        // 5.08 us against 5.91 us for 8 CTAs x 2 tiles of the 16-warp shape; gpurun r2_37)
        if (cw == 8 && !force_cw && S * n_iter > 1 && !(cs == 16 && S == 1 && n_iter <= 2)) continue;
        int stages = force_stages ? min(kL2MaxStages, max(1, force_stages)) : min(cw == 8 ? kL2MaxStages : 4, max(2, S * n_iter));
        auto smem_of = [&](int st) { return lut2_smem_bytes(E, st, cw, pl.smem_cb, smem_x, rows); };
        while (stages > (cw == 8 ? 1 : 2) && smem_of(stages) > smem_cap) --stages;
        const int smem = smem_of(stages);
        if (smem > smem_cap) continue;
        const int64_t clusters = lut2_blocks(part_sizes, n_parts, n, rows) * kg * t;
        if (lut2_blocks(part_sizes, n_parts, n, rows) > 65535 || t > 65535) continue;   // grid.y, grid.z
        const int64_t waves = (clusters + slots1 - 1) / slots1;
        // narrower clusters stream narrower strips of the code matrix and measured ~20 % slower per
        // row iteration; K groups only where they clearly pay (profiles/r1_18_microbench_lut_tma.jsonl)
        double cta;
        if (cw == 16) {
          const double item = 1900.0 * (cs < 8 ? 1.2 : 1.0);
          cta = fixed16 + (cs > 1 ? 1200.0 : 0.0) + (kg > 1 ? 3000.0 : 0.0) +
                S * (2500.0 * E / 256.0 + (rows / 512.0 + 0.3) * item);
        } else {
          const double item = item8 * (cs < 8 ? 1.2 : 1.0) * (per_sm == 2 ? 1.6 : 1.0);
          cta = fixed8 + (per_sm == 2 ? fixed16 - fixed8 : 0) + (cs > 1 ? 1200.0 : 0.0) + (cs > 8 ? 400.0 : 0.0) +
                (kg > 1 ? 3000.0 : 0.0) + S * (build8 * E / 256.0 + (rows / 256.0 + 0.3) * item);
        }
        const double cost = (double)waves * cta * (kg > 1 ? 1.15 : 1.0);
        if (cost < best) {
          best = cost; found = true;
          pl.cw = cw; pl.cs = cs; pl.S = S; pl.kg = kg; pl.rows_per_cluster = rows; pl.stages = stages;
          pl.smem_x = smem_x; pl.smem_total = smem;
        }
      }
    }
  }
  return found;
}

int gemv_lut2_debug_slots(int cw, int cs, int smem) {
  return cw == 8 ? lut2_query_clusters<8>(cs, smem) : lut2_query_clusters<16>(cs, smem);
}

// {cw, cs, S, kg, rows per cluster, stages, smem bytes, n_tiles, grid.x} of the plan (diagnostics, tests)
int gemv_lut2_debug_plan(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                         int n_parts, int* out9) {
  Lut2Plan pl;
  if (!lut2_plan(pl, t, n, k, d, nbits, m, part_sizes, n_parts)) return PQ_ERR_UNSUPPORTED;
  const int v[9] = {pl.cw, pl.cs, pl.S, pl.kg, pl.rows_per_cluster, pl.stages, pl.smem_total, pl.n_tiles,
                    (int)(lut2_blocks(part_sizes, n_parts, n, pl.rows_per_cluster) * pl.kg * pl.cs)};
  for (int i = 0; i < 9; ++i) out9[i] = v[i];
  return PQ_OK;
}

// CTAs that finish rows (= floats per token of the fused epilogue's ssq output), or a negative error
int64_t gemv_lut2_grid_x(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                         int n_parts, int64_t workspace_bytes) {
  Lut2Plan pl;
  const bool have_ws = workspace_bytes >= kL2TicketBytes + (int64_t)kL2MaxKG * t * n * 4;
  if (!lut2_plan(pl, t, n, k, d, nbits, m, part_sizes, n_parts, have_ws)) return PQ_ERR_UNSUPPORTED;
  return lut2_blocks(part_sizes, n_parts, n, pl.rows_per_cluster) * pl.cs;
}

// Upper bound for every plan (the plan may depend on the partition sizes): tickets + KG partials.
int64_t gemv_lut2_workspace(int64_t t, int64_t n, int64_t, int, int, int, int) {
  return kL2TicketBytes + (int64_t)kL2MaxKG * t * n * 4;
}

bool gemv_lut2_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (nbits > 8 || nbits < 4 || t > kL2MaxT || t < 1) return false;   // E >= 16: whole 16-entry MMA blocks
  if (d != 4 && d != 8 && d != 16) return false;
  if (m != 1 && m != 2 && m != 4) return false;
  if (k % d) return false;
  const int64_t gp = k / d * m;
  if (gp % 4 || gp < kL2Tile || n < kL2BoxRows) return false;          // TMA box inside the tensor; cp.async >= 4 bytes
  if (gp > (int64_t)1 << 30 || n > (int64_t)1 << 30) return false;
  Lut2Plan pl;
  return lut2_plan(pl, t, n, k, d, nbits, m, nullptr, PQ_MAX_PARTS);
}

// The fused prologue / epilogue is instantiated for 8-bit codes only.
bool gemv_lut2_fused_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  return nbits == 8 && gemv_lut2_supported(t, n, k, d, nbits, m);
}

// PQ_ALGO_AUTO heuristic: tiny problems are better served by the direct kernel.
bool gemv_lut2_preferred(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  static const bool off = [] { const char* e = getenv("PQ_AUTO_LUT2"); return e && atoi(e) == 0; }();
  return !off && gemv_lut2_supported(t, n, k, d, nbits, m) && n * (k / d * m) >= (1 << 18);
}

typedef CUresult (*PFN_l2TmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                          const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                          CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                          CUtensorMapFloatOOBfill);

static PFN_l2TmapEncodeTiled lut2_tmap_encoder() {
  static PFN_l2TmapEncodeTiled fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (PFN_l2TmapEncodeTiled)p;
  }();
  return fn;
}

// codes [N, Gp] bytes, row-major, as a 2-D tensor map: box = 64 code bytes x 64 rows, no swizzle,
// out-of-bounds rows and positions read as zero.
static int lut2_codes_tmap(CUtensorMap* map, const void* codes, int64_t n, int64_t gp, int64_t pitch) {
  PFN_l2TmapEncodeTiled enc = lut2_tmap_encoder();
  if (!enc) return PQ_ERR_UNSUPPORTED;
  const cuuint64_t gdim[2] = {(cuuint64_t)gp, (cuuint64_t)n};
  const cuuint64_t gstride[1] = {(cuuint64_t)pitch};
  const cuuint32_t box[2] = {(cuuint32_t)kL2Tile, (cuuint32_t)kL2BoxRows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(codes), gdim, gstride, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                         CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? PQ_OK : PQ_ERR_ALIGN;
}

template <typename T, int D, bool MASK, bool COPY, int CW, bool FUSED, bool TRACE = false>
static int launch_lut2_c(const CUtensorMap& map, Lut2Args& a, const Lut2Plan& pl, cudaStream_t stream) {
  auto kern = pq_gemv_lut2_kernel<T, D, MASK, COPY, CW, FUSED, TRACE>;
  static bool attr_set[64] = {};                   // per instantiation, per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return (int)e;
    attr_set[dev & 63] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(pl.kg * pl.cs, a.parts.cta_end[a.parts.n_parts - 1], a.T);
  cfg.blockDim = dim3(l2_threads<CW, COPY>());
  cfg.dynamicSmemBytes = pl.smem_total;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = pl.cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 2 : 1;
  if (a.trace) {                                   // diagnostics only (tools/trace_lut.py)
    g_lut_last_grid[0] = cfg.gridDim.x * cfg.gridDim.y; g_lut_last_grid[1] = cfg.gridDim.z;
    g_lut_last_grid[2] = pl.cs; g_lut_last_grid[3] = pl.smem_total;
  }
  return (int)cudaLaunchKernelEx(&cfg, kern, map, a);
}

template <typename T, int D, bool MASK>
static int launch_lut2(const CUtensorMap& map, Lut2Args& a, const Lut2Plan& pl, cudaStream_t stream) {
  if constexpr (std::is_same<T, __half>::value && D == 8 && !MASK) {      // phase trace: the bench's instantiation only
    if (a.trace != nullptr && !a.copy_w && !a.ep_flags && a.xact != 2)
      return pl.cw == 8 ? launch_lut2_c<T, D, false, false, 8, false, true>(map, a, pl, stream)
                        : launch_lut2_c<T, D, false, false, 16, false, true>(map, a, pl, stream);
  }
  if constexpr (!MASK) {                           // fused steps: 8-bit codes only (gemv_lut2_fused_supported)
    if (a.ep_flags || a.xact == 2) {
      if (pl.cw == 8)
        return a.copy_w ? launch_lut2_c<T, D, false, true, 8, true>(map, a, pl, stream)
                        : launch_lut2_c<T, D, false, false, 8, true>(map, a, pl, stream);
      return a.copy_w ? launch_lut2_c<T, D, false, true, 16, true>(map, a, pl, stream)
                      : launch_lut2_c<T, D, false, false, 16, true>(map, a, pl, stream);
    }
  }
  if (pl.cw == 8)
    return a.copy_w ? launch_lut2_c<T, D, MASK, true, 8, false>(map, a, pl, stream)
                    : launch_lut2_c<T, D, MASK, false, 8, false>(map, a, pl, stream);
  return a.copy_w ? launch_lut2_c<T, D, MASK, true, 16, false>(map, a, pl, stream)
                  : launch_lut2_c<T, D, MASK, false, 16, false>(map, a, pl, stream);
}

template <typename T>
static int launch_lut2_d(const CUtensorMap& map, Lut2Args& a, int d, const Lut2Plan& pl, cudaStream_t stream) {
  if (a.mask) {
    switch (d) {
      case 4: return launch_lut2<T, 4, true>(map, a, pl, stream);
      case 8: return launch_lut2<T, 8, true>(map, a, pl, stream);
      default: return launch_lut2<T, 16, true>(map, a, pl, stream);
    }
  }
  switch (d) {
    case 4: return launch_lut2<T, 4, false>(map, a, pl, stream);
    case 8: return launch_lut2<T, 8, false>(map, a, pl, stream);
    default: return launch_lut2<T, 16, false>(map, a, pl, stream);
  }
}

int gemv_lut2_dispatch(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                       const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                       int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                       int n_parts, int dtype, void* workspace, int64_t workspace_bytes,
                       cudaStream_t stream, int xact, const pq_b200_epilogue_t* ep) {
  if ((uintptr_t)codes % 16 || (uintptr_t)x % 4 || (x_stride * 2) % 4) return PQ_ERR_ALIGN;
  Lut2Plan pl;
  // K groups meet in the workspace; without one (it is optional) the plan keeps each row block in one cluster
  const bool have_ws = workspace && workspace_bytes >= kL2TicketBytes + (int64_t)kL2MaxKG * t * n * 4;
  if (!lut2_plan(pl, t, n, k, d, nbits, m, part_sizes, n_parts, have_ws)) return PQ_ERR_UNSUPPORTED;
  Lut2Args a{};
  a.x = x; a.codebooks = codebooks; a.scales = scales; a.bias = bias;
  a.y = y; a.x_stride = x_stride; a.y_stride = y_stride;
  a.T = (int)t; a.N = (int)n; a.K = (int)k; a.E = 1 << nbits; a.M = m;
  a.m_log2 = m == 1 ? 0 : m == 2 ? 1 : 2;
  a.Gp = (int)(k / d * m); a.n_tiles = pl.n_tiles;
  a.codes = (const unsigned char*)codes;
  a.copy_w = a.Gp % 16 == 0 ? 0 : (a.Gp % 8 == 0 ? 8 : 4);
  a.S = pl.S; a.CS = pl.cs; a.cs_log2 = lut2_log2(pl.cs);
  a.rows_per_cluster = pl.rows_per_cluster; a.stages = pl.stages; a.KG = pl.kg;
  if (pl.kg > 1) {
    const int64_t blocks = lut2_blocks(part_sizes, n_parts, n, pl.rows_per_cluster);
    if (!workspace || workspace_bytes < kL2TicketBytes + (int64_t)pl.kg * t * n * 4 ||
        t * blocks * pl.cs * 4 > kL2TicketBytes)
      return PQ_ERR_WORKSPACE;
    a.ws_tickets = reinterpret_cast<int*>(workspace);
    a.ws_part = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + kL2TicketBytes);
  }
  static const int prefetch = [] { const char* e = getenv("PQ_LUT2_PREFETCH"); return e ? atoi(e) : 8; }();
  a.prefetch = prefetch;

  a.smem_cb = pl.smem_cb; a.smem_x = pl.smem_x; a.mask = nbits < 8 ? 1 : 0;
  a.trace = g_lut_trace; a.xact = xact;
  if (ep != nullptr) {
    if (nbits != 8) return PQ_ERR_UNSUPPORTED;
    const int64_t ssq_slots = lut2_blocks(part_sizes, n_parts, n, pl.rows_per_cluster) * pl.cs;
    if (ep->ar_world > 1) {
      if (ep->ar_world > PQ_AR_MAX_WORLD || ep->ar_rank < 0 || ep->ar_rank >= ep->ar_world) return PQ_ERR_SHAPE;
      if (!ep->ar_peer || !ep->ar_state) return PQ_ERR_NULL;
      if (t * n > ep->ar_slot_elems) return PQ_ERR_WORKSPACE;
      a.ar_world = ep->ar_world; a.ar_rank = ep->ar_rank; a.ar_slot = ep->ar_slot_elems;
      a.ar_state = reinterpret_cast<unsigned int*>(ep->ar_state);
      for (int r = 0; r < ep->ar_world; ++r) {
        if (!ep->ar_peer[r] || ((uintptr_t)ep->ar_peer[r] & 7)) return PQ_ERR_ALIGN;
        a.ar_peer[r] = reinterpret_cast<unsigned long long*>(ep->ar_peer[r]);
      }
    }
    if (ep->residual) { a.resid = ep->residual; a.resid_stride = ep->residual_stride; }
    if (ep->ssq_out) {
      if (ep->ssq_slots != ssq_slots) return PQ_ERR_WORKSPACE;       // pq_b200_gemm_fused_slots() of this problem
      a.ssq = ep->ssq_out;
    }
    if (ep->norm_weight) {
      if (xact || !ep->norm_ssq || ep->norm_slots < 1) return PQ_ERR_NULL;
      if ((uintptr_t)ep->norm_weight % 4) return PQ_ERR_ALIGN;
      a.xact = 2; a.nrm_ssq = ep->norm_ssq; a.nrm_w = ep->norm_weight; a.nrm_slots = ep->norm_slots;
      a.nrm_eps = ep->norm_eps;
    }
  }
  static const int early_ld = lut2_env("PQ_LUT2_EARLY_LD", 0);
  a.early_ld = early_ld;
  a.trace_mask = lut2_env("PQ_LUT2_TRACE_MASK", 0xffff);
  a.ep_flags = (a.ar_world > 1 ? 1 : 0) | (a.resid ? 2 : 0) | (a.ssq ? 4 : 0);
  int rc = make_parts(a.parts, part_sizes, n_parts, n, pl.rows_per_cluster);
  if (rc) return rc;
  CUtensorMap map;
  const bool padded = ep != nullptr && ep->codes_padded != nullptr;
  if (padded) {
    if (ep->codes_pitch % 16 || ep->codes_pitch < a.Gp || ((uintptr_t)ep->codes_padded & 15)) return PQ_ERR_ALIGN;
    a.copy_w = 0;
  }
  const int64_t pitch = padded ? ep->codes_pitch : (a.Gp + 15) / 16 * 16;          // map unused when copy_w
  rc = lut2_codes_tmap(&map, padded ? ep->codes_padded : codes, n, padded ? a.Gp : pitch, pitch);
  if (rc) return rc;
  return dtype == PQ_F16 ? launch_lut2_d<__half>(map, a, d, pl, stream)
                         : launch_lut2_d<__nv_bfloat16>(map, a, d, pl, stream);
}

}  // namespace pq
