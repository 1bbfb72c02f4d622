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
// ONE launch per linear, no workspace.  Work decomposition:
//   cluster  = one block of rows (never straddling an output partition); its CS CTAs split K
//   CTA c    = position tiles {c, c + CS, c + 2 CS, ...} of TP = 64 / TT positions each
//   thread   = 16 consecutive code bytes of one row per (tile, row iteration)
// Pipeline inside a CTA:
//  1. BEFORE griddepcontrol.wait (weights do not depend on the previous kernel): the
//     partition's codebooks and the first kRing x 4 KiB of codes are requested with cp.async
//     (16 B per thread, L1 bypass).  Every thread copies exactly the bytes it will later
//     consume, so the ring needs no barrier: cp.async.wait_group is per thread.  Under
//     programmatic dependent launch this CTA is resident next to the previous layer's CTAs
//     (<= 110 KiB shared memory each) and the HBM stream never stops between layers.
//  2. after the wait: x of this CTA's tiles -> shared memory; per tile the table
//     [E entries][TT tokens][TP positions] fp32 is built on the tensor cores (mma.sync
//     m16n8k8 / k16, operands from shared memory) - a tiny (E x d).(d x TP) GEMM.
//  3. lookups: table row stride is 256 B, so ONE byte-permute builds the whole table offset
//     (code << 8 | column) from the packed code word; column = position mod 32 = bank, and each
//     lane walks its 16 bytes from a rotated start so the 32 lanes of a warp always hit 32
//     distinct banks.  Per code: PRMT + LDS + FADD.
//  4. split-K: per-row partial sums live in each CTA's shared memory; after a cluster barrier
//     CTA c sums rows [c, c+1) * rows/CS of all CS peers over distributed shared memory in rank
//     order (deterministic), applies scale and bias in fp32 and writes y.
#include <cstdlib>
#include <type_traits>

#include "pq_common.cuh"

namespace pq {

constexpr int kLutThreads = 256;
constexpr int kLutWarps = kLutThreads / 32;
constexpr int kLutRowFloats = 64;                  // table row: 256 B = TT tokens x TP positions
constexpr int kRing = 4;                           // code chunks (8 KiB each) in flight per CTA
constexpr int kRingVec = kLutThreads * 16;         // one 16-byte vector per thread
constexpr int kRingChunk = 2 * kRingVec;           // up to two vectors per thread and item
constexpr int kLutMaxRowsum = 16 * 1024;           // bytes of per-row partial sums
constexpr int kLutMaxX = 16 * 1024;                // bytes of staged x
constexpr int kLutMaxCluster = 8;

struct LutArgs {
  const void* x; const unsigned char* codes; const void* codebooks; const void* scales; const void* bias;
  void* y;
  long long x_stride, y_stride;
  int T, N, K, E, M;
  int Gp;                 // positions (= code bytes) per row
  int n_tiles;            // ceil(Gp / TP)
  int S;                  // tiles per CTA = ceil(n_tiles / CS)
  int CS;                 // cluster size
  int rows_per_cluster;
  int smem_cb, smem_x;    // bytes
  unsigned long long* trace;   // optional per-CTA phase timestamps (tools/trace_lut.py); NULL in production
  Parts parts;            // cta_end counts clusters
};

constexpr int kTraceSlots = 16;
__device__ __forceinline__ void trace_mark(const LutArgs& a, int slot) {
  if (a.trace != nullptr && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.trace[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * kTraceSlots + slot] = t;
  }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, uint32_t rank) {
  uint32_t ra; float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra));
  return v;
}

template <int OFF> __device__ __forceinline__ float lds_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(OFF));
  return v;
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

// Table for 32 positions of one token: lut[e * 64 + q] = <cb[book(q)][e], x[group(q)]>, e < E.
// cb: this partition's codebooks [M][E][D] in shared memory; xg: the x groups of these 32
// positions ([32 / M][D], zero beyond K) in shared memory.
// MMA mapping (m16n8): rows = 16 entries, columns = 8 positions of ONE codebook.
// n-block j (0..3) uses book j % M; its column c = 2*tq + i is position 8*tq + slot(j, i) with
// slot(j, i) = j % M + M * (2 * (j / M) + i), so that after the four n-blocks lane (g, tq) holds
// positions 8*tq .. 8*tq + 7 of entries g and g + 8: two 128-bit stores per entry row.
template <typename T, int D, int M>
__device__ __forceinline__ void build_lut_32(float* lut, const T* cb, const T* xg, int E, int warp, int lane) {
  const int g = lane >> 2, tq = lane & 3;
  constexpr int EW = D / 2;                       // u32 words per codebook entry / x group
  const uint32_t* cb32 = reinterpret_cast<const uint32_t*>(cb);
  const uint32_t* x32 = reinterpret_cast<const uint32_t*>(xg);

  uint32_t b0[4], b1[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int slot = (j % M) + M * (2 * (j / M) + (g & 1));
    const int grp = (8 * (g >> 1) + slot) / M;
    b0[j] = 0u; b1[j] = 0u;
    if constexpr (D == 4) { if (tq < 2) b0[j] = x32[grp * EW + tq]; }
    else if constexpr (D == 8) { b0[j] = x32[grp * EW + tq]; }
    else { b0[j] = x32[grp * EW + tq]; b1[j] = x32[grp * EW + tq + 4]; }
  }
  for (int eb = warp; eb * 16 < E; eb += kLutWarps) {
    const int e0 = eb * 16 + g, e1 = e0 + 8;
    uint32_t fa[M][4];
#pragma unroll
    for (int m = 0; m < M; ++m) {
      const uint32_t* ca = cb32 + (m * E + e0) * EW;
      const uint32_t* cbp = ca + 8 * EW;
      const bool ok = (D != 4 || tq < 2);
      fa[m][0] = ok ? ca[tq] : 0u;
      fa[m][1] = ok ? cbp[tq] : 0u;
      if constexpr (D == 16) { fa[m][2] = ca[tq + 4]; fa[m][3] = cbp[tq + 4]; }
    }
    float out0[8], out1[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = j % M;
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      if constexpr (D == 16) mma_k16<T>(c, fa[m][0], fa[m][1], fa[m][2], fa[m][3], b0[j], b1[j]);
      else mma_k8<T>(c, fa[m][0], fa[m][1], b0[j]);
      const int s0 = (j % M) + M * (2 * (j / M)), s1 = s0 + M;
      out0[s0] = c[0]; out0[s1] = c[1]; out1[s0] = c[2]; out1[s1] = c[3];
    }
    float4* r0 = reinterpret_cast<float4*>(lut + e0 * kLutRowFloats + 8 * tq);
    float4* r1 = reinterpret_cast<float4*>(lut + e1 * kLutRowFloats + 8 * tq);
    // odd entry rows store their halves in the other order: the 8 lanes of a 128-bit store phase
    // (g in {2k, 2k+1} x tq) then cover all 32 banks
    const int h = g & 1;
    r0[h] = h ? make_float4(out0[4], out0[5], out0[6], out0[7]) : make_float4(out0[0], out0[1], out0[2], out0[3]);
    r0[h ^ 1] = h ? make_float4(out0[0], out0[1], out0[2], out0[3]) : make_float4(out0[4], out0[5], out0[6], out0[7]);
    r1[h] = h ? make_float4(out1[4], out1[5], out1[6], out1[7]) : make_float4(out1[0], out1[1], out1[2], out1[3]);
    r1[h ^ 1] = h ? make_float4(out1[0], out1[1], out1[2], out1[3]) : make_float4(out1[4], out1[5], out1[6], out1[7]);
  }
}

template <typename T, int D, int TT>
__global__ void __launch_bounds__(kLutThreads, 2)
pq_gemv_lut_kernel(const LutArgs a) {
  constexpr int TP = kLutRowFloats / TT;          // positions per tile (64 or 32)
  constexpr int VPL = TP / 32;                    // 16-byte code vectors per lane and item
  constexpr int RPI = 16 * kLutWarps;             // rows per CTA iteration: 2 lanes per row
  extern __shared__ __align__(256) unsigned char lut_smem[];
  float* lut = reinterpret_cast<float*>(lut_smem);
  unsigned char* ring = lut_smem + (size_t)a.E * (kLutRowFloats * 4);
  T* cbs = reinterpret_cast<T*>(ring + kRing * kRingChunk);
  T* xs = reinterpret_cast<T*>(reinterpret_cast<unsigned char*>(cbs) + a.smem_cb);
  float* rowsum = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(xs) + a.smem_x);
  float* recv = rowsum + a.rows_per_cluster * TT;  // [CS][chunk * TT] partial sums pushed by the peers

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rank = a.CS > 1 ? (int)cluster_ctarank() : 0;
  const int z = blockIdx.y;
  trace_mark(a, 0);
  const RowBlock rb = locate_row_block(a.parts, blockIdx.x / a.CS, a.rows_per_cluster);
  const int rows = rb.row1 - rb.row0;
  const int n_iter = (rows + RPI - 1) / RPI;
  const int sub = lane & 1, rw = lane >> 1;
  const int row_in_iter = warp * 16 + rw;
  const int my_iters = row_in_iter < rows ? (rows - row_in_iter + RPI - 1) / RPI : 0;

  // ---- weights: codebooks + the first kRing code chunks, independent of the previous kernel ----
  {
    const unsigned char* cb_g = reinterpret_cast<const unsigned char*>(a.codebooks) +
                                (size_t)rb.part * a.smem_cb;
    const uint32_t dst = smem_u32(cbs);
    for (int i = tid; i < a.smem_cb / 16; i += kLutThreads) cp_async_16(dst + i * 16, cb_g + (size_t)i * 16);
    cp_async_commit();
  }
  const uint32_t my_slot0 = smem_u32(ring) + tid * 16;
  const unsigned char* my_codes = a.codes + (size_t)(rb.row0 + row_in_iter) * a.Gp + sub * 16;
  const size_t iter_stride = (size_t)RPI * a.Gp;
  int is = 0, iit = 0, islot = 0;                  // issue cursor: (tile step, row iteration, ring slot)
  const unsigned char* isrc = my_codes + (size_t)rank * TP;
  int iq = rank * TP + sub * 16;                   // first position of this lane's vector 0
  auto issue = [&]() {
    if (is < a.S) {
      if (iit < my_iters) {
#pragma unroll
        for (int v = 0; v < VPL; ++v)
          if (iq + 32 * v < a.Gp) cp_async_16(my_slot0 + islot * kRingChunk + kRingVec * v, isrc + 32 * v);
      }
      isrc += iter_stride;
      if (++iit == n_iter) {
        iit = 0; ++is;
        iq += a.CS * TP;
        isrc = my_codes + (size_t)(is * a.CS + rank) * TP;
      }
      islot = (islot + 1 == kRing) ? 0 : islot + 1;
    }
    cp_async_commit();                             // always: keeps the group count uniform
  };
#pragma unroll
  for (int j = 0; j < kRing; ++j) issue();
  for (int i = tid; i < a.rows_per_cluster * TT; i += kLutThreads) rowsum[i] = 0.f;

  // the rows this CTA finishes: scale and bias are weights too - fetch them now
  const int chunk = (rows + a.CS - 1) / a.CS;
  const int fin0 = rank * chunk, fin1 = min(rows, fin0 + chunk);
  unsigned short sc0_raw = 0, bi0_raw = 0;         // raw bits: converting here would stall on the load
  if (fin0 + tid < fin1) {
    sc0_raw = __ldg(reinterpret_cast<const unsigned short*>(a.scales) + rb.row0 + fin0 + tid);
    if (a.bias) bi0_raw = __ldg(reinterpret_cast<const unsigned short*>(a.bias) + rb.row0 + fin0 + tid);
  }

  trace_mark(a, 1);
  pdl_launch_dependents();                         // the next layer may start fetching its weights
  pdl_wait();                                      // x comes from the previous kernel
  trace_mark(a, 2);

  // ---- x of this CTA's tiles -> shared memory (zero beyond K / T) ----
  {
    const int wpt = TP / a.M * D / 2;              // u32 words per (tile, token)
    uint32_t* xs32 = reinterpret_cast<uint32_t*>(xs);
    const T* xg = reinterpret_cast<const T*>(a.x);
    for (int s = 0; s < a.S; ++s) {
      const int tile = s * a.CS + rank;
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        const int tok = z * TT + t;
        for (int w = tid; w < wpt; w += kLutThreads) {
          const int elem = tile * (TP / a.M) * D + 2 * w;
          uint32_t v = 0u;
          if (tile < a.n_tiles && tok < a.T && elem < a.K)
            v = __ldg(reinterpret_cast<const uint32_t*>(xg + (size_t)tok * a.x_stride + elem));
          xs32[(s * TT + t) * wpt + w] = v;
        }
      }
    }
  }
  cp_async_wait<kRing>();                          // codebooks have landed
  __syncthreads();
  trace_mark(a, 3);

  // lane-constant table columns: step k of vector v reads position 32*v + 16*sub + (k + rw) % 16;
  // bank = 16*sub + (k + rw) % 16: the 32 lanes (16 rows x 2) of a warp hit 32 distinct banks
  uint32_t col[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) col[k] = (uint32_t)(16 * sub + ((k + rw) & 15)) * 4u;
  const uint32_t lut_base = smem_u32(lut);
  const uint32_t emask4 = ((uint32_t)a.E - 1u) * 0x01010101u;

  int cslot = 0;
  for (int s = 0; s < a.S; ++s) {
    const int tile = s * a.CS + rank;
    if (s > 0) __syncthreads();                    // everyone is done with the previous table
    if (tile < a.n_tiles) {
#pragma unroll
      for (int t = 0; t < TT; ++t)
#pragma unroll
        for (int hb = 0; hb < VPL; ++hb) {
          float* lt = lut + t * TP + hb * 32;
          const T* xt = xs + ((size_t)(s * TT + t) * (TP / a.M) + hb * 32 / a.M) * D;
          if (a.M == 1) build_lut_32<T, D, 1>(lt, cbs, xt, a.E, warp, lane);
          else if (a.M == 2) build_lut_32<T, D, 2>(lt, cbs, xt, a.E, warp, lane);
          else build_lut_32<T, D, 4>(lt, cbs, xt, a.E, warp, lane);
        }
    }
    __syncthreads();
    if (s == 0) trace_mark(a, 4);
    const int q0 = tile * TP + sub * 16;
    for (int it = 0; it < n_iter; ++it) {
      cp_async_wait<kRing - 1>();                  // this thread's vectors of item (s, it) have landed
      float acc[2 * TT];
#pragma unroll
      for (int t = 0; t < 2 * TT; ++t) acc[t] = 0.f;
      if (it < my_iters && tile < a.n_tiles) {
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          if (q0 + 32 * v < a.Gp) {
            uint4 cv;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(cv.x), "=r"(cv.y), "=r"(cv.z), "=r"(cv.w)
                         : "r"(my_slot0 + cslot * kRingChunk + kRingVec * v));
            uint32_t u[4];
            rotate_bytes(cv, rw, emask4, u);
#pragma unroll
            for (int k = 0; k < 16; ++k) {
              // one PRMT: (code << 8) | column offset - the whole table offset (row stride 256 B)
              const uint32_t off = __byte_perm(col[k], u[k >> 2], 0x1140u + ((k & 3) << 4));
              if constexpr (TT == 1) {
                if (v == 0) acc[k & 1] += lds_f32<0>(lut_base + off);        // two add chains
                else acc[k & 1] += lds_f32<128>(lut_base + off);
              } else {
                acc[(k & 1) * 2] += lds_f32<0>(lut_base + off);
                acc[(k & 1) * 2 + 1] += lds_f32<TP * 4>(lut_base + off);
              }
            }
          }
        }
      }
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        float r = acc[t] + acc[TT + t];
        r += __shfl_xor_sync(0xffffffffu, r, 1);
        if (sub == 0 && it < my_iters) rowsum[t * a.rows_per_cluster + it * RPI + row_in_iter] += r;
      }
      issue();
      cslot = (cslot + 1 == kRing) ? 0 : cslot + 1;
    }
  }
  __syncthreads();
  trace_mark(a, 5);

  // ---- split-K over the cluster: every CTA pushes its partial sums of rows [o, o+1) * chunk into
  // CTA o's receive buffer (distributed shared memory), ONE cluster barrier, then each CTA adds
  // the CS partials of its own rows in rank order (deterministic), scales and stores ----
  const float* fin = rowsum;                       // CS == 1: the sums are already final
  int fin_stride = a.rows_per_cluster, fin_parts = 1, part_stride = 0, fin_off = 0;
  if (a.CS > 1) {
    const uint32_t recv_addr = smem_u32(recv);
    for (int i = tid; i < rows; i += kLutThreads) {
      const int o = i / chunk, j = i - o * chunk;
#pragma unroll
      for (int t = 0; t < TT; ++t) {
        uint32_t ra;
        const uint32_t la = recv_addr + (uint32_t)((rank * TT + t) * chunk + j) * 4u;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(o));
        asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(ra), "f"(rowsum[t * a.rows_per_cluster + i]) : "memory");
      }
    }
    cluster_sync_all();
    fin = recv; fin_stride = chunk; fin_parts = a.CS; part_stride = TT * chunk; fin_off = -fin0;
  }
  trace_mark(a, 6);
  for (int i = fin0 + tid; i < fin1; i += kLutThreads) {
    const int row = rb.row0 + i;
    float sc = Traits<T>::to_f(*reinterpret_cast<const T*>(&sc0_raw));
    float b = Traits<T>::to_f(*reinterpret_cast<const T*>(&bi0_raw));
    if (i != fin0 + tid) {                         // chunk > blockDim: later rows fetch theirs here
      sc = Traits<T>::to_f(reinterpret_cast<const T*>(a.scales)[row]);
      b = a.bias ? Traits<T>::to_f(reinterpret_cast<const T*>(a.bias)[row]) : 0.f;
    }
#pragma unroll
    for (int t = 0; t < TT; ++t) {
      const int tok = z * TT + t;
      float sum = 0.f;
      for (int r = 0; r < fin_parts; ++r) sum += fin[r * part_stride + t * fin_stride + i + fin_off];
      if (tok < a.T)
        reinterpret_cast<T*>(a.y)[(size_t)tok * a.y_stride + row] = Traits<T>::from_f(fmaf(sum, sc, b));
    }
  }
  trace_mark(a, 7);
}

// ------------------------------------------------------------------ host side
extern bool g_use_pdl;
unsigned long long* g_lut_trace = nullptr;
int g_lut_last_grid[4] = {0, 0, 0, 0};           // grid.x, grid.y, cluster size, shared memory of the last launch

static int lut_token_tile(int64_t t) { return t >= 2 ? 2 : 1; }

struct LutPlan {
  int tt, tp, n_tiles, cs, S, rows_per_cluster, z_tiles;
  int smem_cb, smem_x, smem_total;
};

static size_t lut_smem_bytes(int E, int smem_cb, int smem_x, int rows_per_cluster, int tt) {
  return (size_t)E * kLutRowFloats * 4 + (size_t)kRing * kRingChunk + smem_cb + smem_x +
         2 * (size_t)(rows_per_cluster + kLutMaxCluster) * tt * 4;   // own sums + receive buffer
}

int lut_max_active_clusters(int cs, int smem);

// Co-resident clusters of `cs` CTAs at `smem` bytes per CTA (cudaOccupancyMaxActiveClusters, cached;
// GPC granularity strands a few SMs for cs >= 4).  Without a device: B200 figures.
static int lut_cluster_slots(int cs, int smem) {
  static int cache[16][4][32] = {};               // device, log2(cs), smem / 8 KiB
  int dev = 0;
  const bool have_dev = cudaGetDevice(&dev) == cudaSuccess && dev < 16;
  const int l2 = cs == 1 ? 0 : cs == 2 ? 1 : cs == 4 ? 2 : 3, sb = min(31, smem / 8192);
  if (have_dev && cache[dev][l2][sb] > 0) return cache[dev][l2][sb];
  int n = have_dev ? lut_max_active_clusters(cs, (sb + 1) * 8192) : -1;
  if (n <= 0) {
    cudaGetLastError();
    const int per_sm = smem <= 112 * 1024 ? 2 : 1;
    static const int one[4] = {148, 74, 33, 15}, two[4] = {296, 148, 71, 33};
    n = per_sm == 2 ? two[l2] : one[l2];
  }
  if (have_dev) cache[dev][l2][sb] = n;
  return n;
}

// Pick the cluster size and the rows per cluster with a small time model calibrated on B200 phase
// traces (tools/trace_lut.py): a CTA pays a fixed cost, one table build per tile and one lookup
// pass per (tile, row iteration); two CTAs on one SM slow each other down.
static bool lut_plan(LutPlan& pl, int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts) {
  const int E = 1 << nbits;
  pl.tt = lut_token_tile(t);
  pl.tp = kLutRowFloats / pl.tt;
  pl.z_tiles = (int)((t + pl.tt - 1) / pl.tt);
  const int64_t gp = k / d * m;
  pl.n_tiles = (int)((gp + pl.tp - 1) / pl.tp);
  pl.smem_cb = m * E * d * 2;
  const int rpi = 16 * kLutWarps;
  const int sms = sm_count();
  static const int force_cs = [] { const char* e = getenv("PQ_LUT_CS"); return e ? atoi(e) : 0; }();
  static const int force_occ = [] { const char* e = getenv("PQ_LUT_OCC"); return e ? atoi(e) : 0; }();
  double best = 1e30;
  bool found = false;
  for (int cs = 1; cs <= kLutMaxCluster; cs *= 2) {
    if (cs > pl.n_tiles && cs > 1) break;
    if (force_cs && cs != force_cs) continue;      // tuning knob
    const int S = (pl.n_tiles + cs - 1) / cs;
    const int smem_x = S * pl.tt * (pl.tp / m) * d * 2;
    if (smem_x > kLutMaxX) continue;
    for (int occ = 1; occ <= 2; ++occ) {           // CTAs of this launch per SM
      if (force_occ && occ != force_occ) continue;
      int slots = lut_cluster_slots(cs, occ == 2 ? 100 * 1024 : 120 * 1024);
      if (occ == 1 && slots * cs > sms) slots = sms / cs;
      int clusters = slots / pl.z_tiles;
      if (clusters < 1) clusters = 1;
      int64_t rows = (n + clusters - 1) / clusters;
      rows = (rows + rpi - 1) / rpi * rpi;
      const int64_t max_rows = kLutMaxRowsum / (8 * pl.tt) / rpi * rpi;
      if (rows > max_rows) rows = max_rows;
      const int smem = (int)lut_smem_bytes(E, pl.smem_cb, smem_x, (int)rows, pl.tt);
      if (smem > (occ == 2 ? 112 : 200) * 1024) continue;
      const int64_t blocks = (n + rows - 1) / rows + (n_parts - 1);   // partition-aligned blocks
      const int64_t waves = (blocks * pl.z_tiles + slots - 1) / slots;
      const double slow = (occ == 2 && blocks * pl.z_tiles * cs > sms) ? 1.6 : 1.0;
      const double build = 0.9 * pl.tt * (E / 256.0) * slow + 0.1;
      const double item = 0.30 * pl.tt * slow;
      const double cta = 1.7 + S * (build + (double)(rows / rpi) * item) + (cs > 1 ? 0.8 : 0.2);
      const double cost = (double)waves * cta;
      if (cost < best) {
        best = cost; found = true;
        pl.cs = cs; pl.S = S; pl.rows_per_cluster = (int)rows; pl.smem_x = smem_x; pl.smem_total = smem;
      }
    }
  }
  return found;
}

bool gemv_lut_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (nbits > 8 || nbits < 4 || t > 8 || t < 1) return false;       // E >= 16: whole 16-entry MMA blocks
  if (d != 4 && d != 8 && d != 16) return false;
  if (m != 1 && m != 2 && m != 4) return false;
  if (k % d) return false;
  const int64_t gp = k / d * m;
  if (gp % 16) return false;                        // 16-byte code vectors, 16-byte aligned rows
  LutPlan pl;
  return lut_plan(pl, t, n, k, d, nbits, m, PQ_MAX_PARTS);
}

// PQ_ALGO_AUTO heuristic: tiny problems are better served by the direct kernel.
bool gemv_lut_preferred(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  return gemv_lut_supported(t, n, k, d, nbits, m) && n * (k / d * m) >= (1 << 18);
}

// The table kernel needs no global workspace (split-K runs over distributed shared memory).
int64_t gemv_lut_workspace(int64_t, int64_t, int64_t, int, int, int, int) { return 0; }

template <typename T, int D, int TT>
static int launch_lut(LutArgs& a, const LutPlan& pl, cudaStream_t stream) {
  auto kern = pq_gemv_lut_kernel<T, D, TT>;
  static bool attr_set[64] = {};                   // per instantiation, per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr_set[dev & 63] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(a.parts.cta_end[a.parts.n_parts - 1] * pl.cs, pl.z_tiles, 1);
  cfg.blockDim = dim3(kLutThreads);
  cfg.dynamicSmemBytes = pl.smem_total;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = pl.cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 2 : 1;
  g_lut_last_grid[0] = cfg.gridDim.x; g_lut_last_grid[1] = cfg.gridDim.y;
  g_lut_last_grid[2] = pl.cs; g_lut_last_grid[3] = pl.smem_total;
  return (int)cudaLaunchKernelEx(&cfg, kern, a);
}

// Co-resident clusters of size cs for the fp16 d=8 T=1 kernel at `smem` bytes per CTA.
int lut_max_active_clusters(int cs, int smem) {
  auto kern = pq_gemv_lut_kernel<__half, 8, 1>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) return -1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(cs * 1024, 1, 1); cfg.blockDim = dim3(kLutThreads); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { cudaGetLastError(); return -1; }
  return n;
}

template <typename T, int D>
static int launch_lut_tt(LutArgs& a, const LutPlan& pl, cudaStream_t stream) {
  return pl.tt == 1 ? launch_lut<T, D, 1>(a, pl, stream) : launch_lut<T, D, 2>(a, pl, stream);
}

template <typename T>
static int launch_lut_d(LutArgs& a, int d, const LutPlan& pl, cudaStream_t stream) {
  switch (d) {
    case 4: return launch_lut_tt<T, 4>(a, pl, stream);
    case 8: return launch_lut_tt<T, 8>(a, pl, stream);
    default: return launch_lut_tt<T, 16>(a, pl, stream);
  }
}

int gemv_lut_dispatch(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                      const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                      int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                      int n_parts, int dtype, void* workspace, int64_t workspace_bytes,
                      cudaStream_t stream) {
  (void)workspace; (void)workspace_bytes;
  if ((uintptr_t)codes % 16 || (uintptr_t)x % 4 || (x_stride * 2) % 4) return PQ_ERR_ALIGN;
  LutPlan pl;
  if (!lut_plan(pl, t, n, k, d, nbits, m, n_parts)) return PQ_ERR_UNSUPPORTED;
  LutArgs a{};
  a.x = x; a.codes = (const unsigned char*)codes; a.codebooks = codebooks; a.scales = scales; a.bias = bias;
  a.y = y; a.x_stride = x_stride; a.y_stride = y_stride;
  a.T = (int)t; a.N = (int)n; a.K = (int)k; a.E = 1 << nbits; a.M = m;
  a.Gp = (int)(k / d * m); a.n_tiles = pl.n_tiles; a.S = pl.S; a.CS = pl.cs;
  a.rows_per_cluster = pl.rows_per_cluster; a.smem_cb = pl.smem_cb; a.smem_x = pl.smem_x;
  a.trace = g_lut_trace;
  int rc = make_parts(a.parts, part_sizes, n_parts, n, pl.rows_per_cluster);
  if (rc) return rc;
  return dtype == PQ_F16 ? launch_lut_d<__half>(a, d, pl, stream)
                         : launch_lut_d<__nv_bfloat16>(a, d, pl, stream);
}

}  // namespace pq
