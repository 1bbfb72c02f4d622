// extern "C" surface of libpq_b200.so (see include/pq_b200.h).
// Validates arguments, picks the kernel family, launches on the caller's stream.
#include <cstdlib>

#include "pq_common.cuh"

namespace pq {
int dequant_dispatch(const void*, const void*, const void*, void*, int64_t, int64_t, int, int, int,
                     const int64_t*, int, int, cudaStream_t);
int gemv_direct_dispatch(const void*, int64_t, const void*, const void*, const void*, const void*,
                         void*, int64_t, int64_t, int64_t, int64_t, int, int, int, const int64_t*,
                         int, int, cudaStream_t);
int gemv_lut_dispatch(const void*, int64_t, const void*, const void*, const void*, const void*,
                      void*, int64_t, int64_t, int64_t, int64_t, int, int, int, const int64_t*,
                      int, int, void*, int64_t, cudaStream_t);
int64_t gemv_lut_workspace(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts);
bool gemv_lut_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
bool gemv_lut_preferred(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
int gemv_lut2_dispatch(const void*, int64_t, const void*, const void*, const void*, const void*,
                       void*, int64_t, int64_t, int64_t, int64_t, int, int, int, const int64_t*,
                       int, int, void*, int64_t, cudaStream_t, int xact, const pq_b200_epilogue_t* ep);
int64_t gemv_lut2_grid_x(int64_t, int64_t, int64_t, int, int, int, const int64_t*, int, int64_t);
int embedding_dispatch(const long long*, int64_t, const void*, const void*, const void*, void*, int64_t, int64_t,
                       int64_t, int, int, int, int, cudaStream_t);
int add_rms_norm_dispatch(void*, int64_t, void*, int64_t, const void*, float, int64_t, int64_t, int, cudaStream_t);
int silu_and_mul_dispatch(const void*, int64_t, void*, int64_t, int64_t, int64_t, int, cudaStream_t);
int64_t gemv_lut2_workspace(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts);
bool gemv_lut2_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
bool gemv_lut2_preferred(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
bool gemv_lut2_fused_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
int gemm_tc_dispatch(const void*, int64_t, const void*, const void*, const void*, const void*,
                     void*, int64_t, int64_t, int64_t, int64_t, int, int, int, const int64_t*,
                     int, int, void*, int64_t, cudaStream_t);
int64_t gemm_tc_workspace(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts);
bool gemm_tc_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
int gemm_mma_dispatch(const void*, int64_t, const void*, const void*, const void*, const void*,
                      void*, int64_t, int64_t, int64_t, int64_t, int, int, int, const int64_t*,
                      int, int, cudaStream_t);
bool gemm_mma_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
int allreduce_oneshot_dispatch(const void*, void*, int64_t, int, void* const*, void* const*, int, int,
                               int64_t, cudaStream_t);
int allreduce_twostage_dispatch(const void*, void*, int64_t, int, void* const*, void* const*, void* const*, int, int,
                                int64_t, cudaStream_t);
extern bool g_use_pdl;
extern unsigned long long* g_lut_trace;
extern int g_lut_last_grid[4];
extern int g_mm_last_plan[8];
int lut_max_active_clusters(int cs, int smem);
int gemv_lut2_debug_plan(int64_t, int64_t, int64_t, int, int, int, const int64_t*, int, int*);
int gemv_lut2_debug_slots(int cw, int cs, int smem);
}  // namespace pq

static int check_format(int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                        int n_parts, int dtype) {
  if (dtype != PQ_F16 && dtype != PQ_BF16) return PQ_ERR_DTYPE;
  if (nbits < 1 || nbits > 16 || d < 1 || d > 64 || m < 1 || m > 16) return PQ_ERR_FORMAT;
  if (n <= 0 || k <= 0 || k % d != 0 || n > INT32_MAX || k > INT32_MAX) return PQ_ERR_SHAPE;
  if (part_sizes == nullptr) return PQ_ERR_NULL;
  if (n_parts < 1 || n_parts > PQ_MAX_PARTS) return PQ_ERR_SHAPE;
  int64_t s = 0;
  for (int i = 0; i < n_parts; ++i) {
    if (part_sizes[i] <= 0) return PQ_ERR_SHAPE;
    s += part_sizes[i];
  }
  return s == n ? PQ_OK : PQ_ERR_SHAPE;
}

extern "C" {

int pq_b200_abi_version(void) { return PQ_B200_ABI_VERSION; }

const char* pq_b200_strerror(int code) {
  switch (code) {
    case PQ_OK: return "ok";
    case PQ_ERR_DTYPE: return "dtype must be PQ_F16 or PQ_BF16";
    case PQ_ERR_FORMAT: return "PQ format (nbits, d, m) not supported";
    case PQ_ERR_SHAPE: return "bad shape: k % d != 0, non-positive size, or partition sizes do not sum to n";
    case PQ_ERR_ALIGN: return "pointer or stride not aligned to 16 bytes";
    case PQ_ERR_WORKSPACE: return "workspace too small";
    case PQ_ERR_NULL: return "required pointer is NULL";
    case PQ_ERR_UNSUPPORTED: return "no kernel for this request";
    case PQ_ERR_DEVICE: return "device is not sm_100";
    case PQ_ERR_STATIC_SCRATCH:
      return "problem needs more scratch than the static array of pq_b200_gemv_static holds "
             "(PQ_STATIC_SCRATCH_BYTES); use pq_b200_gemm with a caller workspace";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown error";
  }
}

void pq_b200_set_pdl(int enabled) { pq::g_use_pdl = enabled != 0; }

void pq_b200_debug_set_trace(void* device_buffer) { pq::g_lut_trace = (unsigned long long*)device_buffer; }
void pq_b200_debug_last_launch(int* out4) { for (int i = 0; i < 4; ++i) out4[i] = pq::g_lut_last_grid[i]; }
void pq_b200_debug_last_mma_plan(int* out8) { for (int i = 0; i < 8; ++i) out8[i] = pq::g_mm_last_plan[i]; }
int pq_b200_debug_lut_tma_plan(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                               int n_parts, int* out9) {
  return pq::gemv_lut2_debug_plan(t, n, k, d, nbits, m, part_sizes, n_parts, out9);
}
int pq_b200_debug_lut_tma_slots(int consumer_warps, int cluster_size, int smem_bytes) {
  return pq::gemv_lut2_debug_slots(consumer_warps, cluster_size, smem_bytes);
}
int pq_b200_debug_max_active_clusters(int cluster_size, int smem_bytes) {
  return pq::lut_max_active_clusters(cluster_size, smem_bytes);
}

int pq_b200_dequant(const void* codes, const void* codebooks, const void* scales, void* w_out,
                    int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                    int n_parts, int dtype, void* stream) {
  int rc = check_format(n, k, d, nbits, m, part_sizes, n_parts, dtype);
  if (rc) return rc;
  if (!codes || !codebooks || !w_out) return PQ_ERR_NULL;
  if (((uintptr_t)codebooks | (uintptr_t)w_out) & 15) return PQ_ERR_ALIGN;
  return pq::dequant_dispatch(codes, codebooks, scales, w_out, n, k, d, nbits, m, part_sizes, n_parts,
                              dtype, (cudaStream_t)stream);
}

static int pick_algo_impl(int64_t t, int64_t n, int64_t k, int d, int nbits, int m, bool allow_lut_tma) {
  // Crossovers measured on B200 (profiles/r1_03_*): the register-fragment mma.sync kernel wins for
  // 2..16 tokens everywhere and at 1 token / 32-64 tokens on the small-N layers (o_proj, down_proj),
  // where the table GEMV's per-CTA table build and the tcgen05 kernel's per-tile dequant are not
  // amortised.  Tuning knobs: PQ_T1_ALGO forces the batch-1 kernel, PQ_MMA_MAX_T caps the mma range.
  static const int t1_algo = [] { const char* e = getenv("PQ_T1_ALGO"); return e ? atoi(e) : 0; }();
  static const int mma_max_t = [] { const char* e = getenv("PQ_MMA_MAX_T"); return e ? atoi(e) : 64; }();
  if (t == 1 && t1_algo && (allow_lut_tma || t1_algo != PQ_ALGO_GEMV_LUT_TMA) &&
      pq_b200_algo_supported(t1_algo, t, n, k, d, nbits, m)) return t1_algo;
  // one token: the TMA-fed table GEMV wins on every Llama projection (profiles/r1_13_microbench_lut2.jsonl)
  if (allow_lut_tma && t == 1 && pq::gemv_lut2_preferred(t, n, k, d, nbits, m)) return PQ_ALGO_GEMV_LUT_TMA;
  const bool mma_ok = t <= mma_max_t && pq::gemm_mma_supported(t, n, k, d, nbits, m);
  const bool tc_ok = pq::gemm_tc_supported(t, n, k, d, nbits, m);
  // Crossovers re-measured in r2 with the split-K tcgen05 kernel (profiles/r2_24, r2_25):
  //  * 8-bit codes: the mma.sync kernel up to 16 tokens; one codebook keeps it up to 32 tokens on the
  //    small-N layers (o_proj, down_proj); beyond that the tcgen05 kernel wins everywhere;
  //  * 16-bit codes (L2-resident 65536-entry books): the gather GEMV for one token and for 2..7
  //    tokens on layers below 2^27 weights, the tcgen05 kernel otherwise - it gathers each codeword
  //    once per 128-row tile instead of once per token.
  if (mma_ok) {
    if (t <= 16) return PQ_ALGO_GEMM_MMA;
    if (t <= 32 && m == 1 && n < 8192) return PQ_ALGO_GEMM_MMA;
    if (!tc_ok) return PQ_ALGO_GEMM_MMA;
  }
  if (tc_ok) {
    if (nbits > 8 && (t >= 8 || (t >= 2 && n * k >= ((int64_t)1 << 27)))) return PQ_ALGO_GEMM_TC;
    if (nbits <= 8 && t > 8) return PQ_ALGO_GEMM_TC;
  }
  if (pq::gemv_lut_preferred(t, n, k, d, nbits, m)) return PQ_ALGO_GEMV_LUT;
  if (mma_ok) return PQ_ALGO_GEMM_MMA;
  return PQ_ALGO_GEMV_DIRECT;
}

int pq_b200_pick_algo(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  return pick_algo_impl(t, n, k, d, nbits, m, true);
}

int pq_b200_algo_supported(int algo, int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (t < 1 || n < 1 || k < 1 || d < 1 || k % d || nbits < 1 || nbits > 16 || m < 1) return 0;
  switch (algo) {
    case PQ_ALGO_AUTO: case PQ_ALGO_GEMV_DIRECT: return 1;
    case PQ_ALGO_GEMV_LUT: return pq::gemv_lut_supported(t, n, k, d, nbits, m) ? 1 : 0;
    case PQ_ALGO_GEMM_TC: return pq::gemm_tc_supported(t, n, k, d, nbits, m) ? 1 : 0;
    case PQ_ALGO_GEMM_MMA: return pq::gemm_mma_supported(t, n, k, d, nbits, m) ? 1 : 0;
    case PQ_ALGO_GEMV_LUT_TMA: return pq::gemv_lut2_supported(t, n, k, d, nbits, m) ? 1 : 0;
    default: return 0;
  }
}

int64_t pq_b200_workspace_bytes(int algo, int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                                int n_parts) {
  if (algo == PQ_ALGO_AUTO) algo = pq_b200_pick_algo(t, n, k, d, nbits, m);
  if (algo == PQ_ALGO_GEMV_LUT) return pq::gemv_lut_workspace(t, n, k, d, nbits, m, n_parts);
  if (algo == PQ_ALGO_GEMM_TC) return pq::gemm_tc_workspace(t, n, k, d, nbits, m, n_parts);
  if (algo == PQ_ALGO_GEMV_LUT_TMA) return pq::gemv_lut2_workspace(t, n, k, d, nbits, m, n_parts);
  return 0;
}

}  // extern "C"

// xact = 1: x is [t, 2k] (gate | up) and the layer's input is SiLU(gate) * up.  The TMA table GEMV
// takes the activation on its x load; every other kernel gets it from a silu_and_mul launch into
// the head of the workspace.
static int64_t act_scratch_bytes(int64_t t, int64_t k) { return (t * k * 2 + 255) / 256 * 256; }
// Offset of the activations in the workspace: behind the kernel's own part and never inside the
// first 64 KiB, which hold the split-K tickets that every launch leaves at zero for the next
// one - a caller may hand the same buffer to every call.
static int64_t act_offset(int algo, int64_t t, int64_t n, int64_t k, int d, int nbits, int m, int n_parts) {
  const int64_t own = (pq_b200_workspace_bytes(algo, t, n, k, d, nbits, m, n_parts) + 255) / 256 * 256;
  return own > (64 << 10) ? own : (64 << 10);
}

static int gemm_impl(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                     const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                     int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                     int n_parts, int dtype, int algo, void* workspace, int64_t workspace_bytes,
                     void* stream, int xact) {
  int rc = check_format(n, k, d, nbits, m, part_sizes, n_parts, dtype);
  if (rc) return rc;
  if (!x || !codes || !codebooks || !scales || !y) return PQ_ERR_NULL;
  if (t <= 0 || t > INT32_MAX || x_stride < (xact ? 2 * k : k) || y_stride < n) return PQ_ERR_SHAPE;
  if (((uintptr_t)codebooks) & 15) return PQ_ERR_ALIGN;
  const bool was_auto = algo == PQ_ALGO_AUTO;
  if (was_auto) algo = pq_b200_pick_algo(t, n, k, d, nbits, m);
  cudaStream_t s = (cudaStream_t)stream;
  if (xact && !(algo == PQ_ALGO_GEMV_LUT_TMA && pq::gemv_lut2_supported(t, n, k, d, nbits, m))) {
    const int64_t own = act_offset(algo, t, n, k, d, nbits, m, n_parts);
    if (!workspace || workspace_bytes < own + act_scratch_bytes(t, k) || ((uintptr_t)workspace & 15))
      return PQ_ERR_WORKSPACE;
    void* act = (char*)workspace + own;
    rc = pq::silu_and_mul_dispatch(x, x_stride, act, k, t, k, dtype, s);
    if (rc) return rc;
    return gemm_impl(act, k, codes, codebooks, scales, bias, y, y_stride, t, n, k, d, nbits, m,
                     part_sizes, n_parts, dtype, algo, workspace, own, stream, 0);
  }
  if (was_auto && algo == PQ_ALGO_GEMV_LUT_TMA) {
    // conditions only the dispatcher can see (alignment of this call's pointers, the plan for these
    // partition sizes): on PQ_ERR_UNSUPPORTED fall back to the other kernels
    rc = pq::gemv_lut2_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d,
                                nbits, m, part_sizes, n_parts, dtype, workspace, workspace_bytes, s, xact, nullptr);
    if (rc != PQ_ERR_UNSUPPORTED) return rc;
    algo = pick_algo_impl(t, n, k, d, nbits, m, false);
    if (xact)      // the dispatcher refused (partition sizes): activation as its own launch
      return gemm_impl(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d, nbits, m,
                       part_sizes, n_parts, dtype, algo, workspace, workspace_bytes, stream, 1);
  }
  switch (algo) {
    case PQ_ALGO_GEMV_DIRECT:
      return pq::gemv_direct_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k,
                                      d, nbits, m, part_sizes, n_parts, dtype, s);
    case PQ_ALGO_GEMV_LUT:
      if (!pq::gemv_lut_supported(t, n, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
      return pq::gemv_lut_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d,
                                   nbits, m, part_sizes, n_parts, dtype, workspace, workspace_bytes, s);
    case PQ_ALGO_GEMM_TC:
      if (!pq::gemm_tc_supported(t, n, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
      // the workspace is optional (include/pq_b200.h): without it a few-tile problem runs unsplit
      return pq::gemm_tc_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d,
                                  nbits, m, part_sizes, n_parts, dtype, workspace, workspace_bytes, s);
    case PQ_ALGO_GEMV_LUT_TMA:
      if (!pq::gemv_lut2_supported(t, n, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
      return pq::gemv_lut2_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d,
                                    nbits, m, part_sizes, n_parts, dtype, workspace, workspace_bytes, s, xact, nullptr);
    case PQ_ALGO_GEMM_MMA:
      if (!pq::gemm_mma_supported(t, n, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
      return pq::gemm_mma_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d,
                                   nbits, m, part_sizes, n_parts, dtype, s);
    default:
      return PQ_ERR_UNSUPPORTED;
  }
}

extern "C" {

int pq_b200_gemm(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                 const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                 int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                 int n_parts, int dtype, int algo, void* workspace, int64_t workspace_bytes,
                 void* stream) {
  return gemm_impl(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d, nbits, m,
                   part_sizes, n_parts, dtype, algo, workspace, workspace_bytes, stream, 0);
}

int pq_b200_gemm_silu_mul(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                          const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                          int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                          int n_parts, int dtype, int algo, void* workspace, int64_t workspace_bytes,
                          void* stream) {
  return gemm_impl(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d, nbits, m,
                   part_sizes, n_parts, dtype, algo, workspace, workspace_bytes, stream, 1);
}

int64_t pq_b200_gemm_silu_mul_workspace_bytes(int algo, int64_t t, int64_t n, int64_t k, int d, int nbits,
                                              int m, int n_parts) {
  if (algo == PQ_ALGO_AUTO) algo = pq_b200_pick_algo(t, n, k, d, nbits, m);
  return act_scratch_bytes(t, k) + act_offset(algo, t, n, k, d, nbits, m, n_parts);
}

int pq_b200_silu_and_mul(const void* x, int64_t x_stride, void* out, int64_t out_stride, int64_t t,
                         int64_t k, int dtype, void* stream) {
  if (dtype != PQ_F16 && dtype != PQ_BF16) return PQ_ERR_DTYPE;
  if (!x || !out) return PQ_ERR_NULL;
  if (t <= 0 || k <= 0 || x_stride < 2 * k || out_stride < k) return PQ_ERR_SHAPE;
  return pq::silu_and_mul_dispatch(x, x_stride, out, out_stride, t, k, dtype, (cudaStream_t)stream);
}

int pq_b200_gemm_fused(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                       const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                       int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                       int n_parts, int dtype, void* workspace, int64_t workspace_bytes, void* stream,
                       const pq_b200_epilogue_t* ep) {
  int rc = check_format(n, k, d, nbits, m, part_sizes, n_parts, dtype);
  if (rc) return rc;
  if (!x || !codes || !codebooks || !scales || !y) return PQ_ERR_NULL;
  const int xact = (ep && ep->input_silu_mul) ? 1 : 0;
  if (t <= 0 || t > INT32_MAX || x_stride < (xact ? 2 * k : k) || y_stride < n) return PQ_ERR_SHAPE;
  if (((uintptr_t)codebooks) & 15) return PQ_ERR_ALIGN;
  if (!pq::gemv_lut2_fused_supported(t, n, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
  return pq::gemv_lut2_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d, nbits, m,
                                part_sizes, n_parts, dtype, workspace, workspace_bytes, (cudaStream_t)stream, xact, ep);
}

int pq_b200_gemm_fused_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m) {
  if (t < 1 || n < 1 || k < 1 || d < 1 || k % d || m < 1) return 0;
  return pq::gemv_lut2_fused_supported(t, n, k, d, nbits, m) ? 1 : 0;
}

int64_t pq_b200_gemm_fused_slots(int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                                 const int64_t* part_sizes, int n_parts, int64_t workspace_bytes) {
  if (check_format(n, k, d, nbits, m, part_sizes, n_parts, PQ_F16)) return PQ_ERR_SHAPE;
  if (!pq::gemv_lut2_fused_supported(t, n, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
  return pq::gemv_lut2_grid_x(t, n, k, d, nbits, m, part_sizes, n_parts, workspace_bytes);
}

int pq_b200_embedding(const int64_t* ids, int64_t n_ids, const void* codes, const void* codebooks, const void* scales,
                      void* out, int64_t out_stride, int64_t n, int64_t k, int d, int nbits, int m, int dtype,
                      void* stream) {
  const int64_t one = n;
  int rc = check_format(n, k, d, nbits, m, &one, 1, dtype);
  if (rc) return rc;
  if (!ids || !codes || !codebooks || !scales || !out) return PQ_ERR_NULL;
  if (n_ids < 0 || out_stride < k) return PQ_ERR_SHAPE;
  if (d != 4 && d != 8 && d != 16) return PQ_ERR_FORMAT;
  if ((((uintptr_t)codebooks | (uintptr_t)out) & 15) || (out_stride * 2) % 16) return PQ_ERR_ALIGN;
  if (n_ids == 0) return PQ_OK;
  return pq::embedding_dispatch(reinterpret_cast<const long long*>(ids), n_ids, codes, codebooks, scales, out,
                                out_stride, n, k, d, nbits, m, dtype, (cudaStream_t)stream);
}

int pq_b200_fused_add_rms_norm(void* input, int64_t input_stride, void* residual, int64_t residual_stride,
                               const void* weight, float eps, int64_t t, int64_t hidden, int dtype, void* stream) {
  if (dtype != PQ_F16 && dtype != PQ_BF16) return PQ_ERR_DTYPE;
  if (!input || !weight) return PQ_ERR_NULL;
  if (t <= 0 || hidden <= 0 || input_stride < hidden || (residual && residual_stride < hidden)) return PQ_ERR_SHAPE;
  return pq::add_rms_norm_dispatch(input, input_stride, residual, residual_stride, weight, eps, t, hidden, dtype,
                                   (cudaStream_t)stream);
}

int pq_b200_allreduce_oneshot(const void* in, void* out, int64_t n_elems, int dtype,
                              void* const* peer_data, void* const* peer_flags, int rank, int world,
                              int64_t slot_bytes, void* stream) {
  if (dtype != PQ_F16 && dtype != PQ_BF16) return PQ_ERR_DTYPE;
  if (!in || !out || !peer_data || !peer_flags) return PQ_ERR_NULL;
  if (world < 2 || world > PQ_AR_MAX_WORLD || rank < 0 || rank >= world || n_elems <= 0) return PQ_ERR_SHAPE;
  if ((n_elems * 2) % 16 || slot_bytes % 16 || ((uintptr_t)in | (uintptr_t)out) & 15) return PQ_ERR_ALIGN;
  if (n_elems * 2 > slot_bytes) return PQ_ERR_WORKSPACE;
  return pq::allreduce_oneshot_dispatch(in, out, n_elems, dtype, peer_data, peer_flags, rank, world, slot_bytes,
                                        (cudaStream_t)stream);
}

int pq_b200_allreduce_twostage(const void* in, void* out, int64_t n_elems, int dtype, void* const* peer_stage,
                               void* const* peer_res, void* const* peer_flags, int rank, int world,
                               int64_t max_bytes, void* stream) {
  if (dtype != PQ_F16 && dtype != PQ_BF16) return PQ_ERR_DTYPE;
  if (!in || !out || !peer_stage || !peer_res || !peer_flags) return PQ_ERR_NULL;
  if (world < 2 || world > PQ_AR_MAX_WORLD || rank < 0 || rank >= world || n_elems <= 0) return PQ_ERR_SHAPE;
  if ((n_elems * 2) % 16 || max_bytes % 256 || ((uintptr_t)in | (uintptr_t)out) & 15) return PQ_ERR_ALIGN;
  if (n_elems * 2 > max_bytes || n_elems * 2 / 16 > INT32_MAX) return PQ_ERR_WORKSPACE;
  return pq::allreduce_twostage_dispatch(in, out, n_elems, dtype, peer_stage, peer_res, peer_flags, rank, world,
                                         max_bytes, (cudaStream_t)stream);
}

}  // extern "C"

// ---- pq_b200_gemv_static: the table GEMV with its scratch in a static device array ------------------
// No counterpart in the reference (include/pq_b200.h).  The array's NAME is a host-side stand-in in host
// code, not a device address: the address is resolved once per device, with that device current, by
// cudaGetSymbolAddress, and only that pointer is handed to the dispatcher / kernel.
static __device__ __align__(256) unsigned char g_static_scratch[PQ_STATIC_SCRATCH_BYTES];

static int static_scratch_ptr(void** out) {
  static void* ptrs[64] = {};                       // per device ordinal; a benign race writes the same value
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return PQ_ERR_DEVICE;
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return PQ_ERR_DEVICE;
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return PQ_OK;
}

extern "C" int pq_b200_gemv_static(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                                   const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                                   int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                                   int n_parts, int dtype, void* stream) {
  int rc = check_format(n, k, d, nbits, m, part_sizes, n_parts, dtype);
  if (rc) return rc;
  if (!x || !codes || !codebooks || !scales || !y) return PQ_ERR_NULL;
  if (t <= 0 || t > INT32_MAX || x_stride < k || y_stride < n) return PQ_ERR_SHAPE;
  if (((uintptr_t)codebooks) & 15) return PQ_ERR_ALIGN;
  if (!pq::gemv_lut2_supported(t, n, k, d, nbits, m)) return PQ_ERR_UNSUPPORTED;
  if (pq::gemv_lut2_workspace(t, n, k, d, nbits, m, n_parts) > (int64_t)PQ_STATIC_SCRATCH_BYTES)
    return PQ_ERR_STATIC_SCRATCH;
  void* scratch = nullptr;
  rc = static_scratch_ptr(&scratch);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  // the ticket counters at the head of the scratch must start at zero, and the array keeps whatever the
  // last call (possibly an aborted or a racing one) left there; the partial sums need no clearing
  cudaError_t e = cudaMemsetAsync(scratch, 0, 64 * 1024, s);
  if (e != cudaSuccess) return (int)e;
  return pq::gemv_lut2_dispatch(x, x_stride, codes, codebooks, scales, bias, y, y_stride, t, n, k, d, nbits, m,
                                part_sizes, n_parts, dtype, scratch, (int64_t)PQ_STATIC_SCRATCH_BYTES, s, 0,
                                nullptr);
}
