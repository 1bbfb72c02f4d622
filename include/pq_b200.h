/*
 * pq_b200.h - C ABI of libpq_b200.so: the B200 (sm_100a) product-quantization
 * linear kernels.  This is the drop-in boundary for the reference's native PQ
 * (AQLM) ops.  Plain pointers and sizes only; no torch types; no allocation and
 * no synchronisation inside, so every call is CUDA-graph capturable.
 *
 * Reference interfaces replaced (paths relative to /root/reference):
 *   pq_b200_dequant  <- aqlm_dequant   csrc/ops.h:129-131, csrc/torch_bindings.cpp:169-173,
 *                                      csrc/quantization/aqlm/gemm_kernels.cu:544-597
 *   pq_b200_gemm     <- aqlm_gemm      csrc/ops.h:123-127, csrc/torch_bindings.cpp:163-167,
 *                                      csrc/quantization/aqlm/gemm_kernels.cu:519-542
 *                       and, for tokens > 6, optimized_dequantize_gemm
 *                       (vllm/model_executor/layers/quantization/aqlm.py:132-156)
 *   pq_b200_allreduce_* <- CustomAllreduce  csrc/custom_all_reduce.cuh:214-506
 *
 * Tensor layouts are the reference checkpoint layouts (aqlm.py:253-316):
 *   codes      [n, k/d, m]      int8 (nbits <= 8) or int16 (nbits <= 16), signed storage,
 *                               index = (unsigned)code mod 2^nbits      (aqlm.py:32-34)
 *   codebooks  [n_parts*m, 2^nbits, 1, d]  fp16 / bf16
 *   scales     [n] (checkpoint shape [n,1,1,1])  fp16 / bf16
 *   x          [t, k] row stride x_stride elements;  y [t, n] row stride y_stride
 * Partition p owns rows [sum(part_sizes[:p]), +part_sizes[p]) and codebooks
 * [p*m, (p+1)*m)                                               (aqlm.py:110-126).
 *
 * Every function returns 0 on success, a negative PQ_ERR_* on a bad argument,
 * or a positive cudaError_t from the launch.
 */
#ifndef PQ_B200_H_
#define PQ_B200_H_

#include <stdint.h>

#if defined(__GNUC__)
#define PQ_API __attribute__((visibility("default")))
#else
#define PQ_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define PQ_B200_ABI_VERSION 2
#define PQ_MAX_PARTS 8

enum { PQ_F16 = 0, PQ_BF16 = 1 };

enum {
  PQ_OK = 0,
  PQ_ERR_DTYPE = -1,       /* dtype not PQ_F16 / PQ_BF16 */
  PQ_ERR_FORMAT = -2,      /* (nbits, d, m) outside the supported grid */
  PQ_ERR_SHAPE = -3,       /* k % d != 0, sizes <= 0, partition sizes do not sum to n */
  PQ_ERR_ALIGN = -4,       /* a pointer or stride is not aligned as documented */
  PQ_ERR_WORKSPACE = -5,   /* workspace too small (see pq_b200_workspace_bytes) */
  PQ_ERR_NULL = -6,        /* required pointer is NULL */
  PQ_ERR_UNSUPPORTED = -7, /* valid request this build has no kernel for */
  PQ_ERR_DEVICE = -8,      /* not an sm_100 device */
  PQ_ERR_STATIC_SCRATCH = -9 /* pq_b200_gemv_static: the problem needs more than PQ_STATIC_SCRATCH_BYTES */
};

/* which kernel family pq_b200_gemm uses */
enum {
  PQ_ALGO_AUTO = 0,
  PQ_ALGO_GEMV_DIRECT = 1, /* codebook gather in shared memory / L2, CUDA cores            */
  PQ_ALGO_GEMV_LUT = 2,    /* per-token x.centroid lookup tables (nbits <= 8)               */
  PQ_ALGO_GEMM_TC = 3,     /* dequant tile -> smem -> tcgen05.mma, TMEM accumulator         */
  PQ_ALGO_GEMM_MMA = 4,    /* codeword gather straight into mma.sync fragments (t <= 64)    */
  PQ_ALGO_GEMV_LUT_TMA = 5 /* lookup tables, TMA code stream, one row per lane (nbits <= 8)  */
};

PQ_API int pq_b200_abi_version(void);
/* Launch decode kernels with programmatic dependent launch (off by default). */
PQ_API void pq_b200_set_pdl(int enabled);
PQ_API const char* pq_b200_strerror(int code);

/* Diagnostics (tools/trace_lut.py); not part of the reference boundary.
 * set_trace: device buffer of 16 x u64 per CTA that the table GEMV fills with %globaltimer phase
 * stamps (NULL = off, the default).  last_launch: {grid.x, grid.y, cluster size, smem bytes}.
 * max_active_clusters: cudaOccupancyMaxActiveClusters for the table GEMV. */
PQ_API void pq_b200_debug_set_trace(void* device_buffer);
PQ_API void pq_b200_debug_last_launch(int* out4);
/* {token blocks, row tiles per step, cluster size, warps along K, superblocks per x phase,
 * rows per cluster, grid.x, smem bytes} of the last PQ_ALGO_GEMM_MMA launch. */
PQ_API void pq_b200_debug_last_mma_plan(int* out8);
PQ_API int pq_b200_debug_max_active_clusters(int cluster_size, int smem_bytes);
/* The table GEMV's launch plan for a problem: {consumer warps per CTA, cluster size, tiles per CTA,
 * K groups, rows per cluster, ring stages, shared-memory bytes, 64-position tiles, grid.x}. */
PQ_API int pq_b200_debug_lut_tma_plan(int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                                      const int64_t* part_sizes, int n_parts, int* out9);
/* cudaOccupancyMaxActiveClusters of the TMA table GEMV's CTA shape (8 or 16 consumer warps). */
PQ_API int pq_b200_debug_lut_tma_slots(int consumer_warps, int cluster_size, int smem_bytes);

/* Unscaled (scales == NULL; what aqlm_dequant returns, gemm_kernels.cu:563-592) or
 * scaled (aqlm.py:70-72) weights w_out[n, k].  Sum over the m codebooks is
 * accumulated in fp32 in codebook order and rounded to nearest even once; the
 * scale multiply rounds once more.  Bit-exact against aqlm.py:37-74.
 * Alignment: codebooks, w_out 16 bytes. */
PQ_API int pq_b200_dequant(const void* codes, const void* codebooks, const void* scales,
                    void* w_out, int64_t n, int64_t k, int d, int nbits, int m,
                    const int64_t* part_sizes, int n_parts, int dtype, void* stream);

/* Workspace (bytes) pq_b200_gemm can use for this problem with `algo`; it is OPTIONAL: given NULL
 * (or fewer bytes) every kernel still runs - the table GEMV then keeps each row block inside one
 * cluster instead of splitting K across clusters, which costs long-K layers some speed.  The first 64 KiB of a
 * non-empty workspace hold split-K ticket counters: they must be ZERO before the first call and
 * every call leaves them zero, so one zero-initialised buffer can be reused for all calls that
 * are ordered on a stream (the rest of the workspace is scratch).  256-byte aligned. */
PQ_API int64_t pq_b200_workspace_bytes(int algo, int64_t t, int64_t n, int64_t k, int d,
                                int nbits, int m, int n_parts);

/* y[t, n] = x[t, k] . W^T * scales + bias, W as in pq_b200_dequant.
 * bias may be NULL.  workspace may be NULL when pq_b200_workspace_bytes() == 0.
 * algo: PQ_ALGO_AUTO picks by (t, format).  Accumulation is fp32; scale and bias
 * are applied in fp32 before the single rounding of y.
 * Alignment: x, y, codebooks 16 bytes; x_stride, y_stride in elements. */
PQ_API int pq_b200_gemm(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                 const void* scales, const void* bias, void* y, int64_t y_stride,
                 int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                 const int64_t* part_sizes, int n_parts, int dtype, int algo,
                 void* workspace, int64_t workspace_bytes, void* stream);

/* y = pq_linear(SiLU(x[:, :k]) * x[:, k:]): x is [t, 2k] (gate | up, row stride x_stride >= 2k),
 * everything else as pq_b200_gemm.  Replaces silu_and_mul + aqlm_gemm of the MLP's down_proj
 * (vllm/model_executor/models/llama.py:89-93, layers/activation.py:65-84,
 * csrc/activation_kernels.cu): PQ_ALGO_GEMV_LUT_TMA applies the activation while it loads x;
 * the other kernels get it from one pq_b200_silu_and_mul launch into the workspace, which must
 * then hold pq_b200_gemm_silu_mul_workspace_bytes().  Rounding follows the reference kernel:
 * SiLU in fp32 rounded to the dtype, product rounded to the dtype. */
PQ_API int pq_b200_gemm_silu_mul(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                          const void* scales, const void* bias, void* y, int64_t y_stride,
                          int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                          const int64_t* part_sizes, int n_parts, int dtype, int algo,
                          void* workspace, int64_t workspace_bytes, void* stream);
PQ_API int64_t pq_b200_gemm_silu_mul_workspace_bytes(int algo, int64_t t, int64_t n, int64_t k, int d,
                                                     int nbits, int m, int n_parts);
/* out[t, k] = SiLU(x[:, :k]) * x[:, k:]  <- silu_and_mul  csrc/activation_kernels.cu,
 * csrc/torch_bindings.cpp (silu_and_mul).  k even; 16-byte loads when k, strides and pointers allow. */
PQ_API int pq_b200_silu_and_mul(const void* x, int64_t x_stride, void* out, int64_t out_stride, int64_t t,
                                int64_t k, int dtype, void* stream);

/* out[i, :] = row ids[i] of the scaled dequantised [n, k] weight (one output partition): the embedding
 * lookup of a PQ-quantised VocabParallelEmbedding  <- QuantizeMethodBase.embedding
 * (vllm/model_executor/layers/quantization/base_config.py:30-35, layers/vocab_parallel_embedding.py:41-43,
 * 229-246,395-404).  Bit-identical to the rows of pq_b200_dequant with scales.  ids: int64 on the device;
 * an id outside [0, n) gives a zero row.  d in {4, 8, 16}.  Alignment: codebooks, out 16 bytes. */
PQ_API int pq_b200_embedding(const int64_t* ids, int64_t n_ids, const void* codes, const void* codebooks,
                             const void* scales, void* out, int64_t out_stride, int64_t n, int64_t k, int d,
                             int nbits, int m, int dtype, void* stream);

/* ---- fused epilogue / prologue of the decode table GEMV ------------------------
 * What the reference runs as separate launches around a row-parallel linear at decode
 *   apply -> tensor_model_parallel_all_reduce -> fused_add_rms_norm -> next linear
 *   (vllm/model_executor/layers/linear.py:1209-1213, csrc/custom_all_reduce.cuh:214-237,
 *    csrc/custom_all_reduce.cu:75-77, csrc/layernorm_kernels.cu:52-98, models/llama.py:272-286)
 * becomes part of the two GEMV launches it sits between:
 *   producer (o_proj / down_proj): each finished row is pushed to every peer as one 8-byte
 *     {fp32 value, epoch} packet, summed in rank order (bitwise identical on all ranks), rounded
 *     once, added to the residual, stored, and its square added to the CTA's slot of ssq_out;
 *   consumer (qkv_proj / gate_up_proj): x = RMSNorm(z) * weight on the x load, rstd from norm_ssq.
 * All fields optional (0 / NULL = off).  Buffers of the all-reduce (per rank, cudaMalloc'd, exchanged
 * by CUDA IPC like pq_b200_allreduce_oneshot's): ar_peer[r] = rank r's receive buffer of
 * 2 * ar_world * ar_slot_elems 8-byte packets, zeroed once; ar_state = this rank's 2 x uint32
 * {epoch, ticket}, zeroed once.  Every rank must issue the same sequence of fused launches. */
typedef struct {
  int32_t ar_world, ar_rank;     /* ar_world > 1: all-reduce y over the TP group                */
  void* const* ar_peer;          /* [ar_world] device pointers (peer-mapped)                     */
  void* ar_state;
  int64_t ar_slot_elems;         /* >= t * n                                                     */
  const void* residual;          /* [t, n] or NULL: y = T(T(sum) + residual)                     */
  int64_t residual_stride;
  float* ssq_out;                /* [t, ssq_slots] or NULL: per-CTA sums of y^2                  */
  int64_t ssq_slots;             /* must equal pq_b200_gemm_fused_slots() of this problem        */
  const float* norm_ssq;         /* consumer: [t, norm_slots] from the producer, or NULL         */
  const void* norm_weight;       /* [k] RMSNorm weight; non-NULL turns the norm-on-load on       */
  int32_t norm_slots;
  float norm_eps;
  int32_t input_silu_mul;        /* 1: x is [t, 2k] = (gate | up), the input is SiLU(gate) * up   */
  /* Private repacked copy of `codes` built once by process_weights_after_loading (base_config.py:37-42,
   * model_loader/loader.py:157-168): the same bytes with every row padded to codes_pitch bytes, a
   * multiple of 16, so that rows whose code bytes are not (the TP = 4 / 8 shards of Llama-2-7B's
   * down_proj: 344 / 172 bytes) can be streamed by TMA.  NULL: `codes` itself is streamed. */
  const void* codes_padded;
  int64_t codes_pitch;
} pq_b200_epilogue_t;

/* pq_b200_gemm on the table GEMV (PQ_ALGO_GEMV_LUT_TMA) with the extras above; PQ_ERR_UNSUPPORTED when
 * pq_b200_gemm_fused_supported() is 0 (t > 8, nbits != 8, ...): run the steps separately then
 * (pq_b200_gemm, pq_b200_allreduce_oneshot, pq_b200_fused_add_rms_norm). */
PQ_API int pq_b200_gemm_fused(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                              const void* scales, const void* bias, void* y, int64_t y_stride,
                              int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                              const int64_t* part_sizes, int n_parts, int dtype,
                              void* workspace, int64_t workspace_bytes, void* stream,
                              const pq_b200_epilogue_t* ep);
/* 1 if pq_b200_gemm_fused covers the problem (1..8 tokens, 8-bit codes, shapes of the table GEMV). */
PQ_API int pq_b200_gemm_fused_supported(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);
/* Floats per token of ssq_out for this problem (the CTAs that finish rows), or PQ_ERR_UNSUPPORTED. */
PQ_API int64_t pq_b200_gemm_fused_slots(int64_t t, int64_t n, int64_t k, int d, int nbits, int m,
                                        const int64_t* part_sizes, int n_parts, int64_t workspace_bytes);
/* In place: residual += input (rounded to the dtype), input = RMSNorm(residual) * weight; or, with
 * residual == NULL, input = RMSNorm(input) * weight.  <- fused_add_rms_norm / rms_norm,
 * csrc/layernorm_kernels.cu:17-98, same roundings.  One CTA per token. */
PQ_API int pq_b200_fused_add_rms_norm(void* input, int64_t input_stride, void* residual, int64_t residual_stride,
                                      const void* weight, float eps, int64_t t, int64_t hidden, int dtype,
                                      void* stream);

/* The table GEMV (PQ_ALGO_GEMV_LUT_TMA, 1..8 tokens, nbits <= 8) for callers that cannot allocate and have
 * no workspace to hand over (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arguments and result as pq_b200_gemm(..., PQ_ALGO_GEMV_LUT_TMA, workspace, PQ_STATIC_SCRATCH_BYTES, ...)
 * minus the workspace: the K-group partial sums and ticket counters live in a file-scope
 * `static __device__` array of PQ_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the
 * library holds it until exit).  The call allocates nothing - no cudaMalloc / cudaFree, no allocating
 * library routine; its only extra work is one cudaMemsetAsync of the 64 KiB ticket area on `stream`.
 *   * a problem whose pq_b200_workspace_bytes(PQ_ALGO_GEMV_LUT_TMA, ...) exceeds PQ_STATIC_SCRATCH_BYTES
 *     (t * n above ~124 K outputs) is rejected with PQ_ERR_STATIC_SCRATCH, nothing is launched;
 *   * every call on a device shares the one array, which keeps the previous call's contents: two calls
 *     that OVERLAP on the GPU (different streams, or graphs replayed concurrently) race on it - order
 *     them on one stream or with events;
 *   * the array's address is resolved on the first call on each device (cudaGetSymbolAddress, with that
 *     device current): make that first call eagerly, before capturing a graph - like every kernel of this
 *     library, whose first launch sets function attributes. */
#define PQ_STATIC_SCRATCH_BYTES (4 << 20)
PQ_API int pq_b200_gemv_static(const void* x, int64_t x_stride, const void* codes, const void* codebooks,
                               const void* scales, const void* bias, void* y, int64_t y_stride, int64_t t,
                               int64_t n, int64_t k, int d, int nbits, int m, const int64_t* part_sizes,
                               int n_parts, int dtype, void* stream);

/* 1 if kernel family `algo` covers this problem, else 0 (PQ_ALGO_AUTO: always 1). */
PQ_API int pq_b200_algo_supported(int algo, int64_t t, int64_t n, int64_t k, int d, int nbits, int m);

/* Which PQ_ALGO_* would PQ_ALGO_AUTO choose. */
PQ_API int pq_b200_pick_algo(int64_t t, int64_t n, int64_t k, int d, int nbits, int m);

/* ---- tensor-parallel reduction hook (row-parallel O / down projections) -------------
 * One-shot all-reduce over peer-mapped buffers (NVLink P2P).  The caller (host side,
 * vllm_pq_b200/distributed.py) allocates one staging buffer and one flag buffer per rank
 * with pq_b200_malloc, exchanges CUDA IPC handles and passes the world's mapped pointers:
 *   peer_data[r]  : rank r's staging buffer, 2 * world * slot_bytes bytes
 *   peer_flags[r] : rank r's flag buffer, PQ_AR_FLAG_BYTES bytes, zeroed once at start
 * n_elems * 2 <= slot_bytes; n_elems * 2 a multiple of 16; world in {2, 4, 8}.
 * Replaces cross_device_reduce_1stage (custom_all_reduce.cuh:214-230) plus the staging
 * memcpy of custom_all_reduce.cu:75-77: the kernel pushes the input straight into every
 * peer's buffer. The epoch counter lives in the flag buffer and is advanced on the device,
 * so a launch is CUDA-graph replayable. Sum order is rank 0..world-1 in fp32 on every
 * rank: results are bitwise identical across ranks (custom_all_reduce.cuh:204-212). */
#define PQ_AR_FLAG_BYTES 4096
#define PQ_AR_MAX_WORLD 8
PQ_API int pq_b200_allreduce_oneshot(const void* in, void* out, int64_t n_elems, int dtype,
                                     void* const* peer_data, void* const* peer_flags,
                                     int rank, int world, int64_t slot_bytes, void* stream);

/* Two-stage all-reduce (reduce-scatter + all-gather over peer memory) for messages above the one-shot
 * limit - the prefill partial sums of the row-parallel layers.  Replaces cross_device_reduce_2stage
 * (custom_all_reduce.cuh:239-290,478-491) and, under CUDA-graph capture, the ctypes NCCL path of
 * device_communicators/pynccl.py:109-132: one launch, device-side epoch, replayable.  Per rank (cudaMalloc'd,
 * exchanged by CUDA IPC): peer_stage[r] max_bytes; peer_res[r] 2 * round_up(max_bytes / world, 256) bytes;
 * peer_flags[r] PQ_AR2_FLAG_BYTES, zeroed once.  n_elems * 2 <= max_bytes, a multiple of 16; max_bytes a
 * multiple of 256; world in {2, 4, 8}.  Bitwise identical results on every rank. */
#define PQ_AR2_FLAG_BYTES 8192
PQ_API int pq_b200_allreduce_twostage(const void* in, void* out, int64_t n_elems, int dtype,
                                      void* const* peer_stage, void* const* peer_res, void* const* peer_flags,
                                      int rank, int world, int64_t max_bytes, void* stream);

/* Device buffers whose base address is what cudaIpcGetMemHandle maps (torch's caching
 * allocator sub-allocates, so its pointers cannot be shared this way), and the 64-byte
 * cudaIpcMemHandle_t exchange (reference: custom_all_reduce.py:178-201). */
PQ_API int pq_b200_malloc(int64_t bytes, void** out);
PQ_API int pq_b200_free(void* p);
PQ_API int pq_b200_memset_async(void* p, int value, int64_t bytes, void* stream);
PQ_API int pq_b200_ipc_get_handle(void* ptr, void* handle64);
PQ_API int pq_b200_ipc_open_handle(void* handle64, void** out);
PQ_API int pq_b200_ipc_close_handle(void* p);

#ifdef __cplusplus
}
#endif
#endif /* PQ_B200_H_ */
