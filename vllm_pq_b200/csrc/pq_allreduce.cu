// One-shot all-reduce over NVLink peer memory: the tensor-parallel reduction hook
// after the row-parallel PQ linears (o_proj, down_proj).
//
// Replaces cross_device_reduce_1stage + multi_gpu_barrier + the eager-mode staging
// memcpy of the reference (csrc/custom_all_reduce.cuh:178-230, custom_all_reduce.cu:75-77):
//   push   every rank writes its vector straight into slot [rank] of every peer's staging
//          buffer with 128-bit stores (no separate memcpy, no registration of the input)
//   signal per-block flag, st.release.sys, value = epoch
//   wait   ld.acquire.sys on this rank's flags until all peers show the epoch
//   sum    local read of world slots, fp32 accumulate in rank order 0..world-1 on every
//          rank -> bitwise identical results (custom_all_reduce.cuh:204-212)
// The epoch lives in device memory and is advanced by the kernel itself, so a launch
// carries no host-side state and can be replayed from a CUDA graph. Two staging slots
// alternate by epoch parity: a rank can only run one epoch ahead of its slowest peer
// (it needs that peer's flag to finish), so the slot it overwrites is never being read.
//
// Decode messages are 8-16 KiB: latency-bound, one CTA per 8 KiB.
#include <string.h>

#include "pq_common.cuh"

namespace pq {

extern bool g_use_pdl;

constexpr int kArThreads = 512;
constexpr int kArMaxBlocks = 36;
// flag buffer layout (uint32 words): [block * world + peer] ... then
constexpr int kArCounterWord = 1000;   // this rank's epoch counter
constexpr int kArTicketWord = 1001;    // block-completion ticket

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

struct ArPeers {
  unsigned char* data[PQ_AR_MAX_WORLD];
  uint32_t* flags[PQ_AR_MAX_WORLD];
};

template <typename T, int WORLD>
__global__ void __launch_bounds__(kArThreads)
pq_allreduce_oneshot_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int n_vec,
                            ArPeers peers, int rank, long long slot_bytes) {
  // Programmatic dependent launch: the next layer's GEMV may start its weight prefetch now; this
  // kernel itself must not touch anything (not even its epoch counter, which the previous
  // all-reduce two launches back advances at its very end) before the GEMV it follows is complete.
  pdl_launch_dependents();
  pdl_wait();
  uint32_t* my_flags = peers.flags[rank];
  const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(my_flags + kArCounterWord) + 1u;
  const long long slot_off = (long long)(epoch & 1u) * WORLD * slot_bytes;
  const int per_block = (n_vec + gridDim.x - 1) / gridDim.x;
  const int v0 = blockIdx.x * per_block;
  const int v1 = min(v0 + per_block, n_vec);

  // push this block's range into every peer (self included: a local copy)
  for (int v = v0 + threadIdx.x; v < v1; v += kArThreads) {
    const uint4 val = in[v];
#pragma unroll
    for (int p = 0; p < WORLD; ++p) {
      const int peer = (rank + p) % WORLD;               // spread the links
      reinterpret_cast<uint4*>(peers.data[peer] + slot_off + (long long)rank * slot_bytes)[v] = val;
    }
  }
  __syncthreads();
  if (threadIdx.x < WORLD) {
    __threadfence_system();
    st_release_sys(peers.flags[threadIdx.x] + blockIdx.x * WORLD + rank, epoch);
    // monotonic wait: a peer that already finished this epoch may have overwritten its flag with
    // epoch + 1 (it can run at most one epoch ahead) - an equality test would then spin forever
    while ((int32_t)(ld_acquire_sys(my_flags + blockIdx.x * WORLD + threadIdx.x) - epoch) < 0) {}
  }
  __syncthreads();

  const unsigned char* mine = peers.data[rank] + slot_off;
  for (int v = v0 + threadIdx.x; v < v1; v += kArThreads) {
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
    for (int q = 0; q < WORLD; ++q) {
      const uint4 val = ld_cg_16(reinterpret_cast<const uint4*>(mine + (long long)q * slot_bytes) + v);
      const uint32_t w[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = Traits<T>::to_f2(w[i]);
        acc[2 * i] += f.x;
        acc[2 * i + 1] += f.y;
      }
    }
    out[v] = make_uint4(Traits<T>::pack(acc[0], acc[1]), Traits<T>::pack(acc[2], acc[3]),
                        Traits<T>::pack(acc[4], acc[5]), Traits<T>::pack(acc[6], acc[7]));
  }

  // last block to finish advances the epoch (every block read it before arriving here)
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(my_flags + kArTicketWord, 1u);
    if (done == gridDim.x - 1) {
      my_flags[kArTicketWord] = 0u;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(my_flags + kArCounterWord) = epoch;
    }
  }
}

template <typename T>
static int launch_ar(const void* in, void* out, int n_vec, const ArPeers& peers, int rank, int world,
                     long long slot_bytes, cudaStream_t stream) {
  int blocks = (n_vec + kArThreads - 1) / kArThreads;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  if (blocks < 1) blocks = 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(kArThreads); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 1 : 0;
  int rc = 0;
#define PQ_AR(W)                                                                              \
  rc = (int)cudaLaunchKernelEx(&cfg, pq_allreduce_oneshot_kernel<T, W>, (const uint4*)in,     \
                               (uint4*)out, n_vec, peers, rank, (long long)slot_bytes)
  switch (world) {
    case 2: PQ_AR(2); break;
    case 4: PQ_AR(4); break;
    case 8: PQ_AR(8); break;
    default: return PQ_ERR_UNSUPPORTED;
  }
#undef PQ_AR
  return rc;
}

// ---- two-stage all-reduce for messages above the one-shot limit (prefill: [T, hidden] partial sums) ----
// Replaces cross_device_reduce_2stage (csrc/custom_all_reduce.cuh:239-290, dispatch :478-491) and, for
// CUDA-graph capture, the ctypes NCCL binding the reference needs for large messages
// (device_communicators/pynccl.py:109-132): one launch, no host state, epoch on the device.
//   A  copy the input into this rank's peer-mapped staging buffer              (local)
//      barrier 1 (per block: the element -> block map is the same on every rank)
//   B  reduce-scatter: rank r sums chunk r of every rank's staging buffer in rank order, fp32, and
//      writes the rounded chunk into its own result buffer                    (W - 1 remote reads)
//      barrier 2
//   C  all-gather: copy chunk q from rank q's result buffer into the output    (W - 1 remote reads)
// Every rank adds the same fp32 values in the same order and every rank reads the same rounded
// chunks: bitwise identical results across ranks.  The result buffer alternates by launch parity (a
// fast rank may write launch n + 1's chunk while a slow one still gathers launch n's).
constexpr int kAr2Threads = 512;
constexpr int kAr2MaxBlocks = 128;
constexpr int kAr2CounterWord = 2000;  // flag buffer: [block * world + peer] (<= 1024 words), then
constexpr int kAr2TicketWord = 2001;   // this rank's epoch counter and block ticket

struct Ar2Peers {
  unsigned char* stage[PQ_AR_MAX_WORLD];   // [max_bytes] staging copies of the inputs
  unsigned char* res[PQ_AR_MAX_WORLD];     // [2][max_bytes / world] reduced chunks, by launch parity
  uint32_t* flags[PQ_AR_MAX_WORLD];
};

template <int WORLD>
__device__ __forceinline__ void ar2_barrier(const Ar2Peers& peers, int rank, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < WORLD) {
    __threadfence_system();
    st_release_sys(peers.flags[threadIdx.x] + blockIdx.x * WORLD + rank, epoch);
    while ((int32_t)(ld_acquire_sys(peers.flags[rank] + blockIdx.x * WORLD + threadIdx.x) - epoch) < 0) {}
  }
  __syncthreads();
}

template <typename T, int WORLD>
__global__ void __launch_bounds__(kAr2Threads)
pq_allreduce_twostage_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int n_vec, Ar2Peers peers,
                             int rank, long long res_parity_bytes) {
  pdl_launch_dependents();
  pdl_wait();
  uint32_t* my_flags = peers.flags[rank];
  const uint32_t base = *reinterpret_cast<volatile uint32_t*>(my_flags + kAr2CounterWord);
  const int chunk = (n_vec + WORLD - 1) / WORLD;                  // vectors per rank's chunk
  const int per_block = (chunk + gridDim.x - 1) / gridDim.x;
  const int i0 = blockIdx.x * per_block, i1 = min(i0 + per_block, chunk);
  const long long parity_off = (long long)((base >> 1) & 1u) * res_parity_bytes;

  // A: this block's slice of every chunk -> staging
  uint4* my_stage = reinterpret_cast<uint4*>(peers.stage[rank]);
#pragma unroll 1
  for (int q = 0; q < WORLD; ++q)
    for (int i = i0 + threadIdx.x; i < i1; i += kAr2Threads) {
      const int v = q * chunk + i;
      if (v < n_vec) my_stage[v] = in[v];
    }
  ar2_barrier<WORLD>(peers, rank, base + 1u);

  // B: reduce chunk `rank`
  uint4* my_res = reinterpret_cast<uint4*>(peers.res[rank] + parity_off);
  for (int i = i0 + threadIdx.x; i < i1; i += kAr2Threads) {
    const int v = rank * chunk + i;
    if (v >= n_vec) continue;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int q = 0; q < WORLD; ++q) {
      const uint4 val = ld_cg_16(reinterpret_cast<const uint4*>(peers.stage[q]) + v);
      const uint32_t w[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = Traits<T>::to_f2(w[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
    my_res[i] = make_uint4(Traits<T>::pack(acc[0], acc[1]), Traits<T>::pack(acc[2], acc[3]),
                           Traits<T>::pack(acc[4], acc[5]), Traits<T>::pack(acc[6], acc[7]));
  }
  ar2_barrier<WORLD>(peers, rank, base + 2u);

  // C: gather every rank's reduced chunk
#pragma unroll 1
  for (int p = 0; p < WORLD; ++p) {
    const int q = (rank + p) % WORLD;                               // spread the links
    const uint4* src = reinterpret_cast<const uint4*>(peers.res[q] + parity_off);
    for (int i = i0 + threadIdx.x; i < i1; i += kAr2Threads) {
      const int v = q * chunk + i;
      if (v < n_vec) out[v] = ld_cg_16(src + i);
    }
  }

  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(my_flags + kAr2TicketWord, 1u);
    if (done == gridDim.x - 1) {
      my_flags[kAr2TicketWord] = 0u;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(my_flags + kAr2CounterWord) = base + 2u;
    }
  }
}

int allreduce_twostage_dispatch(const void* in, void* out, int64_t n_elems, int dtype, void* const* peer_stage,
                                void* const* peer_res, void* const* peer_flags, int rank, int world,
                                int64_t max_bytes, cudaStream_t stream) {
  Ar2Peers peers{};
  for (int r = 0; r < world; ++r) {
    if (!peer_stage[r] || !peer_res[r] || !peer_flags[r]) return PQ_ERR_NULL;
    peers.stage[r] = (unsigned char*)peer_stage[r];
    peers.res[r] = (unsigned char*)peer_res[r];
    peers.flags[r] = (uint32_t*)peer_flags[r];
  }
  const int n_vec = (int)(n_elems * 2 / 16);
  const int chunk = (n_vec + world - 1) / world;
  int blocks = (chunk + kAr2Threads * 4 - 1) / (kAr2Threads * 4);       // >= 4 vectors per thread
  if (blocks > kAr2MaxBlocks) blocks = kAr2MaxBlocks;
  if (blocks < 1) blocks = 1;
  const long long res_parity_bytes = ((max_bytes / world + 255) / 256) * 256;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(blocks); cfg.blockDim = dim3(kAr2Threads); cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = g_use_pdl ? 1 : 0;
  int rc = 0;
#define PQ_AR2(TT, W)                                                                              \
  rc = (int)cudaLaunchKernelEx(&cfg, pq_allreduce_twostage_kernel<TT, W>, (const uint4*)in, (uint4*)out, \
                               n_vec, peers, rank, res_parity_bytes)
#define PQ_AR2_W(TT)                               \
  switch (world) {                                  \
    case 2: PQ_AR2(TT, 2); break;                   \
    case 4: PQ_AR2(TT, 4); break;                   \
    case 8: PQ_AR2(TT, 8); break;                   \
    default: return PQ_ERR_UNSUPPORTED;             \
  }
  if (dtype == PQ_F16) { PQ_AR2_W(__half) } else { PQ_AR2_W(__nv_bfloat16) }
#undef PQ_AR2_W
#undef PQ_AR2
  return rc;
}

int allreduce_oneshot_dispatch(const void* in, void* out, int64_t n_elems, int dtype,
                               void* const* peer_data, void* const* peer_flags, int rank, int world,
                               int64_t slot_bytes_in, cudaStream_t stream) {
  ArPeers peers{};
  for (int r = 0; r < world; ++r) {
    if (!peer_data[r] || !peer_flags[r]) return PQ_ERR_NULL;
    peers.data[r] = (unsigned char*)peer_data[r];
    peers.flags[r] = (uint32_t*)peer_flags[r];
  }
  const int n_vec = (int)(n_elems * 2 / 16);
  const long long slot_bytes = (long long)slot_bytes_in;
  return dtype == PQ_F16 ? launch_ar<__half>(in, out, n_vec, peers, rank, world, slot_bytes, stream)
                         : launch_ar<__nv_bfloat16>(in, out, n_vec, peers, rank, world, slot_bytes, stream);
}

}  // namespace pq

// ---- buffer + IPC helpers (host) ----------------------------------------------------------
extern "C" {

PQ_API int pq_b200_malloc(int64_t bytes, void** out) {
  if (!out || bytes <= 0) return PQ_ERR_NULL;
  return (int)cudaMalloc(out, (size_t)bytes);
}
PQ_API int pq_b200_free(void* p) { return (int)cudaFree(p); }
PQ_API int pq_b200_memset_async(void* p, int value, int64_t bytes, void* stream) {
  return (int)cudaMemsetAsync(p, value, (size_t)bytes, (cudaStream_t)stream);
}
PQ_API int pq_b200_ipc_get_handle(void* ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, ptr);
  if (e != cudaSuccess) return (int)e;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}
PQ_API int pq_b200_ipc_open_handle(void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  return (int)cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess);
}
PQ_API int pq_b200_ipc_close_handle(void* p) { return (int)cudaIpcCloseMemHandle(p); }

}  // extern "C"
